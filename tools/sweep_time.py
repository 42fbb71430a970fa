"""c4 MVOE sweep timing: python tools/sweep_time.py  (accel 0 / 1 / 2)"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccmpc_b200 import ops
from tools.prof_driver import c4_inputs
dev = torch.device("cuda:0")
cov, mu, ref, chi = c4_inputs(dev)
M = cov.numel() // 3
for accel in (0, 1, 2):
    out = None
    for _ in range(3):
        out = ops.mvoe_halfplane_sweep(cov, mu, ref, chi, accel=accel, out=out)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        ops.mvoe_halfplane_sweep(cov, mu, ref, chi, accel=accel, out=out)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    print(f"accel {accel}: {ms:.3f} ms  {M / ms * 1e-6:.2f} G fits/s  {M * 90 / ms * 1e-6 / 6548.8:.3f} of HBM (90 B/fit)")
