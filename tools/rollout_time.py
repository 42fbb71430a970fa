"""c3 stand-alone rollout timing: python tools/rollout_time.py   (CCMPC_ROLLOUT_NO_TMA=1: the LDS + STG flush)"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccmpc_b200 import ops, synth
dev = torch.device("cuda:0")
for T in (16, 8):
    O, K, n = 8, 4, 250000
    C = O * K
    sc = synth.make_scene(3, O, K, T + 1)
    m, c, x = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in sc.moments())
    m, c, x = m.reshape(C, T + 1, 2), c.reshape(C, T + 1, 2, 2), x.reshape(C, T + 1, T + 1, 2, 2)
    traj = torch.empty((C, n, T, 2), dtype=torch.float64, device=dev)
    for i in range(3):
        ops.mc_rollout(m, c, x, None, n, T, seed=i, out=traj)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(10):
        ops.mc_rollout(m, c, x, None, n, T, seed=3 + i, out=traj)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    print(f"T={T} rollout {ms:.3f} ms  {C*n*T*16/ms*1e-6/6548.8:.3f} of HBM  no_tma={os.environ.get('CCMPC_ROLLOUT_NO_TMA')}")
