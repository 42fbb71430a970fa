"""compute_mvoe alone (ccmpc_mvoe_fit) on 13.1 M SPD pairs: both forms, both iterate modes, GB/s per variant."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccmpc_b200 import ops
dev = torch.device("cuda:0")
M = 4096 * 8 * 16 * 25
g = torch.Generator(device=dev).manual_seed(1)
def spd():
    sx = torch.rand(M, generator=g, device=dev, dtype=torch.float64) * 2.9 + 0.1
    sy = torch.rand(M, generator=g, device=dev, dtype=torch.float64) * 2.9 + 0.1
    rho = torch.rand(M, generator=g, device=dev, dtype=torch.float64) * 1.6 - 0.8
    return torch.stack([sx * sx, rho * sx * sy, sy * sy], dim=-1).contiguous()
s1, s2 = spd(), spd()
for iso in (False, True):
    for accel in (0, 2):
        for want in (True, False):
            f = lambda: ops.mvoe_fit(s1, None if iso else s2, accel=accel, want_iters=want)
            for _ in range(3): f()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(10): f()
            b.record(); torch.cuda.synchronize()
            ms = a.elapsed_time(b) / 10
            byt = (24 if iso else 48) + 24 + 8 + (5 if want else 0)
            print(f"iso={iso} accel={accel} iters/status={want}: {ms:.3f} ms  {M/ms/1e6:.1f} Gfit/s  {M*byt/ms/1e6:.0f} GB/s")
