#!/usr/bin/env python
"""Where does the end-to-end leg of bench.py spend its time beyond the kernels?  Host timestamps around the calls of one
step (H2D enqueue, MidlevelConstraints.mc_collision_risk, D2H enqueue, stream sync) and CUDA events around the device work."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccmpc_b200 import midlevel, ops, synth

dev = torch.device("cuda:0")
O, K, T, n = 8, 4, 16, 250_000
C = O * K
sc = synth.make_scene(3, O, K, T + 1)
m, c, x = sc.moments()
parts = (m.reshape(C, T + 1, 2), c.reshape(C, T + 1, 2, 2), x.reshape(C, T + 1, T + 1, 2, 2), synth.reference_trajectory(T))
sizes = [int(np.prod(a.shape)) for a in parts]
arena_p = torch.empty(sum(sizes), dtype=torch.float64).pin_memory()
arena_d = torch.empty(sum(sizes), dtype=torch.float64, device=dev)
o, views = 0, []
for a, sz in zip(parts, sizes):
    arena_p[o:o + sz].copy_(torch.from_numpy(np.ascontiguousarray(a)).reshape(-1))
    views.append(arena_d[o:o + sz].view(a.shape))
    o += sz
m_, c_, x_, e_ = views
mc = midlevel.MidlevelConstraints(T, out_dir=None, device=dev)
mom = midlevel.Moments(m_, c_, x_, np.repeat(np.arange(O), K), np.tile(np.arange(K), O), [K] * O, T + 1)
sums_host = torch.empty((C, ops.sums_len(T)), dtype=torch.float64).pin_memory()
rows = []
for i in range(30):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    arena_d.copy_(arena_p, non_blocking=True)
    t1 = time.perf_counter()
    res = mc.mc_collision_risk(mom, [K] * O, T, e_, n_samples=n, seed=1234 + i, sample_offset=0, want_moments=True)
    t2 = time.perf_counter()
    sums_host.copy_(res["sums"], non_blocking=True)
    e1.record()
    t3 = time.perf_counter()
    torch.cuda.current_stream().synchronize()
    t4 = time.perf_counter()
    rows.append(((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3, (t4 - t0) * 1e3, e0.elapsed_time(e1)))
r = np.median(np.asarray(rows[5:]), axis=0)
print("median ms: H2D enqueue %.3f | mc_collision_risk (host) %.3f | D2H enqueue %.3f | sync wait %.3f | step %.3f | device span %.3f"
      % tuple(r))
