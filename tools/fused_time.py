import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccmpc_b200 import ops
from oracle import synth
dev = torch.device("cuda:0")
O, K, T, n = 8, 4, int(os.environ.get("FT_T", "16")), int(os.environ.get("FT_N", "250000"))
C = O * K
sc = synth.make_scene(3, O, K, T + 1)
m, c, x = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in sc.moments())
m, c, x = m.reshape(C, T + 1, 2), c.reshape(C, T + 1, 2, 2), x.reshape(C, T + 1, T + 1, 2, 2)
ego = torch.from_numpy(synth.reference_trajectory(T)).to(dev)
traj = torch.empty((C, n, T, 2), dtype=torch.float64, device=dev)
pivot = m[:, 1:T + 1].contiguous()
ws = ops.mc_rollout_sums_workspace(C, n, T, dev)
sums = torch.empty((C, ops.sums_len(T)), dtype=torch.float64, device=dev)
import ctypes
from ccmpc_b200 import _lib
lib = _lib.load()
exps = [int(a) for a in sys.argv[1:]] or [0]
for exp in exps:
  lib.ccmpc_debug_ws_exp(ctypes.c_int(exp))
  for mat in (True, False):
      f = lambda i: ops.mc_rollout_sums(m, c, x, None, n, T, pivot, ego, seed=i, materialize=mat, workspace=ws, out=traj, sums_out=sums)
      for i in range(3): f(i)
      torch.cuda.synchronize()
      a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      a.record()
      for i in range(10): f(3 + i)
      b.record(); torch.cuda.synchronize()
      print("exp", exp, "fused materialize", mat, f"{a.elapsed_time(b)/10:.3f} ms")
