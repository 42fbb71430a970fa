"""ncu driver for the fused kernel with an experiment flag: python tools/fused_prof.py <exp>"""
import ctypes, os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccmpc_b200 import ops, _lib
from oracle import synth
lib = _lib.load()
lib.ccmpc_debug_ws_exp(ctypes.c_int(int(sys.argv[1]) if len(sys.argv) > 1 else 0))
dev = torch.device("cuda:0")
O, K, T, n = 8, 4, 16, 250_000
C = O * K
sc = synth.make_scene(3, O, K, T + 1)
m, c, x = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in sc.moments())
m, c, x = m.reshape(C, T + 1, 2), c.reshape(C, T + 1, 2, 2), x.reshape(C, T + 1, T + 1, 2, 2)
ego = torch.from_numpy(synth.reference_trajectory(T)).to(dev)
traj = torch.empty((C, n, T, 2), dtype=torch.float64, device=dev)
pv = m[:, 1:T + 1].contiguous()
ws = ops.mc_rollout_sums_workspace(C, n, T, dev)
for i in range(4):
    ops.mc_rollout_sums(m, c, x, None, n, T, pv, ego, seed=i, workspace=ws, out=traj)
torch.cuda.synchronize()
print("ok")
