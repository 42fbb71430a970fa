"""c1 shrinking step 2 through the planner class, a few times (for an ncu launch list)."""
import os, sys, types, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccmpc_b200 import midlevel, synth
dev = torch.device("cuda:0")
sc1 = synth.make_scene(1, 1, 3, 8)
z1, pos1 = synth.sample_scene(sc1, 5000, 1)
pp = [np.ascontiguousarray(pos1[0][z1[0] == k]) for k in range(3)]
ov1 = [types.SimpleNamespace(pred_positions=pp, pred_yaws=[np.zeros(p.shape[:2]) for p in pp], n_states=3, bbox=np.array([3.7, 1.79]))]
mc1 = midlevel.MidlevelConstraints(8, out_dir=None, device=dev)
mc1.compute_obstacle_constraints_GMM_Minkowski_idealprediction(types.SimpleNamespace(frame=100), ov1, Tsh=8, ref_traj=sc1.ref_traj)
for _ in range(4):
    mc1.compute_obstacle_constraints_GMM_Minkowski_idealprediction(types.SimpleNamespace(frame=110), ov1, Tsh=7, ref_traj=sc1.ref_traj)
torch.cuda.synchronize()
print("ok")
