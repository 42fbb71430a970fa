"""Where does the planner-class c2 step spend its host time?  python tools/class_profile.py"""
import cProfile, os, pstats, sys, time, types
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccmpc_b200 import midlevel, synth
dev = torch.device("cuda:0")
O, K, T, N = 4, 25, 8, 5000
sc = synth.make_scene(2, O, K, T)
z, pos = synth.sample_scene(sc, N, 2)
ovs = []
for o in range(O):
    pp = [np.ascontiguousarray(pos[o][z[o] == k]) for k in range(K)]
    ovs.append(types.SimpleNamespace(pred_positions=pp, pred_yaws=[np.zeros(p.shape[:2]) for p in pp], n_states=K,
                                     bbox=np.array([3.7, 1.79])))
mc = midlevel.MidlevelConstraints(T, out_dir=None, device=dev)
prm = types.SimpleNamespace(frame=100)
f = lambda: mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(prm, ovs, Tsh=T, ref_traj=sc.ref_traj)
for _ in range(5): f()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(50): f()
print("ms per call", (time.perf_counter() - t0) * 20)
pr = cProfile.Profile(); pr.enable()
for _ in range(50): f()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)

# ---- c1 shrinking step 2 (Tsh = 7, 3 x 1e6 ideal samples)
sc1 = synth.make_scene(1, 1, 3, 8)
z1, pos1 = synth.sample_scene(sc1, 5000, 1)
pp = [np.ascontiguousarray(pos1[0][z1[0] == k]) for k in range(3)]
ov1 = [types.SimpleNamespace(pred_positions=pp, pred_yaws=[np.zeros(p.shape[:2]) for p in pp], n_states=3,
                             bbox=np.array([3.7, 1.79]))]
mc1 = midlevel.MidlevelConstraints(8, out_dir=None, device=dev)
mc1.compute_obstacle_constraints_GMM_Minkowski_idealprediction(types.SimpleNamespace(frame=100), ov1, Tsh=8, ref_traj=sc1.ref_traj)
g = lambda: mc1.compute_obstacle_constraints_GMM_Minkowski_idealprediction(types.SimpleNamespace(frame=110), ov1, Tsh=7, ref_traj=sc1.ref_traj)
for _ in range(5): g()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(50): g()
print("c1 step 2: ms per call", (time.perf_counter() - t0) * 20)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); g(); b.record(); torch.cuda.synchronize(); print("device span ms", a.elapsed_time(b))
pr = cProfile.Profile(); pr.enable()
for _ in range(50): g()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
