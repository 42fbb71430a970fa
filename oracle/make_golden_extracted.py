"""Goldens of reference functions whose MODULES do not import here but whose bodies run as they are.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Run in the build container:

    python -m oracle.make_golden_extracted

Each function is compiled on its own from the reference file by ``ref_loader.extract`` (no module
import, no copy) and executed on seeded inputs:

  ref_l4.npz             compute_L4_outerapproximation          midlevel/util.py:171-200
  ref_moments.npz        MidlevelAgent.save_moments             v8ideal/__init__.py:2575-2618
  ref_ideal_rollout.npz  MidlevelAgent.predict_ideal            v8ideal/__init__.py:2620-2711
  ref_grouping.npz       OVehicle.from_trajectron               midlevel/ovehicle.py:23-117
                         (fed by the bucket loop of make_ovehicles, v8ideal:486-494, restated inline)
  ref_cost.npz           getData_singleEpisode.cost_final       dataForCost_ref.py:198-232
"""
from __future__ import annotations

import contextlib
import io
import os
import pickle
import sys
import tempfile
import types

import numpy as np
import scipy.spatial

from . import ref_loader, ref_planner as rp, restate, synth

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
N_IDEAL = 1_000_000


def gen_l4(rng):
    ns = ref_loader.extract(ref_loader.MIDLEVEL_UTIL, ["compute_L4_outerapproximation"], {"np": np})
    fn = ns["compute_L4_outerapproximation"]
    npu = ref_loader.npu()
    M, N = 24, 150
    centers = rng.uniform(-200, 200, size=(M, 1, 2)) + rng.normal(0, 2.0, size=(M, N, 2))
    yaw = rng.uniform(-np.pi, np.pi, size=(M, 1)) + rng.normal(0, 0.3, size=(M, N))
    theta = yaw.mean(axis=1)
    lw = np.array([3.7, 1.79])
    At = np.zeros((M, 4, 2)); b0 = np.zeros((M, 4))
    for i in range(M):
        At[i], b0[i] = fn(theta[i], npu.vertices_of_bboxes(centers[i], yaw[i], lw))
    np.savez(os.path.join(OUT, "ref_l4.npz"), centers=centers, yaw=yaw, theta=theta, lw=lw, At=At, b0=b0)


def gen_moments_and_rollout():
    """save_moments on a small scene, then predict_ideal from the pickle it wrote."""
    O, K, T, N = 2, 2, 6, 400
    sc = synth.make_scene(31, O, K, T)
    z, pos = synth.sample_scene(sc, N, 31)
    ovs = []
    for o in range(O):
        pp = [pos[o][z[o] == k] for k in range(K)]
        ovs.append(rp.FakeOVehicle(pp, [np.zeros(p.shape[:2]) for p in pp], (3.7, 1.79)))
    agent = rp.make_agent(T, 0, seed=404)
    params = rp.make_params(ovs, 50)
    agent.save_moments(ovs, params.O, params.K, T, T, 0, params)
    with open(os.path.join("out", "data", "agent0_frame50_moments"), "rb") as f:
        d = pickle.load(f)
    flat, offs, co, ck = synth.cells_from_samples(z, pos, K)
    C = len(co)
    mean = np.zeros((C, T, 2)); cov = np.zeros((C, T, 2, 2)); cross = np.zeros((C, T, T, 2, 2))
    for c in range(C):
        for t in range(T):
            mean[c, t] = d["mean_p0p1"][co[c]][ck[c]][t]
            cov[c, t] = d["cov_p0p1"][co[c]][ck[c]][t]
            for tau in range(t):
                cross[c, t, tau] = d["cross_cov"][co[c]][ck[c]][t][tau]
    np.savez(os.path.join(OUT, "ref_moments.npz"), samples=flat, offsets=offs, cell_obstacle=co, cell_mode=ck,
             mean=mean, cov=cov, cross=cross)
    # predict_ideal (frame 60 reads frame 50), Tsh = T - 1: n_samples = 1e6 is a literal of the reference
    Tsh = T - 1
    traj_all = agent.predict_ideal(ovs, Tsh, 0, rp.make_params(ovs, 60))
    keep = 512
    head = np.stack([traj_all[co[c]][ck[c]][:keep] for c in range(C)])
    tail = np.stack([traj_all[co[c]][ck[c]][-keep:] for c in range(C)])
    tmean = np.stack([traj_all[co[c]][ck[c]].mean(axis=0) for c in range(C)])
    x0 = np.concatenate(agent.rnd.x0_draws, axis=0)
    # the same stream again, to record which normals those rows received
    rs = np.random.RandomState(404)
    Zh = np.zeros((C, keep, Tsh, 2)); Zt = np.zeros((C, keep, Tsh, 2))
    for c in range(C):
        rs.standard_normal((1, 2))
        for t in range(Tsh):
            blk = rs.standard_normal((N_IDEAL, 2))
            Zh[c, :, t], Zt[c, :, t] = blk[:keep], blk[-keep:]
    np.savez(os.path.join(OUT, "ref_ideal_rollout.npz"), mean=mean, cov=cov, cross=cross, Tsh=Tsh, x0=x0,
             Z_head=Zh, Z_tail=Zt, traj_head=head, traj_tail=tail, traj_mean=tmean, noise_seed=404,
             n_ideal=N_IDEAL)


def gen_grouping():
    class _Rec:
        def __init__(self, node, T, past, ground_truth, latent_pmf, pred_positions, pred_yaws, init_center, bbox):
            self.latent_pmf, self.pred_positions, self.pred_yaws = latent_pmf, pred_positions, pred_yaws
            self.init_center = init_center
    ns = {"np": np, "scipy": scipy, "DEFAULT_BBOX": np.array([4.5, 2.5])}
    ref_loader.extract(ref_loader.OVEHICLE, ["from_trajectron"], ns, class_name="OVehicle")
    from_trajectron = ns["OVehicle"].__dict__["from_trajectron"].__func__
    out = {}
    cases = [(41, 700, 6, 25, 0.1), (42, 300, 8, 25, 0.1), (43, 64, 4, 8, 0.2)]
    for ci, (seed, N, T, S, filt) in enumerate(cases):
        rng = np.random.default_rng(seed)
        pmf = rng.dirichlet(np.full(S, 0.25))
        z = rng.choice(S, size=N, p=pmf)
        centers = rng.uniform(-30, 30, size=(S, 1, 2)) + np.cumsum(rng.normal(0, 2, size=(S, T, 2)), axis=1)
        pred = (centers[z] + rng.normal(0, 0.7, size=(N, T, 2))).astype(np.float32)
        minpos = np.array([100.0, -150.0])
        pos_last = minpos + rng.uniform(-5, 5, size=2)
        veh_predict = pred + minpos                                     # v8ideal:486
        buckets = [[] for _ in range(S)]                                # v8ideal:489-494
        for j, p in enumerate(veh_predict):
            buckets[z[j]].append(p)
        buckets = [np.array(b) for b in buckets]
        with contextlib.redirect_stderr(io.StringIO()):
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                ov = from_trajectron(_Rec, None, T, None, pos_last[None], pmf, buckets, filter_pmf=filt)
        Kc = len(ov.pred_positions)
        out.update({f"c{ci}_pred": pred, f"c{ci}_z": z.astype(np.int32), f"c{ci}_pmf": pmf, f"c{ci}_minpos": minpos,
                    f"c{ci}_pos_last": pos_last, f"c{ci}_filter": filt, f"c{ci}_K": Kc,
                    f"c{ci}_offsets": np.cumsum([0] + [p.shape[0] for p in ov.pred_positions]),
                    f"c{ci}_positions": np.concatenate(ov.pred_positions), f"c{ci}_yaws": np.concatenate(ov.pred_yaws),
                    f"c{ci}_init_center": ov.init_center, f"c{ci}_latent_pmf": ov.latent_pmf})
    out["n_cases"] = len(cases)
    np.savez_compressed(os.path.join(OUT, "ref_grouping.npz"), **out)


def gen_cost(rng):
    ns = {"np": np, "util": types.SimpleNamespace(AttrDict=lambda **kw: types.SimpleNamespace(**kw))}
    ref_loader.extract(ref_loader.DATA_FOR_COST, ["cost_final"], ns, class_name="getData_singleEpisode")
    obj = ns["getData_singleEpisode"]()
    E, H = 12, 9
    X = rng.uniform([150, -100, -3, 0], [200, -60, 3, 10], size=(E, H, 4))
    U = rng.uniform(-2, 2, size=(E, H, 2))
    lens = rng.integers(2, H + 1, size=E)
    lens[0], lens[1] = 2, H
    cost = np.zeros(E); costU = np.zeros(E)
    with contextlib.redirect_stdout(io.StringIO()):                     # cost_final prints (:230)
        for e in range(E):
            cost[e], costU[e] = obj.cost_final([x for x in X[e, :lens[e]]], [u for u in U[e, :lens[e]]])
    np.savez(os.path.join(OUT, "ref_cost.npz"), X=X, U=U, lens=lens, cost=cost, costU=costU)


def main():
    if not ref_loader.available():
        sys.exit("reference not found under %s" % ref_loader.REFERENCE_ROOT)
    rng = np.random.default_rng(20261105)
    gen_l4(rng)
    gen_grouping()
    gen_cost(rng)
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.makedirs(os.path.join(tmp, "out", "data"))
        os.chdir(tmp)
        try:
            gen_moments_and_rollout()
        finally:
            os.chdir(cwd)
    for f in ("ref_l4.npz", "ref_moments.npz", "ref_ideal_rollout.npz", "ref_grouping.npz", "ref_cost.npz"):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
