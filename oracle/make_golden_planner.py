"""Goldens of the PLANNER-LEVEL methods, produced by the reference's own code.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Run in the build container:

    python -m oracle.make_golden_planner

Runs, through oracle/ref_planner.py (the reference's unmodified MidlevelAgent
methods compiled from /root/reference, cvxpy and np.random shimmed):

  ref_planner_c1.npz     config c1 (O=1, ph8, N=5000): Minkowski builder at
                         Tsh = 8 (frame 100), 7 (frame 110, predict_ideal from the
                         frame-100 ``_moments`` pickle) and 6 (frame 120, from the
                         moments of the ideal trajectories), and the affine builder.
  ref_planner_kgrow.npz  O=2: step 1 with K = [2, 2], step 2 with K = [3, 2]
                         (the "K changed" fallback of v8ideal:2652-2656).
  ref_planner_c2.npz     config c2 (O=4, K=25, ph8, N=5000): Minkowski at Tsh = 8
                         and the affine builder.

Inputs are stored as the float32 local-frame predictions the planner receives
(prediction.py:104) + minpos, so ``pred.astype(float64) + minpos`` reproduces
the float64 samples exactly (v8ideal:486).  The predict_ideal noise is NOT
stored: it comes from ``np.random.RandomState(seed)`` (frozen legacy stream),
one ``standard_normal((1, 2))`` per cell for the x0 draw followed by
``standard_normal((1_000_000, 2))`` per step; the tests replay it and check the
stored checksums.  The x0 draws themselves are stored (they go through an SVD).
"""
from __future__ import annotations

import logging
import os
import pickle
import sys
import tempfile

import numpy as np

from . import ref_loader, ref_planner as rp, restate, synth

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
N_IDEAL = 1_000_000        # v8ideal:2640 (a literal inside predict_ideal)
BBOX = (3.7, 1.79)


class _Capture(logging.Handler):
    """Catches ``logging.info(f"prob_lower_save:{...}")`` (v8ideal:963), the only place the
    shrinking-horizon prob_lower_save of a Tsh < ph step is visible."""

    def __init__(self):
        super().__init__(level=logging.INFO)
        self.last = None

    def emit(self, record):
        msg = record.getMessage()
        if msg.startswith("prob_lower_save:"):
            self.last = eval(msg[len("prob_lower_save:"):], {"np": np, "nan": float("nan"), "None": None})


def ovehicles_from(pred_f32, z, pos_last, K_per_obstacle, minpos):
    """Fake OVehicles: cell k of obstacle o = its samples with z == k, in sample order (v8ideal:489-494)."""
    ovs = []
    pos = pred_f32.astype(np.float64) + minpos
    for o in range(pos.shape[0]):
        pp = [pos[o][z[o] == k] for k in range(K_per_obstacle[o])]
        yy = [restate.yaws_from_positions(p, pos_last[o]) for p in pp]
        ovs.append(rp.FakeOVehicle(pp, yy, BBOX, past=pos_last[o][None]))
    return ovs


def _dense_ok(nested, O, Kmax):
    out = np.full((O, Kmax), np.nan)
    for o in range(O):
        for k in range(Kmax):
            if nested[o][k] is not None:
                out[o, k] = nested[o][k]
    return out


def _moments_dense(d):
    mean_l, cov_l, cross_l = d["mean_p0p1"], d["cov_p0p1"], d["cross_cov"]
    O, Kmax, T = len(mean_l), len(mean_l[0]), len(mean_l[0][0])
    mean = np.full((O, Kmax, T, 2), np.nan)
    cov = np.full((O, Kmax, T, 2, 2), np.nan)
    cross = np.full((O, Kmax, T, max(T - 1, 1), 2, 2), np.nan)
    for o in range(O):
        for k in range(Kmax):
            for t in range(T):
                if mean_l[o][k][t] is None:
                    continue
                mean[o, k, t] = mean_l[o][k][t]
                cov[o, k, t] = cov_l[o][k][t]
                for tau in range(t):
                    cross[o, k, t, tau] = cross_l[o][k][t][tau]
    return mean, cov, cross


def run_step(agent, cap, ovs, frame, Tsh, ref_traj, builder="minkowski", prefix=""):
    """One builder call of the reference; returns a flat dict of arrays keyed ``prefix + name``."""
    Tpred = agent._MidlevelAgent__prediction_horizon
    params = rp.make_params(ovs, frame)
    eps_ura = rp.eps_ura_of(params)
    cap.last = None
    x0_before = len(agent.rnd.x0_draws)
    fn = (agent.compute_obstacle_constraints_GMM_Minkowski_idealprediction if builder == "minkowski"
          else agent.compute_obstacle_constraints_GMM_affine)
    (rows, vertices, A_union, b_union, OVconstraint, direct, ov_mean, ov_cov, last) = fn(
        params, ovs, None, None, rp.make_temp_x(Tpred), eps_ura, None, Tsh, ref_traj)
    n, t, side, rhs = rp.rows_to_arrays(rows)
    O, Kmax = params.O, int(max(params.K))
    out = {"n": n, "t": t, "side": side, "rhs": rhs, "K": params.K.astype(np.int64), "eps_ura": eps_ura,
           "Tsh": np.int64(Tsh), "frame": np.int64(frame),
           "A_union": rp.nested_to_array(A_union, (4, 2), Tpred, Kmax, O),
           "b_union": rp.nested_to_array(b_union, (4,), Tpred, Kmax, O),
           "vertices_t0k0o0": np.asarray(vertices[0][0][0]),
           "OVconstraint": np.bool_(OVconstraint), "direct_all_none": np.bool_(all(d is None for d in direct)),
           "last": np.int64(last),
           "ov_mean": np.stack([_dense_ok(a, O, Kmax) for a in ov_mean]),
           "ov_cov": np.stack([_dense_ok(a, O, Kmax) for a in ov_cov])}
    if builder == "minkowski":
        out["prob_lower_save"] = np.asarray(cap.last, dtype=np.float64)
        path = os.path.join("out", "data", f"agent0_frame{frame}_moments")
        with open(path, "rb") as f:
            raw = f.read()
        mean, cov, cross = _moments_dense(pickle.loads(raw))
        out.update(mom_mean=mean, mom_cov=cov, mom_cross=cross, mom_pickle=np.frombuffer(raw, dtype=np.uint8))
        if Tsh < Tpred:
            out["x0"] = np.concatenate(agent.rnd.x0_draws[x0_before:], axis=0)          # (C, 2), cell order
    return {prefix + k: v for k, v in out.items()}


def _inputs(seed, O, K, T, N, flat=True):
    sc = synth.make_scene(seed, O, K, T, flat=flat)
    z, pos = synth.sample_scene(sc, N, seed)
    pred = (pos - synth.MINPOS).astype(np.float32)
    assert np.array_equal(pred.astype(np.float64) + synth.MINPOS, pos)
    return sc, z, pred


def noise_checksums(seed, cells_per_step, steps_T):
    """Replay of the seeded stream: per (step, cell, t) the sum of the Z block - the tests must reproduce it."""
    rs = np.random.RandomState(seed)
    sums = []
    for C, T in zip(cells_per_step, steps_T):
        for _ in range(C):
            rs.standard_normal((1, 2))
            for _t in range(T):
                sums.append(rs.standard_normal((N_IDEAL, 2)).sum())
    return np.asarray(sums)


def gen_c1():
    O, K, T, N, seed = 1, 3, 8, 5000, 1
    sc, z, pred = _inputs(seed, O, K, T, N)
    ovs = ovehicles_from(pred, z, sc.pos_last, [K] * O, synth.MINPOS)
    agent, cap = rp.make_agent(T, 0, seed=101), _Capture()
    logging.getLogger().addHandler(cap)
    logging.getLogger().setLevel(logging.INFO)
    out = dict(pred=pred, z=z.astype(np.int32), pos_last=sc.pos_last, ref_traj=sc.ref_traj, minpos=synth.MINPOS,
               bbox=np.asarray(BBOX), noise_seed=np.int64(101), n_ideal=np.int64(N_IDEAL))
    out.update(run_step(agent, cap, ovs, 100, T, sc.ref_traj, prefix="s1_"))
    out.update(run_step(agent, cap, ovs, 110, T - 1, sc.ref_traj, prefix="s2_"))
    out.update(run_step(agent, cap, ovs, 120, T - 2, sc.ref_traj, prefix="s3_"))
    out.update(run_step(agent, cap, ovs, 100, T, sc.ref_traj, builder="affine", prefix="af_"))
    out.update(run_step(agent, cap, ovs, 100, 5, sc.ref_traj, builder="affine", prefix="af5_"))
    out["noise_sums"] = noise_checksums(101, [K * O, K * O], [T - 1, T - 2])
    logging.getLogger().removeHandler(cap)
    np.savez_compressed(os.path.join(OUT, "ref_planner_c1.npz"), **out)


def gen_kgrow():
    O, T, N, seed = 2, 8, 1200, 21
    sc, z3, pred = _inputs(seed, O, 3, T, N)
    # step 1 sees two modes per obstacle (latent 2 folded into latent 1), step 2 sees K = [3, 2]
    z_a = np.minimum(z3, 1)
    z_b = z3.copy()
    z_b[1] = np.minimum(z3[1], 1)
    ovs_a = ovehicles_from(pred, z_a, sc.pos_last, [2, 2], synth.MINPOS)
    ovs_b = ovehicles_from(pred, z_b, sc.pos_last, [3, 2], synth.MINPOS)
    agent, cap = rp.make_agent(T, 0, seed=202), _Capture()
    logging.getLogger().addHandler(cap)
    logging.getLogger().setLevel(logging.INFO)
    out = dict(pred=pred, z_a=z_a.astype(np.int32), z_b=z_b.astype(np.int32), pos_last=sc.pos_last,
               ref_traj=sc.ref_traj, minpos=synth.MINPOS, bbox=np.asarray(BBOX), noise_seed=np.int64(202),
               n_ideal=np.int64(N_IDEAL))
    out.update(run_step(agent, cap, ovs_a, 100, T, sc.ref_traj, prefix="s1_"))
    out.update(run_step(agent, cap, ovs_b, 110, T - 1, sc.ref_traj, prefix="s2_"))
    out["noise_sums"] = noise_checksums(202, [5], [T - 1])
    logging.getLogger().removeHandler(cap)
    np.savez_compressed(os.path.join(OUT, "ref_planner_kgrow.npz"), **out)


def gen_c2():
    O, K, T, N, seed = 4, 25, 8, 5000, 2
    sc, z, pred = _inputs(seed, O, K, T, N)
    ovs = ovehicles_from(pred, z, sc.pos_last, [K] * O, synth.MINPOS)
    agent, cap = rp.make_agent(T, 0, seed=303), _Capture()
    logging.getLogger().addHandler(cap)
    logging.getLogger().setLevel(logging.INFO)
    out = dict(pred=pred, z=z.astype(np.int32), pos_last=sc.pos_last, ref_traj=sc.ref_traj, minpos=synth.MINPOS,
               bbox=np.asarray(BBOX))
    s1 = run_step(agent, cap, ovs, 100, T, sc.ref_traj, prefix="s1_")
    s1.pop("s1_mom_pickle")                      # 100 cells of nested lists: keep the dense arrays only
    out.update(s1)
    out.update(run_step(agent, cap, ovs, 100, T, sc.ref_traj, builder="affine", prefix="af_"))
    logging.getLogger().removeHandler(cap)
    np.savez_compressed(os.path.join(OUT, "ref_planner_c2.npz"), **out)


def main(which=("c1", "kgrow", "c2")):
    if not ref_loader.available():
        sys.exit("reference not found under %s" % ref_loader.REFERENCE_ROOT)
    os.makedirs(OUT, exist_ok=True)
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:       # the reference writes to ./out/data (v8ideal:2613)
        os.makedirs(os.path.join(tmp, "out", "data"))
        os.chdir(tmp)
        try:
            for w in which:
                {"c1": gen_c1, "kgrow": gen_kgrow, "c2": gen_c2}[w]()
        finally:
            os.chdir(cwd)
    for f in sorted(os.listdir(OUT)):
        if f.startswith("ref_planner"):
            print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main(tuple(sys.argv[1:]) or ("c1", "kgrow", "c2"))
