"""NumPy restatement of the PLANNER-LEVEL drivers around the per-cell loops of restate.py.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Pinned by tests/golden/ref_planner_*.npz, which the
reference's own methods produced (oracle/make_golden_planner.py).

  minkowski_step   compute_obstacle_constraints_GMM_Minkowski_idealprediction   v8ideal:781-964
  affine_step      compute_obstacle_constraints_GMM_affine                      v8ideal:1378-1539
  predict_ideal    predict_ideal with supplied noise                            v8ideal:2620-2711
"""
from __future__ import annotations

import numpy as np

from . import restate


def ov_state(cells, yaws, Tpred):
    """[tau+1|tau] bookkeeping (v8ideal:864-875) and OV_injuction (:831-842, the LAST latent decides)."""
    O = len(cells)
    Kmax = max(len(c) for c in cells)
    mean = np.full((3, O, Kmax), np.nan)
    cov = np.full((3, O, Kmax), np.nan)
    inj = [None] * O
    for o in range(O):
        for k, pos in enumerate(cells[o]):
            x, y, yw = pos[:, 0, 0], pos[:, 0, 1], yaws[o][k][:, 0]
            mean[:, o, k] = np.mean(x), np.mean(y), np.mean(yw)
            cov[:, o, k] = np.cov(x), np.cov(y), np.cov(yw)
            inj[o] = not (np.mean(x) >= 190 or np.mean(y) <= -80)
    return mean, cov, inj


def l4_union(cells, yaws, bboxes, Tpred):
    """__compute_vertices + __compute_overapproximations (v8ideal:627-640, :694-736): dense
    A_union (T, Kmax, O, 4, 2), b_union (T, Kmax, O, 4), NaN in unused slots."""
    O = len(cells)
    Kmax = max(len(c) for c in cells)
    A = np.full((Tpred, Kmax, O, 4, 2), np.nan)
    b = np.full((Tpred, Kmax, O, 4), np.nan)
    for o in range(O):
        for k, pos in enumerate(cells[o]):
            At, b0, _ = restate.l4_cell(pos, yaws[o][k], bboxes[o])
            A[:, k, o], b[:, k, o] = At, b0
    return A, b


def predict_ideal(moments, K_per_obstacle, Tsh, x0, Z):
    """v8ideal:2641-2711 with the noise supplied: ``x0[c]`` (2,) and ``Z[c]`` (n, Tsh, 2) per cell c in
    (obstacle, latent) order.  ``moments`` is the nested dict of save_moments.  data_idx fallback :2652-2656."""
    out, c = {}, 0
    for o, K_o in enumerate(K_per_obstacle):
        out[o] = {}
        n_latent = len(moments["mean_p0p1"][o])
        for k in range(int(K_o)):
            di = k if k < n_latent else max(n_latent - 1, 0)
            T_saved = len(moments["mean_p0p1"][o][di])
            mean = np.stack([moments["mean_p0p1"][o][di][t] for t in range(T_saved)])
            cov = np.stack([moments["cov_p0p1"][o][di][t] for t in range(T_saved)])
            cross = np.zeros((T_saved, T_saved, 2, 2))
            for t in range(T_saved):
                for tau in range(t):
                    cross[t, tau] = moments["cross_cov"][o][di][t][tau]
            out[o][k] = restate.ideal_rollout(mean, cov, cross, Tsh, Z[c], x0[c])
            c += 1
    return out


def minkowski_step(cells, yaws, bboxes, ref_traj, eps_ura, Tsh, Tpred, moments_prev=None, x0=None, Z=None):
    """One call of the Minkowski builder.  ``cells[o][k]`` (N_k, Tpred, 2), ``yaws[o][k]`` (N_k, Tpred).
    Rows come back flattened in the reference's emission order (o, k, t, tau)."""
    O = len(cells)
    K = [len(c) for c in cells]
    T = int(Tsh)
    work = cells
    if T < Tpred:                                                     # :824-825, :885-888
        ideal = predict_ideal(moments_prev, K, T, x0, Z)
        work = [[ideal[o][k] for k in range(K[o])] for o in range(O)]
    rows = dict(n=[], t=[], side=[], rhs=[], idx=[])
    prob_lower_save = [None] * T
    for o in range(O):
        for k in range(K[o]):
            r = restate.minkowski_cell(work[o][k][:, :T], ref_traj, eps_ura[o, k], Tpred)
            rows["n"].append(r["n"]); rows["t"].append(r["t"]); rows["side"].append(r["side"])
            rows["rhs"].append(r["d"]); rows["idx"].append(r["idx"])
            prob_lower_save = list(r["prob_lower"])                   # :947 overwritten per cell: last wins
    ov_mean, ov_cov, inj = ov_state(cells, yaws, Tpred)
    A_union, b_union = l4_union(cells, yaws, bboxes, Tpred)
    moments = restate.save_moments_dict([[w[:, :T] for w in wo] for wo in work], T)       # :960
    return dict(n=np.concatenate(rows["n"]), t=np.concatenate(rows["t"]), side=np.concatenate(rows["side"]),
                rhs=np.concatenate(rows["rhs"]), idx=np.concatenate(rows["idx"]),
                A_union=A_union, b_union=b_union, ov_mean=ov_mean, ov_cov=ov_cov,
                OVconstraint=any(bool(v) for v in inj), prob_lower_save=np.asarray(prob_lower_save),
                moments=moments, work=work)


def affine_step(cells, yaws, bboxes, ref_traj, eps_ura, Tsh, Tpred):
    """One call of the affine builder (out_junction is hard-coded False, :1424: OVconstraint is always False)."""
    O = len(cells)
    T = int(Tsh)
    rows = dict(n=[], t=[], side=[], rhs=[], idx=[])
    for o in range(O):
        for k in range(len(cells[o])):
            r = restate.affine_cell(cells[o][k][:, :T], ref_traj, eps_ura[o, k], Tpred)
            rows["n"].append(r["n"]); rows["t"].append(np.arange(T, dtype=np.int32)); rows["side"].append(r["side"])
            rows["rhs"].append(r["rhs"]); rows["idx"].append(r["idx"])
    ov_mean, ov_cov, _ = ov_state(cells, yaws, Tpred)
    A_union, b_union = l4_union(cells, yaws, bboxes, Tpred)
    return dict(n=np.concatenate(rows["n"]), t=np.concatenate(rows["t"]), side=np.concatenate(rows["side"]),
                rhs=np.concatenate(rows["rhs"]), idx=np.concatenate(rows["idx"]),
                A_union=A_union, b_union=b_union, ov_mean=ov_mean, ov_cov=ov_cov, OVconstraint=False)


def replay_noise(seed, n_cells, Tsh, n_ideal, rs=None):
    """The normals the seeded np.random shim of oracle/ref_planner.py handed to predict_ideal: per cell one
    ``standard_normal((1, 2))`` (inside multivariate_normal, :2664) then Tsh blocks ``standard_normal((n, 2))``
    (:2700).  Returns (Z (C, n, Tsh, 2), block sums (C*Tsh,), the RandomState for the next step)."""
    rs = np.random.RandomState(seed) if rs is None else rs
    Z = np.empty((n_cells, n_ideal, Tsh, 2))
    sums = []
    for c in range(n_cells):
        rs.standard_normal((1, 2))
        for t in range(Tsh):
            blk = rs.standard_normal((n_ideal, 2))
            Z[c, :, t] = blk
            sums.append(blk.sum())
    return Z, np.asarray(sums), rs
