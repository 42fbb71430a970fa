"""NumPy/SciPy restatement of the CC-MPC prediction -> constraint path.

TEST INFRASTRUCTURE (see oracle/__init__.py).  float64 everywhere, single
thread, written to follow the reference's *arithmetic and operation order*;
every function cites the reference lines it restates (paths relative to
/root/reference; "v8ideal" = collect/in_simulation/midlevel/v8ideal/__init__.py,
"makeconstraint" = collect/in_simulation/midlevel/v8ideal/makeconstraint.py).

Pinning: the makeconstraint.* and npu.vertices_of_bboxes restatements are
checked against tests/golden/ref_*.npz, produced by running the reference's
own functions here (oracle/make_golden.py).  The loop drivers, L4, grouping,
cost and count restatements are "parity unpinned" (their source modules do not
import in this environment).
"""
from __future__ import annotations

import math

import numpy as np
import scipy.linalg
import scipy.spatial
import scipy.stats

# Constants the reference hard-codes on this path.
R_EV_OV = 3.4            # v8ideal:795, :1394  "EV radius + OV radius"
TARGET_P = 0.9999        # v8ideal:912
EPS_TOTAL = 0.05         # v8ideal:2920
BBOX_DEFAULT = (3.7, 1.79)   # v8ideal:792 truck['d']
GOAL = (167.174698, -81.759842)  # dataForCost_ref.py:209


# --------------------------------------------------------------------------
# makeconstraint.py
# --------------------------------------------------------------------------
def mvoe(Sigma1, Sigma2, tol=1e-8, maxiter=1000, return_iters=False):
    """Outer ellipsoid of the Minkowski sum of two centred ellipsoids.

    makeconstraint:7-38.  lambda = eig(Sigma1^-1 Sigma2).real (:21); beta
    starts at 1 and is replaced by sqrt(sum 1/(1+beta lam) / sum lam/(1+beta lam))
    until two successive iterates differ by < tol (:24-32);
    Q = (1+1/beta) Sigma1 + (1+beta) Sigma2 (:35).
    """
    Sigma1 = np.asarray(Sigma1, dtype=np.float64)
    Sigma2 = np.asarray(Sigma2, dtype=np.float64)
    lam = np.linalg.eigvals(scipy.linalg.solve(Sigma1, Sigma2)).real
    beta = 1.0
    iters = 0
    for _ in range(maxiter):
        iters += 1
        w = 1.0 + beta * lam
        nxt = np.sqrt(np.sum(1.0 / w) / np.sum(lam / w))
        close = abs(nxt - beta) < tol
        beta = nxt
        if close:
            break
    Q = (1 + 1.0 / beta) * Sigma1 + (1 + beta) * Sigma2
    if return_iters:
        return beta, Q, iters
    return beta, Q


def pair_moments(p_t_tau):
    """Conditional-moment split of one (t, tau) pair.

    makeconstraint:41-70.  cov = np.cov of the 4 stacked rows [x_t, y_t, x_tau,
    y_tau] (ddof = 1) (:55); Sigma_mu = C Sigma_tau^-1 C^T (:65);
    Sigma_infer = Sigma_t - Sigma_mu (:68).  Raises LinAlgError for singular
    Sigma_tau exactly as np.linalg.inv does.
    """
    c = np.cov(p_t_tau)
    C = c[0:2, 2:4]
    Ct = c[2:4, 0:2]
    S_t = c[0:2, 0:2]
    S_tau = c[2:4, 2:4]
    S_mu = C @ np.linalg.inv(S_tau) @ Ct
    return S_t - S_mu, S_mu, S_t


def tangents_of_slope(mu, Sigma, c, m):
    """Two lines n.x = d of slope m tangent to (x-mu)^T Sigma^-1 (x-mu) = c^2.

    makeconstraint:134-162.  n = [-m, 1]; q = n^T Sigma n; empty list if q <= 0;
    d = n.mu +- c sqrt(q).
    """
    n = np.array([-m, 1.0])
    q = float(n @ (np.asarray(Sigma) @ n))
    if q <= 0:
        return []
    proj = float(n @ np.asarray(mu))
    half = c * np.sqrt(q)
    return [(n, proj + half), (n, proj - half)]


def point_line_distance(n, d, a):
    """|n.a - d| / ||n||_2  (makeconstraint:165-173)."""
    return abs(float(np.asarray(n) @ np.asarray(a)) - d) / np.linalg.norm(n)


def closest_tangent(mu, Sigma, c, m, a, const_idx=None):
    """Pick the tangent nearer to point a (strict <, ties keep index 0).

    makeconstraint:176-207.  Returns (n, d, idx); 4 x None when no tangent
    exists (:191-192 - the reference's asymmetric arity is kept).
    """
    cand = tangents_of_slope(mu, Sigma, c, m)
    if not cand:
        return (None, None, None, None)
    if const_idx is None:
        idx, best = None, np.inf
        for i, (n, d) in enumerate(cand):
            dist = point_line_distance(n, d, a)
            if dist < best:
                best, idx = dist, i
    else:
        idx = const_idx
    n, d = cand[idx]
    return n, d, idx


def scale_factor(cov_infer, cov_mu, cov_t, Gamma_ijt, target_p=TARGET_P):
    """makeconstraint:259-280."""
    base = np.sqrt(np.linalg.norm(cov_t, "fro"))
    alpha = np.sqrt(np.linalg.norm(cov_infer, "fro")) / base
    beta = np.sqrt(np.linalg.norm(cov_mu, "fro")) / base
    chi_p = scipy.stats.chi2.ppf(target_p, df=2)
    return (np.sqrt(chi_p) * beta / Gamma_ijt + alpha) ** 2


def lower_bound(cov_infer, cov_mu, cov_t, eps_t=0.05 / 8):
    """makeconstraint:282-303."""
    base = np.sqrt(np.linalg.norm(cov_t, "fro"))
    alpha = np.sqrt(np.linalg.norm(cov_infer, "fro")) / base
    beta = np.sqrt(np.linalg.norm(cov_mu, "fro")) / base
    Gamma = scipy.stats.norm.ppf(1 - eps_t)
    return scipy.stats.chi2.cdf((Gamma * (1 - alpha) / beta) ** 2, df=2)


# --------------------------------------------------------------------------
# v8ideal: epsilon allocation, Minkowski and affine half-plane loops
# --------------------------------------------------------------------------
def eps_allocation(K_per_obstacle, eps=EPS_TOTAL):
    """eps_ura[o, k] = eps / O for k < K_o, else 0  (v8ideal:2920-2926)."""
    O = len(K_per_obstacle)
    out = np.zeros((O, int(max(K_per_obstacle)) if O else 0))
    for o in range(O):
        for k in range(int(K_per_obstacle[o])):
            out[o, k] = eps / O
    return out


class _Own:
    """Default function table of the loop drivers: this module's restatements.
    make_golden.py passes the reference's makeconstraint module instead, so the
    same driver loop yields goldens computed by the reference's own functions."""
    predict_moments = staticmethod(lambda rows: pair_moments(rows))
    compute_mvoe = staticmethod(lambda a, b: mvoe(a, b))
    choose_closest_tangent = staticmethod(lambda mu, S, c, m, a: closest_tangent(mu, S, c, m, a))
    compute_lower_bound = staticmethod(lambda ci, cm, ct, e: lower_bound(ci, cm, ct, e))


def minkowski_cell(pos, ref_traj, eps_cell, prediction_horizon,
                   R=R_EV_OV, target_p=TARGET_P, impl=_Own):
    """Half-planes of ONE (obstacle, mode) cell, Minkowski variant.

    v8ideal:893-947 for one (ov_idx, latent_idx).  ``pos`` is the cell's
    (N_k, T, 2) sample tensor; the reference's strided views
    poseData[t::Tpred, c] over the (N_k*T, 2) vstack (:890-902) are exactly
    pos[:, t, c].  For every t >= 1 and tau < t:
      eps_ijt = eps_cell / prediction_horizon (:910); chi2 quantiles (:911-913);
      (Sigma_infer, Sigma_mu, Sigma_t) (:907); Q1 = mvoe(Sigma_infer chi_eps,
      Sigma_mu chi_p) (:915); Q = mvoe(Q1, R^2 I) (:917-918);
      m = -(ref_x - mu_x)/(ref_y - mu_y) (:923); tangent (:924); side (:926-939);
      prob_lower[t] = min over tau of lower_bound (:942-947).
    Returns a dict of arrays indexed by the pair number p = t(t-1)/2 + tau.
    """
    pos = np.asarray(pos, dtype=np.float64)
    T = pos.shape[1]
    P = T * (T - 1) // 2
    out = dict(
        n=np.zeros((P, 2)), d=np.zeros(P), idx=np.zeros(P, np.int32),
        side=np.zeros(P, np.int32), A=np.zeros((P, 2)), b=np.zeros(P),
        Q=np.zeros((P, 2, 2)), Q1=np.zeros((P, 2, 2)),
        beta1=np.zeros(P), beta2=np.zeros(P),
        cov_infer=np.zeros((P, 2, 2)), cov_mu=np.zeros((P, 2, 2)),
        prob_pair=np.zeros(P), prob_lower=np.ones(T), mean=np.zeros((T, 2)),
        t=np.zeros(P, np.int32), tau=np.zeros(P, np.int32),
    )
    for t in range(T):
        x_t = pos[:, t, 0]
        y_t = pos[:, t, 1]
        mean = np.mean([x_t, y_t], axis=1)
        out["mean"][t] = mean
        prob_lower = 1.0
        for tau in range(t):
            p = t * (t - 1) // 2 + tau
            rows = [x_t, y_t, pos[:, tau, 0], pos[:, tau, 1]]
            cov_infer, cov_mu, cov_t = impl.predict_moments(rows)
            eps_ijt = eps_cell / prediction_horizon
            chi_eps = scipy.stats.chi2.ppf(1 - eps_ijt, df=2)
            chi_p = scipy.stats.chi2.ppf(target_p, df=2)
            b1, Q1 = impl.compute_mvoe(cov_infer * chi_eps, cov_mu * chi_p)
            b2, Q = impl.compute_mvoe(Q1, np.identity(2) * R ** 2)
            ref_x = [ref_traj[t][0], ref_traj[t][1]]
            m = -(ref_traj[t][0] - mean[0]) / (ref_traj[t][1] - mean[1])
            n_star, d_star, idx = impl.choose_closest_tangent(mean, Q, 1, m, ref_x)
            if n_star @ mean <= d_star:
                side, A, b = 1, -n_star, -d_star      # row is  n.x >= d
            else:
                side, A, b = -1, n_star, d_star       # row is  n.x <= d
            pl = impl.compute_lower_bound(cov_infer, cov_mu, cov_t, eps_ijt)
            prob_lower = min(prob_lower, pl)
            out["n"][p] = n_star
            out["d"][p] = d_star
            out["idx"][p] = idx
            out["side"][p] = side
            out["A"][p] = A
            out["b"][p] = b
            out["Q"][p] = Q
            out["Q1"][p] = Q1
            out["beta1"][p] = b1
            out["beta2"][p] = b2
            out["cov_infer"][p] = cov_infer
            out["cov_mu"][p] = cov_mu
            out["prob_pair"][p] = pl
            out["t"][p] = t
            out["tau"][p] = tau
        out["prob_lower"][t] = prob_lower
    return out


def affine_cell(pos, ref_traj, eps_cell, Tpred, R=R_EV_OV, impl=_Own):
    """Half-planes of ONE (obstacle, mode) cell, affine (receding-horizon) variant.

    v8ideal:1476-1515.  Per t: eps_ijt = eps_cell/Tpred, Gamma = norm.ppf(1-eps)
    (:1481-1482); mu, Sigma = np.cov([x, y]) (:1485-1493); sqrtm(Sigma) (:1494);
    m (:1496); tangent to the radius-R circle (Sigma = I, c = R) (:1499-1502);
    offset = Gamma ||sqrtm(Sigma) [m, -1]^T||_2 (:1506, :1512); row
    n.x >= d + offset when n.mu <= d, else n.x <= d - offset.
    """
    pos = np.asarray(pos, dtype=np.float64)
    T = pos.shape[1]
    out = dict(n=np.zeros((T, 2)), d=np.zeros(T), idx=np.zeros(T, np.int32),
               side=np.zeros(T, np.int32), offset=np.zeros(T),
               rhs=np.zeros(T), mean=np.zeros((T, 2)), cov=np.zeros((T, 2, 2)))
    for t in range(T):
        eps_ijt = eps_cell / Tpred
        Gamma = scipy.stats.norm.ppf(1 - eps_ijt)
        x = pos[:, t, 0]
        y = pos[:, t, 1]
        mean = np.array([np.mean(x), np.mean(y)])
        cov = np.cov([x, y])
        cov_sqrt = scipy.linalg.sqrtm(cov)
        m = -(ref_traj[t][0] - mean[0]) / (ref_traj[t][1] - mean[1])
        M = np.array([m, -1])
        ref_pose = np.array((ref_traj[t][0], ref_traj[t][1]))
        n_star, d_star, idx = impl.choose_closest_tangent(mean, np.identity(2), R, m, ref_pose)
        offset = Gamma * np.linalg.norm(np.real(cov_sqrt) @ M.T, 2)
        if n_star @ mean <= d_star:
            side, rhs = 1, d_star + offset
        else:
            side, rhs = -1, d_star - offset
        out["n"][t] = n_star
        out["d"][t] = d_star
        out["idx"][t] = idx
        out["side"][t] = side
        out["offset"][t] = offset
        out["rhs"][t] = rhs
        out["mean"][t] = mean
        out["cov"][t] = cov
    return out


# --------------------------------------------------------------------------
# v8ideal: save_moments / predict_ideal
# --------------------------------------------------------------------------
def cell_moments(pos):
    """Moments of one cell as save_moments stores them.

    v8ideal:2586-2606: mean_t = np.mean([x_t, y_t], axis=1); cov_t = np.cov;
    cross[t][tau] = np.cov([x_t, y_t, x_tau, y_tau])[0:2, 2:4] for tau < t.
    Returns mean (T,2), cov (T,2,2), cross (T,T,2,2) (zeros where tau >= t).
    """
    pos = np.asarray(pos, dtype=np.float64)
    T = pos.shape[1]
    mean = np.zeros((T, 2))
    cov = np.zeros((T, 2, 2))
    cross = np.zeros((T, T, 2, 2))
    for t in range(T):
        rows_t = [pos[:, t, 0], pos[:, t, 1]]
        mean[t] = np.mean(rows_t, axis=1)
        cov[t] = np.cov(rows_t)
        for tau in range(t):
            c4 = np.cov(rows_t + [pos[:, tau, 0], pos[:, tau, 1]])
            cross[t, tau] = c4[0:2, 2:4]
    return mean, cov, cross


def save_moments_dict(cells_by_obstacle, T):
    """The nested-list payload save_moments pickles (v8ideal:2580-2610).

    ``cells_by_obstacle[o][k]`` is an (N_k, T', 2) array with T' >= T.
    """
    O = len(cells_by_obstacle)
    Kmax = max(len(c) for c in cells_by_obstacle) if O else 0
    mean_p0p1 = np.empty((O, Kmax, T), dtype=object).tolist()
    cov_p0p1 = np.empty((O, Kmax, T), dtype=object).tolist()
    cross_cov = np.empty((O, Kmax, T, T - 1), dtype=object).tolist() if T > 1 else \
        [[[[] for _ in range(T)] for _ in range(Kmax)] for _ in range(O)]
    for o, cells in enumerate(cells_by_obstacle):
        for k, pos in enumerate(cells):
            mean, cov, cross = cell_moments(np.asarray(pos)[:, :T])
            for t in range(T):
                mean_p0p1[o][k][t] = mean[t]
                cov_p0p1[o][k][t] = cov[t]
                for tau in range(t):
                    cross_cov[o][k][t][tau] = cross[t, tau]
    return {"mean_p0p1": mean_p0p1, "cov_p0p1": cov_p0p1, "cross_cov": cross_cov}


def rollout_params(mean, cov, cross, Tsh):
    """Per-step conditional-Gaussian parameters of predict_ideal.

    v8ideal:2671-2696: A_t = C_{t+1,t} Sigma_t^-1 (:2684-2689);
    cond_cov = Sigma_{t+1} - A C^T (:2693); L = cholesky(cond_cov) (:2696).
    Needs moments up to index Tsh (the reference indexes [t+1], :2677-2681).
    """
    A = np.zeros((Tsh, 2, 2))
    L = np.zeros((Tsh, 2, 2))
    for t in range(Tsh):
        C = cross[t + 1, t]
        A[t] = C @ np.linalg.inv(cov[t])
        L[t] = np.linalg.cholesky(cov[t + 1] - A[t] @ C.T)
    return A, L


def ideal_rollout(mean, cov, cross, Tsh, Z, x0):
    """predict_ideal for one (obstacle, mode), noise supplied.

    v8ideal:2662-2709.  ``x0`` (2,) is the ONE draw from N(mean_0, cov_0) the
    reference broadcasts to all samples (:2664-2665); ``Z`` (n, Tsh, 2) are the
    standard normals of (:2700).  Slot t of the output holds x_{t+1} (:2706;
    the x0 written at :2668 is overwritten at t = 0).
    """
    n = Z.shape[0]
    x = np.asarray(x0, dtype=np.float64).reshape(1, 2) * np.ones((n, 2))
    traj = np.zeros((n, Tsh, 2))
    traj[:, 0, :] = x
    for t in range(Tsh):
        C = cross[t + 1, t]
        A = C @ np.linalg.inv(cov[t])
        cond_mean = mean[t + 1] + ((x - mean[t]) @ A.T)
        cond_cov = cov[t + 1] - A @ C.T
        L = np.linalg.cholesky(cond_cov)
        x = cond_mean + (Z[:, t, :] @ L.T)
        traj[:, t, :] = x
    return traj


def ideal_rollout_sampled(mean, cov, cross, Tsh, n_samples, rng):
    """predict_ideal body drawing its own noise (timed CPU baseline).

    Same statements as the reference (:2664-2706) with a caller-supplied
    Generator/RandomState instead of the unseeded per-step RandomState (:2699).
    """
    x0 = rng.multivariate_normal(mean[0], cov[0], size=1)
    x = x0 * np.ones((n_samples, 2))
    traj = np.zeros((n_samples, Tsh, 2))
    traj[:, 0, :] = x
    for t in range(Tsh):
        C = cross[t + 1, t]
        A = C @ np.linalg.inv(cov[t])
        cond_mean = mean[t + 1] + ((x - mean[t]) @ A.T)
        L = np.linalg.cholesky(cov[t + 1] - A @ C.T)
        Z = rng.standard_normal((n_samples, 2))
        x = cond_mean + (Z @ L.T)
        traj[:, t, :] = x
    return traj


# --------------------------------------------------------------------------
# A9 - sample-level counts and MC cost.  NOT IN THE REFERENCE: restated from
# its constants (R = 3.4, v8ideal:795; footprint R^2 I, :917-918; half-plane
# side, :926-939).  Operation order is part of the contract (no FMA):
#   dx = ex - px ; dy = ey - py ; d2 = dx*dx + dy*dy ; hit = d2 < R*R
#   pen = R*R - d2  (summed over hits only)
#   v = n0*px + n1*py ; viol = side*(v - d) > 0
# --------------------------------------------------------------------------
def collision_counts(pos, ego_xy, R=R_EV_OV):
    """Per-timestep and any-time collision counts and the soft collision cost.

    ``pos`` (N, T, 2) float64 sample tensor of one obstacle (or one mode),
    ``ego_xy`` (T, 2) ego plan.  Returns
      coll (T,) int64  = #{n : |ego_t - p_{n,t}|^2 < R^2}
      any  int64       = #{n : exists t with a hit}
      pen  (T,) float64 = sum_n max(0, R^2 - |ego_t - p_{n,t}|^2)
      d2s  (T,) float64 = sum_n |ego_t - p_{n,t}|^2
    """
    pos = np.asarray(pos, dtype=np.float64)
    ego = np.asarray(ego_xy, dtype=np.float64)
    dx = ego[None, :, 0] - pos[:, :, 0]
    dy = ego[None, :, 1] - pos[:, :, 1]
    d2 = dx * dx + dy * dy
    r2 = R * R
    hit = d2 < r2
    coll = hit.sum(axis=0).astype(np.int64)
    any_ = np.int64(hit.any(axis=1).sum())
    pen = np.where(hit, r2 - d2, 0.0).sum(axis=0)
    return coll, any_, pen, d2.sum(axis=0)


def violation_counts(pos, n, d, side, t_of_row):
    """#{samples on the ego's side of each emitted half-plane}.

    Row r constrains timestep t_of_row[r]; the ego must satisfy
    side*(n.x - d) >= 0, so an obstacle sample centre *violates* the
    separation when side*(n.p - d) > 0.
    """
    pos = np.asarray(pos, dtype=np.float64)
    out = np.zeros(len(d), dtype=np.int64)
    for r in range(len(d)):
        px = pos[:, t_of_row[r], 0]
        py = pos[:, t_of_row[r], 1]
        v = n[r][0] * px + n[r][1] * py
        out[r] = np.count_nonzero(side[r] * (v - d[r]) > 0)
    return out


# --------------------------------------------------------------------------
# utility/npu.py and midlevel/util.py - boxes and the L4 outer approximation
# --------------------------------------------------------------------------
def box_vertices(centers, theta, lw):
    """Corners of N rectangles, shape (N, 4, 2).

    python-utility/utility/utility/npu.py:308-350.  With C = cos, S = sin and
    (l, w) = lw the four corners are centre + 0.5 * [(-C l + S w, -S l - C w),
    (-C l - S w, -S l + C w), (C l - S w, S l + C w), (C l + S w, S l - C w)]
    (the einsum of :343-347 written out; summation order l-term + w-term).
    """
    centers = np.asarray(centers, dtype=np.float64)
    N = centers.shape[0]
    lw = np.asarray(lw, dtype=np.float64)
    lws = np.repeat(lw[None], N, axis=0) if lw.ndim == 1 else lw
    th = np.full(N, theta) if np.ndim(theta) == 0 else np.asarray(theta)
    C, S = np.cos(th), np.sin(th)
    l, w = lws[:, 0], lws[:, 1]
    disp = 0.5 * np.stack([
        -C * l + S * w, -S * l + -C * w,
        -C * l + -S * w, -S * l + C * w,
        C * l + -S * w, S * l + C * w,
        C * l + S * w, S * l + -C * w], axis=1)
    return np.reshape(np.tile(centers, (4,)) + disp, (-1, 4, 2))


def l4_outer(theta, vertices):
    """4-face oriented bounding polytope of all box corners.

    midlevel/util.py:171-200.  At = [I; -I] Rot(theta) with
    Rot = [[cos, sin], [-sin, cos]] (:190-193); b0[j] = max over samples and
    the 4 corners of At[j] . v (:195-199).
    """
    Rt = np.array([[np.cos(theta), np.sin(theta)],
                   [-np.sin(theta), np.cos(theta)]])
    At = np.concatenate((np.eye(2), -np.eye(2))) @ Rt
    per_corner = [np.max(At @ vertices[:, c].T, axis=1) for c in range(4)]
    return At, np.max(np.stack(per_corner), axis=0)


def l4_cell(pos, yaws, bbox=BBOX_DEFAULT):
    """Driver loops for one cell: v8ideal:627-640 and :694-736.

    Returns At (T,4,2), b0 (T,4) and the mean yaw (T,) (:723).
    """
    T = pos.shape[1]
    At = np.zeros((T, 4, 2))
    b0 = np.zeros((T, 4))
    th = np.zeros(T)
    for t in range(T):
        verts = box_vertices(pos[:, t], yaws[:, t], np.asarray(bbox))
        th[t] = np.mean(yaws[:, t])
        At[t], b0[t] = l4_outer(th[t], verts)
    return At, b0, th


# --------------------------------------------------------------------------
# midlevel/ovehicle.py + v8ideal.make_ovehicles - mode grouping and yaws
# --------------------------------------------------------------------------
def yaws_from_positions(ps, pos_last):
    """ovehicle.py:72-76: yaw[n,0] = atan2 towards the first prediction from the
    last observed position, yaw[n,t] = atan2 of consecutive predictions."""
    n_p, T = ps.shape[0], ps.shape[1]
    yaws = np.zeros((n_p, T))
    if n_p == 0:
        return yaws
    yaws[:, 0] = np.arctan2(ps[:, 0, 1] - pos_last[1], ps[:, 0, 0] - pos_last[0])
    for t in range(1, T):
        yaws[:, t] = np.arctan2(ps[:, t, 1] - ps[:, t - 1, 1],
                                ps[:, t, 0] - ps[:, t - 1, 0])
    return yaws


def group_by_mode(predict, z, latent_pmf, pos_last, T, filter_pmf=0.1):
    """Bucket one vehicle's samples by latent and fold rare latents.

    v8ideal:489-494 (bucket by z in sample order) then ovehicle.py:55-113:
    keep latents with pmf > filter_pmf in ascending latent order (:58); their
    centre = mean final position (:80); every other latent's samples go to the
    nearest retained centre by final position (:94-95), appended after the
    retained samples, latents visited in ascending order (:82-106);
    pmf = N_k / N (:109-113).
    Returns (pred_positions[list of (N_k,T,2)], pred_yaws[list of (N_k,T)],
             init_center (K,2), latent_pmf (K,), members[list of index arrays]).
    """
    predict = np.asarray(predict, dtype=np.float64)
    z = np.asarray(z)
    n_states = latent_pmf.size
    buckets = [np.nonzero(z == s)[0] for s in range(n_states)]
    keep = np.argwhere(latent_pmf > filter_pmf).ravel()
    pos_l, yaw_l, mem_l = [], [], []
    centers = np.zeros((keep.size, 2))
    total = 0
    for j, s in enumerate(keep):
        ps = predict[buckets[s]].reshape(-1, T, 2)
        pos_l.append(ps)
        yaw_l.append(yaws_from_positions(ps, pos_last))
        mem_l.append(buckets[s])
        total += ps.shape[0]
        centers[j] = np.mean(ps[:, T - 1], axis=0)
    for s in range(n_states):
        if s in keep:
            continue
        ps = predict[buckets[s]].reshape(-1, T, 2)
        if ps.size == 0:
            continue
        yw = yaws_from_positions(ps, pos_last)
        near = np.argmin(scipy.spatial.distance_matrix(ps[:, T - 1, :], centers), axis=1)
        for j in range(keep.size):
            sel = near == j
            if not sel.any():
                continue
            pos_l[j] = np.concatenate((pos_l[j], ps[sel]))
            yaw_l[j] = np.concatenate((yaw_l[j], yw[sel]))
            mem_l[j] = np.concatenate((mem_l[j], buckets[s][sel]))
        total += ps.shape[0]
    pmf = np.array([p.shape[0] / float(total) for p in pos_l])
    return pos_l, yaw_l, centers, pmf, mem_l


# --------------------------------------------------------------------------
# dataForCost_ref.py - closed-loop cost
# --------------------------------------------------------------------------
def gmm2d_sample(latent_pmf, log_pis, mus, log_sigmas, corrs, u_z, u_c, eps):
    """z ~ Cat(latent_pmf), then GMM2D.rsample per (vehicle, sample, timestep) with SUPPLIED randomness.

    Stands at prediction.py:81-86 (`model.latent.sample_p`, `model.p_y_xz` -> `GMM2D.rsample`); that code is in
    the Trajectron++ submodule which is NOT in /root/reference (fireofearth/Trajectron-plus-plus @ 0e44e82), so
    this follows the PUBLISHED Trajectron++ GMM2D (model/components/gmm2d.py): log_pis normalised by
    logsumexp, sigma = exp(log_sigma), L = [[sx, 0], [rho sy, sy sqrt(1 - rho^2)]], y = mu + L eps for the
    component drawn from Cat(pis).  PARITY UNPINNED (no reference code or vector to check against).
    latent_pmf (V,Z); log_pis (V,Z,T,G) or None; mus, log_sigmas (V,Z,T,G,2); corrs (V,Z,T,G);
    u_z (V,N), u_c (V,N,T) uniforms in [0,1); eps (V,N,T,2) normals.  float32 arithmetic like the model.
    -> z (V,N) int32, predictions (V,N,T,2) float32."""
    f = np.float32
    mus, log_sigmas, corrs = np.asarray(mus, f), np.asarray(log_sigmas, f), np.asarray(corrs, f)
    V, Z, T, G = mus.shape[:4]
    N = u_z.shape[1]
    z = np.zeros((V, N), np.int32)
    pred = np.zeros((V, N, T, 2), f)
    for v in range(V):
        w = np.maximum(np.asarray(latent_pmf[v], np.float64), 0.0)
        cum = (np.cumsum(w) / w.sum()).astype(f)
        zz = np.minimum((np.asarray(u_z[v], f)[:, None] >= cum[None, :]).sum(axis=1), Z - 1)   # first k: u < cum_k
        # bins with zero width after the last positive one cannot be hit; the last bin catches round-off
        z[v] = zz
        for t in range(T):
            if G > 1:
                lp = np.asarray(log_pis[v, zz, t], f)                       # (N,G)
                p = np.exp(lp - lp.max(axis=1, keepdims=True), dtype=f)
                acc = np.cumsum(p, axis=1, dtype=f)
                target = (np.asarray(u_c[v, :, t], f) * acc[:, -1]).astype(f)
                g = np.minimum((target[:, None] >= acc[:, :-1]).sum(axis=1), G - 1)
            else:
                g = np.zeros(N, int)
            mu = mus[v, zz, t, g]
            sg = np.exp(log_sigmas[v, zz, t, g], dtype=f)
            rho = corrs[v, zz, t, g]
            e1, e2 = np.asarray(eps[v, :, t, 0], f), np.asarray(eps[v, :, t, 1], f)
            pred[v, :, t, 0] = mu[:, 0] + sg[:, 0] * e1
            pred[v, :, t, 1] = mu[:, 1] + sg[:, 1] * (rho * e1 + np.sqrt(np.maximum(f(1) - rho * rho, f(0))) * e2)
    return z, pred


def integrate_samples(controls, x0, dt, dynamics="unicycle"):
    """`dynamic.integrate_samples` of model.p_y_xz in PREDICT mode (prediction.py:85-86: the control samples of
    a_dist.rsample() become the `predictions` the planner receives).  The code is in the Trajectron++ submodule
    that is NOT in /root/reference; this follows the PUBLISHED model/dynamics/single_integrator.py and
    unicycle.py (`dynamic` + `integrate_samples`), float32 like the model.  PARITY UNPINNED.
    controls (V,N,T,2) float32; x0 (V,4) = (x, y, phi, v) at the current frame (phi includes the model's learned
    heading correction, v = |velocity|); -> positions (V,N,T,2) float32."""
    f = np.float32
    u = np.asarray(controls, f)
    x0 = np.asarray(x0, f)
    V, N, T, _ = u.shape
    dt = f(dt)
    out = np.zeros((V, N, T, 2), f)
    if dynamics == "single_integrator":
        x = np.broadcast_to(x0[:, None, 0], (V, N)).astype(f)
        y = np.broadcast_to(x0[:, None, 1], (V, N)).astype(f)
        for t in range(T):
            x = x + u[:, :, t, 0] * dt
            y = y + u[:, :, t, 1] * dt
            out[:, :, t, 0], out[:, :, t, 1] = x, y
        return out
    if dynamics != "unicycle":
        raise ValueError(dynamics)
    x = np.broadcast_to(x0[:, None, 0], (V, N)).astype(f)
    y = np.broadcast_to(x0[:, None, 1], (V, N)).astype(f)
    phi = np.broadcast_to(x0[:, None, 2], (V, N)).astype(f)
    v = np.broadcast_to(x0[:, None, 3], (V, N)).astype(f)
    for t in range(T):
        dphi, a = u[:, :, t, 0], u[:, :, t, 1]
        small = np.abs(dphi) <= f(1e-2)
        w = np.where(small, f(1), dphi).astype(f)
        phin = phi + w * dt
        s0, c0, s1, c1 = np.sin(phi), np.cos(phi), np.sin(phin), np.cos(phin)
        ds, dc = (s1 - s0) / w, (c1 - c0) / w
        aw = a / w
        x1 = x + aw * dc + v * ds + aw * s1 * dt
        y1 = y - v * dc + aw * ds - aw * c1 * dt
        x2 = x + v * c0 * dt + (a / f(2)) * c0 * (dt * dt)
        y2 = y + v * s0 * dt + (a / f(2)) * s0 * (dt * dt)
        x = np.where(small, x2, x1).astype(f)
        y = np.where(small, y2, y1).astype(f)
        phi = np.where(small, phi, phi + dphi * dt).astype(f)
        v = (v + a * dt).astype(f)
        out[:, :, t, 0], out[:, :, t, 1] = x, y
    return out


def episode_cost(X_star, U_star, goal=GOAL):
    """dataForCost_ref.py:198-232.

    costX = 6 (|x_F - g_x|^2 + |y_F - g_y|^2) on the last stored state;
    costU is *overwritten* by U[tau]^T R1 U[tau] in the first loop (:223-224)
    so only tau = H-2 survives, then the change penalties of tau < H-2 are
    added (:226-227).  With H < 2 the reference raises (costU undefined).
    """
    R1 = np.array([[0.5, 0.2], [0.2, 1.0]])
    R2 = np.array([[0.5, 0.1], [0.1, 2.0]])
    g = np.array(goal)
    H = len(U_star)
    costX = 6.0 * np.absolute(X_star[-1][0] - g[0]) ** 2 \
        + 6.0 * np.absolute(X_star[-1][1] - g[1]) ** 2
    if H < 2:
        raise UnboundLocalError("costU referenced before assignment")
    for tau in range(0, H - 1):
        u = np.asarray(U_star[tau], dtype=np.float64)
        costU = u.T @ R1 @ u
    for tau in range(0, H - 2):
        du = np.asarray(U_star[tau + 1], dtype=np.float64) - np.asarray(U_star[tau], dtype=np.float64)
        costU = costU + du.T @ R2 @ du
    return costX + costU, costU


# --------------------------------------------------------------------------
# closed forms used by the CUDA path, checked against scipy in the tests
# --------------------------------------------------------------------------
def chi2_ppf_df2(p):
    """chi2.ppf(p, df=2) = -2 ln(1 - p)."""
    return -2.0 * math.log1p(-p)


def chi2_cdf_df2(x):
    """chi2.cdf(x, df=2) = 1 - exp(-x/2)."""
    return -math.expm1(-0.5 * x)
