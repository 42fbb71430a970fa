"""Drop-in twin of the reference's ``makeconstraint`` module on libccmpc.

Same function names, argument meaning, return types and error behaviour as
collect/in_simulation/midlevel/v8ideal/makeconstraint.py, so that
``from ccmpc_b200 import makeconstraint`` replaces
``from . import makeconstraint`` in v8ideal/__init__.py.  Each call ships its
NumPy arguments to the GPU, runs the sm_100a kernel behind the C ABI on a
batch of one, and returns NumPy values; the batched forms (leading dims) live
in :mod:`ccmpc_b200.ops` and are what the planner-level builder
(:mod:`ccmpc_b200.midlevel`) uses.  There is no CPU path.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops

__all__ = ["compute_mvoe", "predict_moments", "tangent_lines_of_slope_m", "distance_point_to_line",
           "choose_closest_tangent", "compute_scale", "compute_lower_bound"]


def _dev():
    if not torch.cuda.is_available():
        raise RuntimeError("ccmpc_b200 needs a CUDA device (there is no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def _sym3(S, name):
    S = np.asarray(S, dtype=np.float64)
    if S.shape != (2, 2):
        raise ValueError(f"{name} must be 2x2")
    # the reference feeds symmetric covariances; the (0,1) entry is the one np.cov / eigvals read first
    return torch.tensor([[S[0, 0], S[0, 1], S[1, 1]]], dtype=torch.float64, device=_dev())


def _full(q3):
    q = q3.cpu().numpy().reshape(3)
    return np.array([[q[0], q[1]], [q[1], q[2]]])


def compute_mvoe(Sigma1, Sigma2, tol=1e-8, maxiter=1000):
    """makeconstraint.py:7-38.  Returns (beta_star, Q_star (2,2)).

    A singular Sigma1 raises ``np.linalg.LinAlgError`` the way
    ``scipy.linalg.solve`` does at :20."""
    beta, q, _, status = ops.mvoe_fit(_sym3(Sigma1, "Sigma1"), _sym3(Sigma2, "Sigma2"), tol=tol, maxiter=maxiter)
    if int(status.item()) == ops.ST_SINGULAR:
        raise np.linalg.LinAlgError("Matrix is singular.")
    return float(beta.item()), _full(q)


def predict_moments(p_t_tau):
    """makeconstraint.py:41-70.  ``p_t_tau`` = [x_t, y_t, x_tau, y_tau] (four (N,) arrays).

    Returns (cov_infer, cov_mu, cov_t), each (2,2).  Singular Sigma_tau raises
    ``np.linalg.LinAlgError`` as ``np.linalg.inv`` does at :65."""
    rows = [np.asarray(r, dtype=np.float64).ravel() for r in p_t_tau]
    if len(rows) != 4:
        raise ValueError("p_t_tau must hold 4 rows")
    n = rows[0].size
    # (N, T=2, 2) sample tensor with time 0 = tau, time 1 = t  (pair p = 0 of a T = 2 cell)
    pos = np.empty((n, 2, 2))
    pos[:, 1, 0], pos[:, 1, 1], pos[:, 0, 0], pos[:, 0, 1] = rows
    dev = _dev()
    samples = torch.from_numpy(pos).to(dev)
    offsets = torch.tensor([0, n], dtype=torch.int64, device=dev)
    pivot = ops.cell_pivot(samples, offsets)
    sums = ops.sample_sums(samples, offsets, pivot)
    ref = torch.zeros((2, 2), dtype=torch.float64, device=dev)
    eps = torch.full((1,), 0.05, dtype=torch.float64, device=dev)
    out = ops.minkowski_alloc(1, 2, dev, fields=("cov_infer", "cov_mu", "status"))
    ops.minkowski_halfplanes(sums, pivot, ref, eps, 8, out=out)
    mom = ops.moments_from_sums(sums, pivot, want_cross=False, want_counts=False, want_yaw=False)
    if int(out["status"].reshape(-1)[0].item()) == ops.ST_SINGULAR:
        raise np.linalg.LinAlgError("Singular matrix")
    cov_t = mom["cov"][0, 1].cpu().numpy()
    return _full(out["cov_infer"].reshape(3)), _full(out["cov_mu"].reshape(3)), cov_t


def _tangent(mu, Sigma, c, m, a, const_idx):
    dev = _dev()
    f = lambda v: torch.tensor(np.asarray(v, dtype=np.float64).reshape(1, -1), dtype=torch.float64, device=dev)
    ci = None if const_idx is None else torch.tensor([int(const_idx)], dtype=torch.int32, device=dev)
    d12, dist, d_star, idx = ops.tangent(f(mu), _sym3(Sigma, "Sigma"), f(c).reshape(1), f(m).reshape(1), f(a), ci)
    return d12.cpu().numpy()[0], dist.cpu().numpy()[0], float(d_star.item()), int(idx.item())


def tangent_lines_of_slope_m(mu, Sigma, c, m):
    """makeconstraint.py:134-162.  List of (n, d): two tangents, or [] when n^T Sigma n <= 0."""
    d12, _, _, idx = _tangent(mu, Sigma, c, m, [0.0, 0.0], None)
    if idx < 0:
        return []
    n = np.array([-m, 1.0])
    return [(n, float(d12[0])), (n, float(d12[1]))]


def distance_point_to_line(n, d, a):
    """makeconstraint.py:165-173: |n.a - d| / ||n||.  The kernel normalises lines as n = [-m, 1];
    a general normal is rescaled to that form first (the distance is scale invariant)."""
    n = np.asarray(n, dtype=np.float64)
    if n[1] == 0.0:
        raise ValueError("distance_point_to_line: the CUDA path represents lines as n = [-m, 1] (n[1] != 0)")
    m = -n[0] / n[1]
    # zero-radius "ellipse" through mu with n.mu = d/n[1]: d_plus = d_minus = d / n[1]
    mu = np.array([0.0, d / n[1]])
    _, dist, _, _ = _tangent(mu, np.array([[1.0, 0.0], [0.0, 1.0]]), 0.0, m, a, None)
    return float(dist[0])


def choose_closest_tangent(mu, Sigma, c, m, a, const_idx=None):
    """makeconstraint.py:176-207.  (n_star, d_star, best_idx); 4 x None when no tangent exists (:191-192)."""
    _, _, d_star, idx = _tangent(mu, Sigma, c, m, a, const_idx)
    if idx < 0:
        return None, None, None, None
    return np.array([-m, 1.0]), d_star, idx


def _triples(*mats):
    return [_sym3(M, "cov") for M in mats]


def compute_scale(cov_infer, cov_mu, cov_t, Gamma_ijt, target_p=0.9999):
    """makeconstraint.py:259-280."""
    ci, cm, ct = _triples(cov_infer, cov_mu, cov_t)
    g = torch.tensor([float(Gamma_ijt)], dtype=torch.float64, device=ci.device)
    scale, _ = ops.scale_lower(ci, cm, ct, gamma=g, target_p=target_p)
    return float(scale.item())


def compute_lower_bound(cov_infer, cov_mu, cov_t, eps_t=0.05 / 8):
    """makeconstraint.py:282-303."""
    ci, cm, ct = _triples(cov_infer, cov_mu, cov_t)
    e = torch.tensor([float(eps_t)], dtype=torch.float64, device=ci.device)
    _, lower = ops.scale_lower(ci, cm, ct, eps=e)
    return float(lower.item())
