"""Batched tensor twins of the reference's constraint-path functions.

Thin, allocation-only wrappers over the C ABI of libccmpc.so
(include/ccmpc.h): every function takes torch CUDA tensors, allocates the
outputs, and enqueues the kernels on the *current* torch CUDA stream without
synchronising.  There is no CPU path: a non-CUDA tensor raises.

Reference call sites (relative to the reference root; "makeconstraint" =
collect/in_simulation/midlevel/v8ideal/makeconstraint.py, "v8ideal" =
collect/in_simulation/midlevel/v8ideal/__init__.py) are cited per function.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _lib

R_EV_OV = 3.4        # v8ideal:795
TARGET_P = 0.9999    # v8ideal:912
MAX_T = 16           # CCMPC_MAX_T

ST_OK, ST_SINGULAR, ST_NO_TANGENT, ST_NOT_CONVERGED, ST_DEGENERATE_SLOPE = range(5)


def _p(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


try:                                   # the raw handle of the current stream without building a Stream object
    _raw_stream, _cur_dev = torch._C._cuda_getCurrentRawStream, torch._C._cuda_getDevice
except AttributeError:                 # (private helpers of torch; ~10 x cheaper per call than the public path)
    _raw_stream = _cur_dev = None


def _stream():
    if _raw_stream is not None:
        return ctypes.c_void_p(_raw_stream(_cur_dev()))
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _chk(t: Optional[torch.Tensor], dtype, name: str, shape_tail=None):
    if t is None:
        return None
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError(f"{name}: expected a CUDA tensor (libccmpc has no CPU path)")
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected dtype {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name}: tensor must be contiguous")
    if t.data_ptr() % 16:
        # the kernels move double2 / float4 / 16-byte staged records: a contiguous view that starts at an odd
        # element (e.g. sigma[1:] of an [M,3] tensor) would fault with a sticky misaligned-address error
        raise ValueError(f"{name}: data pointer must be 16-byte aligned (pass a fresh tensor, not an offset view)")
    if shape_tail is not None and tuple(t.shape[-len(shape_tail):]) != tuple(shape_tail):
        raise ValueError(f"{name}: trailing shape {tuple(t.shape)} does not end with {tuple(shape_tail)}")
    return t


def _new(shape, dtype, like: torch.Tensor):
    return torch.empty(shape, dtype=dtype, device=like.device)


F64, I64, I32, I8, U8 = torch.float64, torch.int64, torch.int32, torch.int8, torch.uint8


# ---------------------------------------------------------------------------
# K3  compute_mvoe (makeconstraint:7-38), batched
# ---------------------------------------------------------------------------
def mvoe_fit(sigma1, sigma2=None, footprint_r2: float = R_EV_OV ** 2, tol: float = 1e-8, maxiter: int = 1000,
             accel: bool = False, want_iters: bool = True):
    """sigma1 [..., 3] sym3; sigma2 [..., 3] or None (=> footprint_r2 * I).
    Returns (beta [...], q [..., 3], iters [...] int32 | None, status [...] uint8)."""
    lib = _lib.load()
    _chk(sigma1, F64, "sigma1", (3,))
    if sigma2 is not None:
        _chk(sigma2, F64, "sigma2", (3,))
        if sigma2.shape != sigma1.shape:
            raise ValueError("sigma1 and sigma2 must have the same shape")
    lead = sigma1.shape[:-1]
    M = sigma1.numel() // 3
    beta = _new(lead, F64, sigma1)
    q = _new(tuple(lead) + (3,), F64, sigma1)
    iters = _new(lead, I32, sigma1) if want_iters else None
    status = _new(lead, U8, sigma1)
    _lib.check(lib.ccmpc_mvoe_fit(_p(sigma1), _p(sigma2), float(footprint_r2), M, float(tol), int(maxiter),
                                  int(accel), _p(beta), _p(q), _p(iters), _p(status), _stream()),
               "ccmpc_mvoe_fit")
    return beta, q, iters, status


def mvoe_halfplane_sweep(cov, mu, ref, chi_scale: float, footprint_r2: float = R_EV_OV ** 2, tol: float = 1e-8,
                         maxiter: int = 1000, accel: bool = False, out=None):
    """cov [S,O,T,K,3], mu [S,O,T,K,2], ref [S,T,2] -> (q, nd, idx_side, status) (v8ideal:915-939 per cell)."""
    lib = _lib.load()
    _chk(cov, F64, "cov", (3,))
    _chk(mu, F64, "mu", (2,))
    _chk(ref, F64, "ref", (2,))
    S, O, T, K = cov.shape[:4]
    if tuple(mu.shape) != (S, O, T, K, 2) or tuple(ref.shape) != (S, T, 2):
        raise ValueError("mvoe_halfplane_sweep: shape mismatch")
    if out is None:
        out = (_new((S, O, T, K, 3), F64, cov), _new((S, O, T, K, 3), F64, cov),
               _new((S, O, T, K), I8, cov), _new((S, O, T, K), U8, cov))
    q, nd, idx_side, status = out
    _lib.check(lib.ccmpc_mvoe_halfplane_sweep(_p(cov), _p(mu), _p(ref), S, O, T, K, float(chi_scale),
                                              float(footprint_r2), float(tol), int(maxiter), int(accel),
                                              _p(q), _p(nd), _p(idx_side), _p(status), _stream()),
               "ccmpc_mvoe_halfplane_sweep")
    return q, nd, idx_side, status


# ---------------------------------------------------------------------------
# K4  tangent lines (makeconstraint:134-207), batched
# ---------------------------------------------------------------------------
def tangent(mu, sigma, c, m, a, const_idx=None):
    """-> d12 [M,2] (d+, d-), dist [M,2], d_star [M], idx [M] int32 (-1: no tangent)."""
    lib = _lib.load()
    _chk(mu, F64, "mu", (2,)); _chk(sigma, F64, "sigma", (3,)); _chk(c, F64, "c"); _chk(m, F64, "m")
    _chk(a, F64, "a", (2,)); _chk(const_idx, I32, "const_idx")
    M = m.numel()
    d12 = _new((M, 2), F64, mu); dist = _new((M, 2), F64, mu)
    d_star = _new((M,), F64, mu); idx = _new((M,), I32, mu)
    _lib.check(lib.ccmpc_tangent(_p(mu), _p(sigma), _p(c), _p(m), _p(a), _p(const_idx), M, _p(d12), _p(dist),
                                 _p(d_star), _p(idx), _stream()), "ccmpc_tangent")
    return d12, dist, d_star, idx


def scale_lower(cov_infer, cov_mu, cov_t, gamma=None, eps=None, target_p: float = TARGET_P):
    """compute_scale (makeconstraint:259-280) and compute_lower_bound (:282-303), batched over [M,3] triples."""
    lib = _lib.load()
    for t, n in ((cov_infer, "cov_infer"), (cov_mu, "cov_mu"), (cov_t, "cov_t")):
        _chk(t, F64, n, (3,))
    _chk(gamma, F64, "gamma"); _chk(eps, F64, "eps")
    M = cov_infer.numel() // 3
    scale = _new((M,), F64, cov_infer) if gamma is not None else None
    lower = _new((M,), F64, cov_infer) if eps is not None else None
    _lib.check(lib.ccmpc_scale_lower(_p(cov_infer), _p(cov_mu), _p(cov_t), _p(gamma), _p(eps), float(target_p), M,
                                     _p(scale), _p(lower), _stream()), "ccmpc_scale_lower")
    return scale, lower


# ---------------------------------------------------------------------------
# K2 + K6  one read of the sample tensor -> all sums over samples
# ---------------------------------------------------------------------------
def sums_len(T: int) -> int:
    n = _lib.load().ccmpc_sums_len(int(T))
    if n < 0:
        raise ValueError(f"T = {T} outside 1..{MAX_T}")
    return int(n)


def sums_layout(T: int) -> dict:
    """Offsets into the packed sums vector (mirror of SumsLayout in csrc/common.cuh)."""
    D = 2 * T
    o_s1 = 1
    o_s2 = o_s1 + D
    o_coll = o_s2 + D * (D + 1) // 2
    o_any = o_coll + T
    o_pen = o_any + 1
    o_d2s = o_pen + T
    o_yaw = o_d2s + T
    o_yaw2 = o_yaw + T
    return dict(n=0, s1=o_s1, s2=o_s2, coll=o_coll, any=o_any, pen=o_pen, d2s=o_d2s, yaw=o_yaw, yaw2=o_yaw2,
                len=o_yaw2 + T)


def sample_sums_workspace(n_total: int, C: int, T: int, device) -> torch.Tensor:
    nbytes = _lib.load().ccmpc_sample_sums_workspace_bytes(int(n_total), int(C), int(T))
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)


def cell_pivot(samples, offsets, out=None):
    """pivot[c] = first sample of cell c (zeros for an empty cell): the shift of the one-pass power sums."""
    lib = _lib.load()
    _chk(samples, F64, "samples", (2,)); _chk(offsets, I64, "offsets")
    T = samples.shape[1]
    C = offsets.numel() - 1
    if out is None:
        out = _new((C, T, 2), F64, samples)
    _lib.check(lib.ccmpc_cell_pivot(_p(samples), _p(offsets), C, T, _p(out), _stream()), "ccmpc_cell_pivot")
    return out


def sample_sums(samples, offsets, pivot, ego=None, yaws=None, R: float = R_EV_OV, workspace=None, out=None):
    """samples [N,T,2], offsets [C+1] int64, pivot [C,T,2], ego [T,2] | [C,T,2] | None, yaws [N,T] | None
    -> sums [C, sums_len(T)] (np.mean/np.cov of every (t,tau) pair, makeconstraint:55; counts A9)."""
    lib = _lib.load()
    _chk(samples, F64, "samples", (2,)); _chk(offsets, I64, "offsets"); _chk(pivot, F64, "pivot", (2,))
    _chk(ego, F64, "ego", (2,)); _chk(yaws, F64, "yaws")
    N, T = samples.shape[0], samples.shape[1]
    C = offsets.numel() - 1
    if tuple(pivot.shape) != (C, T, 2):
        raise ValueError("pivot must be [C,T,2]")
    stride = 0
    if ego is not None:
        if tuple(ego.shape) == (T, 2):
            stride = 0
        elif tuple(ego.shape) == (C, T, 2):
            stride = 2 * T
        else:
            raise ValueError("ego must be [T,2] or [C,T,2]")
    if yaws is not None and tuple(yaws.shape) != (N, T):
        raise ValueError("yaws must be [N,T]")
    if workspace is None:
        workspace = sample_sums_workspace(N, C, T, samples.device)
    if out is None:
        out = _new((C, sums_len(T)), F64, samples)
    _lib.check(lib.ccmpc_sample_sums(_p(samples), _p(offsets), N, C, T, _p(pivot), _p(ego), stride, float(R),
                                     _p(yaws), _p(out), _p(workspace), workspace.numel(), _stream()),
               "ccmpc_sample_sums")
    return out


def moments_from_sums(sums, pivot, want_cross=True, want_counts=True, want_yaw=True, out=None):
    """-> dict(mean [C,T,2], cov [C,T,2,2], cross [C,T,T,2,2], counts [C,T+1] int64, yaw_mean [C,T], yaw_var [C,T])
    in the save_moments layout (v8ideal:2586-2606)."""
    lib = _lib.load()
    _chk(sums, F64, "sums"); _chk(pivot, F64, "pivot", (2,))
    C, T = pivot.shape[0], pivot.shape[1]
    if out is None:
        out = dict(mean=_new((C, T, 2), F64, sums), cov=_new((C, T, 2, 2), F64, sums),
                   cross=_new((C, T, T, 2, 2), F64, sums) if want_cross else None,
                   counts=_new((C, T + 1), I64, sums) if want_counts else None,
                   yaw_mean=_new((C, T), F64, sums) if want_yaw else None,
                   yaw_var=_new((C, T), F64, sums) if want_yaw else None)
    _lib.check(lib.ccmpc_moments_from_sums(_p(sums), _p(pivot), C, T, _p(out["mean"]), _p(out["cov"]),
                                           _p(out.get("cross")), _p(out.get("counts")), _p(out.get("yaw_mean")),
                                           _p(out.get("yaw_var")), _stream()), "ccmpc_moments_from_sums")
    return out


MINK_FIELDS = ("nd", "ab", "idx_side", "q", "q1", "betas", "cov_infer", "cov_mu", "prob_pair", "prob_lower", "mean",
               "status")


def minkowski_alloc(C: int, T: int, device, fields=MINK_FIELDS) -> dict:
    P = T * (T - 1) // 2
    shapes = dict(nd=((C, P, 3), F64), ab=((C, P, 3), F64), idx_side=((C, P), I8), q=((C, P, 3), F64),
                  q1=((C, P, 3), F64), betas=((C, P, 2), F64), cov_infer=((C, P, 3), F64), cov_mu=((C, P, 3), F64),
                  prob_pair=((C, P), F64), prob_lower=((C, T), F64), mean=((C, T, 2), F64), status=((C, P), U8))
    return {k: torch.empty(shapes[k][0], dtype=shapes[k][1], device=device) for k in fields}


def minkowski_halfplanes(sums, pivot, ref, eps_cell, prediction_horizon: int, R: float = R_EV_OV,
                         target_p: float = TARGET_P, tol: float = 1e-8, maxiter: int = 1000, accel: bool = False,
                         out: Optional[dict] = None):
    """Minkowski half-plane loop of one MPC step (v8ideal:893-947) for every cell and (t, tau) pair."""
    lib = _lib.load()
    _chk(sums, F64, "sums"); _chk(pivot, F64, "pivot", (2,)); _chk(ref, F64, "ref", (2,)); _chk(eps_cell, F64, "eps")
    C, T = pivot.shape[0], pivot.shape[1]
    if tuple(ref.shape) == (T, 2):
        stride = 0
    elif tuple(ref.shape) == (C, T, 2):
        stride = 2 * T
    else:
        raise ValueError("ref must be [T,2] or [C,T,2]")
    if eps_cell.numel() != C:
        raise ValueError("eps_cell must have C entries")
    if out is None:
        out = minkowski_alloc(C, T, sums.device)
    g = out.get
    _lib.check(lib.ccmpc_minkowski_halfplanes(
        _p(sums), _p(pivot), _p(ref), stride, _p(eps_cell), C, T, int(prediction_horizon), float(R), float(target_p),
        float(tol), int(maxiter), int(accel), _p(g("nd")), _p(g("ab")), _p(g("idx_side")), _p(g("q")),
        _p(g("q1")), _p(g("betas")), _p(g("cov_infer")), _p(g("cov_mu")), _p(g("prob_pair")), _p(g("prob_lower")),
        _p(g("mean")), _p(g("status")), _stream()), "ccmpc_minkowski_halfplanes")
    return out


def affine_halfplanes(sums, pivot, ref, eps_cell, Tpred: int, R: float = R_EV_OV, out: Optional[dict] = None):
    """Affine half-plane loop (v8ideal:1476-1515) per (cell, t)."""
    lib = _lib.load()
    _chk(sums, F64, "sums"); _chk(pivot, F64, "pivot", (2,)); _chk(ref, F64, "ref", (2,)); _chk(eps_cell, F64, "eps")
    C, T = pivot.shape[0], pivot.shape[1]
    stride = 0 if tuple(ref.shape) == (T, 2) else 2 * T
    if out is None:
        dev = sums.device
        out = dict(nd=torch.empty((C, T, 3), dtype=F64, device=dev),
                   offset_rhs=torch.empty((C, T, 2), dtype=F64, device=dev),
                   idx_side=torch.empty((C, T), dtype=I8, device=dev),
                   mean=torch.empty((C, T, 2), dtype=F64, device=dev),
                   cov=torch.empty((C, T, 3), dtype=F64, device=dev),
                   status=torch.empty((C, T), dtype=U8, device=dev))
    _lib.check(lib.ccmpc_affine_halfplanes(_p(sums), _p(pivot), _p(ref), stride, _p(eps_cell), C, T, int(Tpred),
                                           float(R), _p(out["nd"]), _p(out["offset_rhs"]), _p(out["idx_side"]),
                                           _p(out["mean"]), _p(out["cov"]), _p(out["status"]), _stream()),
               "ccmpc_affine_halfplanes")
    return out


def violation_counts(samples, offsets, nd, idx_side, out=None):
    """A9 (ii): #{samples on the ego side of each emitted half-plane} -> [C,P] int64."""
    lib = _lib.load()
    _chk(samples, F64, "samples", (2,)); _chk(offsets, I64, "offsets"); _chk(nd, F64, "nd", (3,))
    _chk(idx_side, I8, "idx_side")
    N, T = samples.shape[0], samples.shape[1]
    C = offsets.numel() - 1
    P = T * (T - 1) // 2
    if out is None:
        out = _new((C, P), I64, samples)
    _lib.check(lib.ccmpc_violation_counts(_p(samples), _p(offsets), N, C, T, _p(nd), _p(idx_side), _p(out),
                                          _stream()), "ccmpc_violation_counts")
    return out


def halfplanes_to_block(nd, rhs, idx_side, t_of_row, T: int, nx: int = 4, nu: int = 2, gamma=None, x_bar=None,
                        u_bar=None, want_state_block: bool = True, out: Optional[dict] = None):
    """Emitted rows -> ONE solver-ready block instead of one cvxpy row each (v8ideal:927-936, :1503-1515).

    nd [..., >=2] (unit rows n0, n1 first), rhs [...] or a strided view column (e.g. nd[..., 2] or
    offset_rhs[..., 1]), idx_side [...] int8, t_of_row [...] int32.
    -> dict(A_x [rows, nx T], b_x [rows]) with ``A_x @ x <= b_x`` in the stacked state of v8ideal:2873-2894,
    plus (A_u [rows, nu T], b_u [rows]) with ``A_u @ u <= b_u`` when the LTV model (gamma, x_bar, u_bar) is given."""
    lib = _lib.load()
    _chk(nd, F64, "nd"); _chk(idx_side, I8, "idx_side"); _chk(t_of_row, I32, "t_of_row")
    nd_stride = int(nd.shape[-1])
    rows = idx_side.numel()
    if nd.numel() != rows * nd_stride or t_of_row.numel() != rows:
        raise ValueError("halfplanes_to_block: nd / idx_side / t_of_row disagree on the number of rows")
    # rhs may be a column view of a [rows, k] tensor: pass base pointer of the column + its stride
    if not (isinstance(rhs, torch.Tensor) and rhs.is_cuda and rhs.dtype == F64 and rhs.numel() == rows):
        raise TypeError("rhs: expected a float64 CUDA tensor with one entry per row")
    flat = rhs.reshape(-1) if rhs.is_contiguous() else rhs
    strides = [st for st, sz in zip(flat.stride(), flat.shape) if sz > 1]
    rhs_stride = 1 if flat.is_contiguous() else (strides[-1] if strides else 1)
    if not flat.is_contiguous():
        # a column of a contiguous [..., k] tensor: consecutive rows are k apart in every leading dimension
        exp, ok = rhs_stride, True
        for st, sz in reversed(list(zip(flat.stride(), flat.shape))):
            ok &= (sz == 1 or st == exp)
            exp *= sz
        if not ok:
            raise ValueError("rhs: only contiguous tensors or a last-axis column of one are supported")
    dev = nd.device
    if out is None:
        out = {}
        if want_state_block:
            out["A_x"] = torch.empty((rows, nx * T), dtype=F64, device=dev)
            out["b_x"] = torch.empty((rows,), dtype=F64, device=dev)
        if gamma is not None:
            out["A_u"] = torch.empty((rows, nu * T), dtype=F64, device=dev)
            out["b_u"] = torch.empty((rows,), dtype=F64, device=dev)
    ldg = 0
    if gamma is not None:
        _chk(gamma, F64, "gamma"); _chk(x_bar, F64, "x_bar"); _chk(u_bar, F64, "u_bar")
        if gamma.dim() != 2 or gamma.shape[0] < nx * T or gamma.shape[1] < nu * T or x_bar.numel() < nx * T \
                or u_bar.numel() < nu * T:
            raise ValueError("gamma must be [>= nx T, >= nu T], x_bar [>= nx T], u_bar [>= nu T]")
        ldg = int(gamma.shape[1])
    _lib.check(lib.ccmpc_halfplanes_to_block(_p(nd), nd_stride, ctypes.c_void_p(rhs.data_ptr()), int(rhs_stride),
                                             _p(idx_side), _p(t_of_row), rows, int(T), int(nx), int(nu), _p(gamma),
                                             ldg, _p(x_bar), _p(u_bar), _p(out.get("A_x")), _p(out.get("b_x")),
                                             _p(out.get("A_u")), _p(out.get("b_u")), _stream()),
               "ccmpc_halfplanes_to_block")
    return out


# ---------------------------------------------------------------------------
# K5  boxes and L4 outer approximation
# ---------------------------------------------------------------------------
def vertices_of_bboxes(centers, theta, lw):
    """utility/npu.py:308-350 on device: centers [N,2], theta [N] or [1], lw (l, w) -> [N,4,2]."""
    lib = _lib.load()
    _chk(centers, F64, "centers", (2,)); _chk(theta, F64, "theta")
    N = centers.shape[0]
    scalar = 1 if theta.numel() == 1 and N != 1 else 0
    if not scalar and theta.numel() != N:
        raise ValueError("theta must have 1 or N entries")
    out = _new((N, 4, 2), F64, centers)
    _lib.check(lib.ccmpc_vertices_of_bboxes(_p(centers), _p(theta), scalar, float(lw[0]), float(lw[1]), N, _p(out),
                                            _stream()), "ccmpc_vertices_of_bboxes")
    return out


def l4_outer(samples, yaws, offsets, yaw_mean, lw=(3.7, 1.79), lw_cell=None, out=None):
    """v8ideal:627-640 + :694-736 + midlevel/util.py:171-200 per (cell, t) -> At [C,T,4,2], b0 [C,T,4]."""
    lib = _lib.load()
    _chk(samples, F64, "samples", (2,)); _chk(yaws, F64, "yaws"); _chk(offsets, I64, "offsets")
    _chk(yaw_mean, F64, "yaw_mean"); _chk(lw_cell, F64, "lw_cell", (2,))
    N, T = samples.shape[0], samples.shape[1]
    C = offsets.numel() - 1
    if out is None:
        out = (_new((C, T, 4, 2), F64, samples), _new((C, T, 4), F64, samples))
    At, b0 = out
    _lib.check(lib.ccmpc_l4_outer(_p(samples), _p(yaws), _p(offsets), N, C, T, _p(yaw_mean), _p(lw_cell),
                                  float(lw[0]), float(lw[1]), _p(At), _p(b0), _stream()), "ccmpc_l4_outer")
    return At, b0


def group_by_mode(predictions, z, latent_pmf, minpos=(0.0, 0.0), filter_pmf: float = 0.1, workspace=None,
                  out: Optional[dict] = None, pos_last=None):
    """A2 on the device (v8ideal:486-494 + ovehicle.py:55-113).  ``predictions`` [V,N,T,2] float32 (minpos is
    added after the upcast) or float64 (minpos already added), ``z`` [V,N] int32, ``latent_pmf`` [V,S] float64.
    -> dict(samples [V*N,T,2] f64 grouped, offsets [V*S+1] i64 (cell = v*S + j), perm [V,N] i64,
            n_states [V] i32, cell_latent [V*S] i32, cell_pmf [V*S], init_center [V*S,2], status [V] u8).
    ``pos_last`` [V,2] (last observed position per vehicle): also returns ``yaws`` [V*N,T] of the grouped rows
    (ovehicle.py:72-76), computed inside the gather instead of by a separate yaws_from_positions pass."""
    lib = _lib.load()
    if predictions.dtype not in (torch.float32, F64):
        raise TypeError("predictions must be float32 or float64")
    _chk(predictions, predictions.dtype, "predictions", (2,))
    _chk(z, torch.int32, "z"); _chk(latent_pmf, F64, "latent_pmf")
    V, N, T = predictions.shape[0], predictions.shape[1], predictions.shape[2]
    S = latent_pmf.shape[1]
    if tuple(z.shape) != (V, N) or latent_pmf.shape[0] != V:
        raise ValueError("z must be [V,N] and latent_pmf [V,S]")
    dev = predictions.device
    if out is None:
        out = dict(samples=torch.empty((V * N, T, 2), dtype=F64, device=dev),
                   offsets=torch.empty(V * S + 1, dtype=I64, device=dev),
                   perm=torch.empty((V, N), dtype=I64, device=dev),
                   n_states=torch.empty(V, dtype=torch.int32, device=dev),
                   cell_latent=torch.empty(V * S, dtype=torch.int32, device=dev),
                   cell_pmf=torch.empty(V * S, dtype=F64, device=dev),
                   init_center=torch.empty((V * S, 2), dtype=F64, device=dev),
                   status=torch.empty(V, dtype=U8, device=dev))
        if pos_last is not None:
            out["yaws"] = torch.empty((V * N, T), dtype=F64, device=dev)
    _chk(pos_last, F64, "pos_last", (2,))
    if pos_last is not None and tuple(pos_last.shape) != (V, 2):
        raise ValueError("pos_last must be [V,2]")
    yaws = out.get("yaws") if pos_last is not None else None
    need = int(lib.ccmpc_group_workspace_bytes(V, N))
    if workspace is None:
        workspace = torch.empty(need, dtype=U8, device=dev)
    f32 = predictions.dtype == torch.float32
    _lib.check(lib.ccmpc_group_by_mode(_p(predictions) if f32 else None, None if f32 else _p(predictions), _p(z),
                                       _p(latent_pmf), V, N, T, S, float(minpos[0]), float(minpos[1]),
                                       float(filter_pmf), _p(out["samples"]), _p(out["offsets"]), _p(out["perm"]),
                                       _p(out["n_states"]), _p(out.get("cell_latent")), _p(out.get("cell_pmf")),
                                       _p(out.get("init_center")), _p(out.get("status")), _p(pos_last), _p(yaws),
                                       _p(workspace),
                                       workspace.numel(), _stream()), "ccmpc_group_by_mode")
    return out


def yaws_from_positions(samples, offsets, pos_last, out=None):
    """midlevel/ovehicle.py:72-76 on device: samples [N,T,2], offsets [C+1], pos_last [C,2] -> yaws [N,T]."""
    lib = _lib.load()
    _chk(samples, F64, "samples", (2,)); _chk(offsets, I64, "offsets"); _chk(pos_last, F64, "pos_last", (2,))
    N, T = samples.shape[0], samples.shape[1]
    C = offsets.numel() - 1
    if out is None:
        out = _new((N, T), F64, samples)
    _lib.check(lib.ccmpc_yaws_from_positions(_p(samples), _p(offsets), N, C, T, _p(pos_last), _p(out), _stream()),
               "ccmpc_yaws_from_positions")
    return out


# ---------------------------------------------------------------------------
# K1  Monte Carlo rollout (predict_ideal, v8ideal:2662-2709)
# ---------------------------------------------------------------------------
def mc_rollout(mean, cov, cross, x0, n: int, Tsh: int, seed: int = 0, sample_offset: int = 0, z=None,
               want_z: bool = False, out=None):
    """mean [C,Tm,2], cov [C,Tm,2,2], cross [C,Tm,Tm,2,2], x0 [C,2] | None (device draw)
    -> traj [C,n,Tsh,2], z_used | None, status [C] uint8, x0_used [C,2]."""
    lib = _lib.load()
    _chk(mean, F64, "mean", (2,)); _chk(cov, F64, "cov", (2, 2)); _chk(cross, F64, "cross", (2, 2))
    _chk(x0, F64, "x0", (2,)); _chk(z, F64, "z", (2,))
    C, Tm = mean.shape[0], mean.shape[1]
    if tuple(cov.shape) != (C, Tm, 2, 2) or tuple(cross.shape) != (C, Tm, Tm, 2, 2) or \
            (x0 is not None and tuple(x0.shape) != (C, 2)):
        raise ValueError("mc_rollout: moment shapes do not agree")
    if z is not None and tuple(z.shape) != (C, n, Tsh, 2):
        raise ValueError("z must be [C,n,Tsh,2]")
    traj = out if out is not None else _new((C, n, Tsh, 2), F64, mean)
    z_out = _new((C, n, Tsh, 2), F64, mean) if want_z else None
    status = _new((C,), U8, mean)
    x0_out = _new((C, 2), F64, mean)
    _lib.check(lib.ccmpc_mc_rollout(_p(mean), _p(cov), _p(cross), Tm, _p(x0), C, int(n), int(Tsh),
                                    int(seed) & 0xFFFFFFFFFFFFFFFF, int(sample_offset), _p(z), _p(traj), _p(z_out),
                                    _p(status), _p(x0_out), _stream()), "ccmpc_mc_rollout")
    return traj, z_out, status, x0_out


def mc_rollout_sums_workspace(C: int, n: int, Tsh: int, device) -> torch.Tensor:
    nbytes = _lib.load().ccmpc_mc_rollout_sums_workspace_bytes(int(C), int(n), int(Tsh))
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)


def mc_rollout_sums(mean, cov, cross, x0, n: int, Tsh: int, pivot, ego, R: float = R_EV_OV, seed: int = 0,
                    sample_offset: int = 0, materialize: bool = True, workspace=None, out=None, sums_out=None):
    """mc_rollout + sample_sums in one kernel (the trajectory tensor is written once and not read back).
    Arguments as mc_rollout plus pivot [C,Tsh,2] and ego [Tsh,2] | [C,Tsh,2]; any Tsh <= 16, Philox noise.
    -> traj [C,n,Tsh,2] | None (materialize=False), sums [C, sums_len(Tsh)], status [C] uint8, x0_used [C,2]."""
    lib = _lib.load()
    _chk(mean, F64, "mean", (2,)); _chk(cov, F64, "cov", (2, 2)); _chk(cross, F64, "cross", (2, 2))
    _chk(x0, F64, "x0", (2,)); _chk(pivot, F64, "pivot", (2,)); _chk(ego, F64, "ego", (2,))
    C, Tm = mean.shape[0], mean.shape[1]
    if tuple(cov.shape) != (C, Tm, 2, 2) or tuple(cross.shape) != (C, Tm, Tm, 2, 2) or \
            (x0 is not None and tuple(x0.shape) != (C, 2)) or tuple(pivot.shape) != (C, Tsh, 2):
        raise ValueError("mc_rollout_sums: shapes do not agree")
    if tuple(ego.shape) == (Tsh, 2):
        stride = 0
    elif tuple(ego.shape) == (C, Tsh, 2):
        stride = 2 * Tsh
    else:
        raise ValueError("ego must be [Tsh,2] or [C,Tsh,2]")
    if workspace is None:
        workspace = mc_rollout_sums_workspace(C, n, Tsh, mean.device)
    traj = (out if out is not None else _new((C, n, Tsh, 2), F64, mean)) if materialize else None
    sums = sums_out if sums_out is not None else _new((C, sums_len(Tsh)), F64, mean)
    status = _new((C,), U8, mean)
    x0_out = _new((C, 2), F64, mean)
    _lib.check(lib.ccmpc_mc_rollout_sums(_p(mean), _p(cov), _p(cross), Tm, _p(x0), C, int(n), int(Tsh),
                                         int(seed) & 0xFFFFFFFFFFFFFFFF, int(sample_offset), _p(pivot), _p(ego),
                                         stride, float(R), _p(traj), _p(sums), _p(status), _p(x0_out), _p(workspace),
                                         workspace.numel(), _stream()), "ccmpc_mc_rollout_sums")
    return traj, sums, status, x0_out


def risk_layout(T: int) -> dict:
    """Offsets into the risk vector of mc_risk: n, coll[T], any, pen[T], d2s[T]."""
    return dict(n=0, coll=1, any=1 + T, pen=2 + T, d2s=2 + 2 * T, len=3 * T + 2)


def mc_risk_workspace(C: int, n: int, Tsh: int, device) -> torch.Tensor:
    nbytes = _lib.load().ccmpc_mc_risk_workspace_bytes(int(C), int(n), int(Tsh))
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)


def mc_risk(mean, cov, cross, x0, n: int, Tsh: int, ego, R: float = R_EV_OV, seed: int = 0, sample_offset: int = 0,
            workspace=None, out=None):
    """Fused predict_ideal rollout (v8ideal:2662-2709) + A9 evaluation, no trajectory tensor.  Arguments as
    mc_rollout plus ego [Tsh,2] | [C,Tsh,2].  -> risk [C, 3 Tsh + 2] (risk_layout), status [C] uint8, x0_used [C,2].
    The samples are the ones mc_rollout would write for the same (seed, sample_offset)."""
    lib = _lib.load()
    _chk(mean, F64, "mean", (2,)); _chk(cov, F64, "cov", (2, 2)); _chk(cross, F64, "cross", (2, 2))
    _chk(x0, F64, "x0", (2,)); _chk(ego, F64, "ego", (2,))
    C, Tm = mean.shape[0], mean.shape[1]
    if tuple(cov.shape) != (C, Tm, 2, 2) or tuple(cross.shape) != (C, Tm, Tm, 2, 2) or \
            (x0 is not None and tuple(x0.shape) != (C, 2)):
        raise ValueError("mc_risk: moment shapes do not agree")
    if tuple(ego.shape) == (Tsh, 2):
        stride = 0
    elif tuple(ego.shape) == (C, Tsh, 2):
        stride = 2 * Tsh
    else:
        raise ValueError("ego must be [Tsh,2] or [C,Tsh,2]")
    if workspace is None:
        workspace = mc_risk_workspace(C, n, Tsh, mean.device)
    risk = out if out is not None else _new((C, 3 * Tsh + 2), F64, mean)
    status = _new((C,), U8, mean)
    x0_out = _new((C, 2), F64, mean)
    _lib.check(lib.ccmpc_mc_risk(_p(mean), _p(cov), _p(cross), Tm, _p(x0), C, int(n), int(Tsh),
                                 int(seed) & 0xFFFFFFFFFFFFFFFF, int(sample_offset), _p(ego), stride, float(R),
                                 _p(risk), _p(status), _p(x0_out), _p(workspace), workspace.numel(), _stream()),
               "ccmpc_mc_risk")
    return risk, status, x0_out


DYNAMICS = {None: 0, "none": 0, "single_integrator": 1, "unicycle": 2}


def gmm2d_sample(latent_pmf, mus, log_sigmas, corrs, n: int, log_pis=None, seed: int = 0, sample_offset: int = 0,
                 noise=None, out=None, seed_dev=None, dynamics=None, x0=None, dt: float = 0.5):
    """Stage 1 on the device (prediction.py:76-104; GMM2D.rsample of the missing Trajectron++ submodule).
    latent_pmf [V,Z] f64, mus / log_sigmas [V,Z,T,G,2] f32, corrs [V,Z,T,G] f32, log_pis [V,Z,T,G] f32 (None iff
    G = 1); noise = (u_z [V,N], u_c [V,N,T], eps [V,N,T,2]) f32 for deterministic parity, else Philox.
    seed_dev: optional int64 [1] device word added to ``seed`` at run time (fresh draws per CUDA-graph replay).
    dynamics: None (the GMM is over positions), "single_integrator" or "unicycle" (the GMM is over controls and
    every sample is integrated from x0 [V,4] f32 = (x, y, phi, v) with step dt: `dynamic.integrate_samples`).
    -> (z [V,N] int32, predictions [V,N,T,2] float32): the inputs of group_by_mode."""
    lib = _lib.load()
    F32 = torch.float32
    _chk(latent_pmf, F64, "latent_pmf"); _chk(mus, F32, "mus", (2,)); _chk(log_sigmas, F32, "log_sigmas", (2,))
    _chk(corrs, F32, "corrs"); _chk(log_pis, F32, "log_pis")
    if mus.dim() != 5:
        raise ValueError("mus must be [V,Z,T,G,2]")
    V, Z, T, G = mus.shape[:4]
    if tuple(latent_pmf.shape) != (V, Z) or tuple(log_sigmas.shape) != (V, Z, T, G, 2) or \
            tuple(corrs.shape) != (V, Z, T, G) or (log_pis is not None and tuple(log_pis.shape) != (V, Z, T, G)):
        raise ValueError("gmm2d_sample: shape mismatch")
    if G > 1 and log_pis is None:
        raise ValueError("gmm2d_sample: G > 1 needs log_pis")
    u_z = u_c = eps = None
    if noise is not None:
        u_z, u_c, eps = noise
        _chk(u_z, F32, "u_z"); _chk(u_c, F32, "u_c"); _chk(eps, F32, "eps", (2,))
        if tuple(u_z.shape) != (V, n) or tuple(u_c.shape) != (V, n, T) or tuple(eps.shape) != (V, n, T, 2):
            raise ValueError("gmm2d_sample: noise must be ([V,N], [V,N,T], [V,N,T,2])")
    if out is None:
        out = (torch.empty((V, n), dtype=torch.int32, device=mus.device),
               torch.empty((V, n, T, 2), dtype=F32, device=mus.device))
    z, pred = out
    if dynamics not in DYNAMICS:
        raise ValueError(f"gmm2d_sample: dynamics must be one of {sorted(k for k in DYNAMICS if k)} or None")
    if DYNAMICS[dynamics]:
        _chk(x0, F32, "x0", (4,))
        if x0 is None or tuple(x0.shape) != (V, 4):
            raise ValueError("gmm2d_sample: dynamics needs x0 [V,4] float32 (x, y, phi, v)")
        _lib.check(lib.ccmpc_gmm2d_sample_dynamics(
            _p(latent_pmf), _p(log_pis), _p(mus), _p(log_sigmas), _p(corrs), DYNAMICS[dynamics], _p(x0), float(dt), V,
            int(n), Z, T, G, int(seed) & (2 ** 64 - 1), _p(seed_dev), int(sample_offset), _p(u_z), _p(u_c), _p(eps),
            _p(z), _p(pred), _stream()), "ccmpc_gmm2d_sample_dynamics")
        return z, pred
    _lib.check(lib.ccmpc_gmm2d_sample(_p(latent_pmf), _p(log_pis), _p(mus), _p(log_sigmas), _p(corrs), V, int(n), Z,
                                      T, G, int(seed) & (2 ** 64 - 1), _p(seed_dev), int(sample_offset), _p(u_z), _p(u_c),
                                      _p(eps),
                                      _p(z), _p(pred), _stream()), "ccmpc_gmm2d_sample")
    return z, pred


def philox4x32(ctr, key):
    """Philox-4x32-10 known-answer hook: ctr [M,4] int32 bits, key [M,2] -> [M,4]."""
    lib = _lib.load()
    _chk(ctr, I32, "ctr", (4,)); _chk(key, I32, "key", (2,))
    M = ctr.shape[0]
    out = torch.empty_like(ctr)
    _lib.check(lib.ccmpc_philox4x32(_p(ctr), _p(key), M, _p(out), _stream()), "ccmpc_philox4x32")
    return out
