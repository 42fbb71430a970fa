"""Planner-level constraint build: the hot methods of the reference's
``MidlevelAgent`` (collect/in_simulation/midlevel/v8ideal/__init__.py, "v8ideal")
on libccmpc.

What is mirrored (same names, argument meaning and quirks):
  compute_obstacle_constraints_GMM_Minkowski_idealprediction   v8ideal:781-964
  compute_obstacle_constraints_GMM_affine                      v8ideal:1378-1539
  predict_ideal                                                v8ideal:2620-2711
  save_moments                                                 v8ideal:2575-2618
  __compute_vertices / __compute_overapproximations            v8ideal:627-640, :694-736
  eps_ura allocation                                           v8ideal:2920-2926
The cvxpy / CPLEX solve, CARLA and plotting stay with the reference planner:
this class hands it half-plane rows (n, d, side) / (A, b), the L4 polytopes and
the bookkeeping tuples it stores in its ``_cov`` pickle.

All arithmetic runs in the sm_100a kernels behind include/ccmpc.h; this file
only packs cells, owns device buffers and converts results to the nested
Python containers the reference planner indexes.  No CPU fallback.
"""
from __future__ import annotations

import os
import pickle
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import ops

EPS_TOTAL = 0.05                   # v8ideal:2920
BBOX_DEFAULT = (3.7, 1.79)         # v8ideal:792
N_IDEAL_SAMPLES = 1_000_000        # v8ideal:2640


class InSimulationException(Exception):
    """Same name as collect/exception.py:11 so planner-side ``except`` clauses keep working."""


def _device(device=None) -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("ccmpc_b200 needs a CUDA device (there is no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


def eps_allocation(K_per_obstacle: Sequence[int], eps: float = EPS_TOTAL) -> np.ndarray:
    """eps_ura[o, k] = eps / O for k < K_o (v8ideal:2920-2926)."""
    O = len(K_per_obstacle)
    out = np.zeros((O, int(max(K_per_obstacle)) if O else 0))
    for o in range(O):
        out[o, :int(K_per_obstacle[o])] = eps / O
    return out


_HOST_STAGE: Dict[int, torch.Tensor] = {}


def _to_host(tensors: Dict[str, torch.Tensor]) -> Dict[str, np.ndarray]:
    """Small device results -> host with ONE stream synchronisation: every tensor is copied asynchronously into its
    slice of a reusable pinned buffer (a `.cpu()` per tensor is a sync each), then the stream is synchronised and the
    slices are copied out (the pinned buffer is reused by the next call)."""
    names = [k for k, v in tensors.items() if v is not None]
    if not names:
        return {}
    sizes = [int(tensors[k].numel() * tensors[k].element_size()) for k in names]
    offs, total = [], 0
    for n in sizes:
        offs.append(total)
        total += (n + 15) // 16 * 16
    cap = 1 << max(12, (total - 1).bit_length())
    stage = _HOST_STAGE.get(cap)
    if stage is None:
        stage = _HOST_STAGE[cap] = torch.empty(cap, dtype=torch.uint8).pin_memory()
    for k, n, o in zip(names, sizes, offs):
        t = tensors[k].contiguous()
        stage[o:o + n].view(t.dtype).view(t.shape).copy_(t, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    host = stage[:total].numpy().copy()
    return {k: host[o:o + n].view(_NP_OF[tensors[k].dtype]).reshape(tuple(tensors[k].shape))
            for k, n, o in zip(names, sizes, offs)}


_NP_OF = {torch.float64: np.float64, torch.float32: np.float32, torch.int64: np.int64, torch.int32: np.int32,
          torch.int8: np.int8, torch.uint8: np.uint8}


def _accel_mode(accel) -> int:
    """0: the reference's iterate sequence from beta = 1 (makeconstraint:24-32; same iterates, same count);
    1: the same fp64 map and stopping rule started from a float32 solve of the stationarity cubic (2-4 iterations);
    2: fp64 Newton on the cubic, accepted by one evaluation of the reference map under its stopping rule."""
    a = int(accel)
    if a not in (0, 1, 2) or (not isinstance(accel, (bool, int, np.integer))):
        raise ValueError(f"accel must be 0, 1 or 2 (got {accel!r})")
    return a


def raise_for_status(status: np.ndarray, populated: Optional[np.ndarray] = None, what: str = "constraint build"):
    """Turn the per-row ``status`` bytes of a builder into the exception the reference raises at the same point.

    ST_SINGULAR            ``np.linalg.inv(cov_tau)`` on a singular matrix (makeconstraint:65; also a cell with
                           fewer than 2 samples, whose np.cov is NaN) -> ``np.linalg.LinAlgError``
    ST_NO_TANGENT          ``choose_closest_tangent`` returned 4 x None and ``n_star @ mean`` fails
                           (makeconstraint:191-192, v8ideal:926) -> ``TypeError``
    ST_DEGENERATE_SLOPE    ``ref_y == mean_y``: the reference divides by zero (v8ideal:923) and hands +-inf / NaN
                           rows to the solver, which rejects them -> ``InSimulationException``
    ST_NOT_CONVERGED       the fixed point did not meet the tolerance in maxiter steps: the reference returns the
                           last iterate silently, and so do we (no exception).
    ``populated`` [C] masks out cells that hold no samples by construction (fixed-layout ingest plans)."""
    st = np.asarray(status)
    if populated is not None:
        st = st[np.asarray(populated, dtype=bool)]
    if not st.size:
        return
    for code, exc, msg in ((ops.ST_SINGULAR, np.linalg.LinAlgError, "Singular matrix"),
                           (ops.ST_NO_TANGENT, TypeError,
                            "unsupported operand type(s) for @: 'NoneType' and 'numpy.ndarray'"),
                           (ops.ST_DEGENERATE_SLOPE, InSimulationException,
                            "reference point and obstacle mean share the y coordinate: tangent slope is infinite")):
        hit = np.argwhere(st == code)
        if hit.size:
            raise exc(f"{msg} ({what}: {hit.shape[0]} row(s), first at index {tuple(int(i) for i in hit[0])})")


_PINNED_STAGE: Dict[str, list] = {}       # device -> [pinned uint8 buffer, event of the last copy out of it]


def _pinned_stage(device, nbytes: int) -> torch.Tensor:
    """``nbytes`` of a grow-only pinned host buffer for one packed H2D copy per step.  The previous copy out of the
    buffer has finished before it is handed out again (event wait); one buffer per device, not thread-safe."""
    ent = _PINNED_STAGE.setdefault(str(device), [None, None])
    if ent[1] is not None:
        ent[1].synchronize()
    if ent[0] is None or ent[0].numel() < nbytes:
        ent[0] = torch.empty(max(int(nbytes) * 5 // 4, 1 << 20), dtype=torch.uint8).pin_memory()
    return ent[0][:max(int(nbytes), 8)]


def _pinned_stage_done(device) -> None:
    """Record that a copy out of the pinned stage was just enqueued on the current stream."""
    ev = torch.cuda.Event()
    ev.record()
    _PINNED_STAGE[str(device)][1] = ev


def _h2d(arr: np.ndarray, device, dtype=None) -> torch.Tensor:
    """Host ndarray -> device on the current stream.  Small arrays go from pageable memory (the driver stages them;
    pinning a fresh buffer per call costs more than the copy), large ones through a pinned buffer."""
    t = torch.from_numpy(np.ascontiguousarray(arr))
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    if t.numel() * t.element_size() >= (16 << 20):
        t = t.pin_memory()
    return t.to(device, non_blocking=True)


@dataclass
class CellBatch:
    """Device-resident samples of all (obstacle, mode) cells of one MPC step.

    ``samples`` is the row-major concatenation of the reference's
    ``ovehicle.pred_positions[latent_idx]`` arrays ((N_k, T, 2) each), obstacle-major,
    mode-minor - the layout of ``np.vstack(...)`` at v8ideal:890 - with CSR ``offsets``."""
    samples: torch.Tensor                  # [N, T, 2] float64
    yaws: Optional[torch.Tensor]           # [N, T] float64
    offsets: torch.Tensor                  # [C+1] int64 (device)
    offsets_host: np.ndarray               # [C+1] int64
    cell_obstacle: np.ndarray              # [C] int
    cell_mode: np.ndarray                  # [C] int
    n_states: List[int]                    # K_o per obstacle
    lw_cell: Optional[torch.Tensor] = None  # [C, 2] float64 (ovehicle.bbox per cell)
    bboxes: Optional[np.ndarray] = None    # [O, 2] host

    @property
    def C(self) -> int:
        return int(self.offsets_host.size - 1)

    @property
    def T(self) -> int:
        return int(self.samples.shape[1])

    @property
    def O(self) -> int:
        return len(self.n_states)

    @property
    def N(self) -> int:
        return int(self.samples.shape[0])

    def cell_index(self, o: int, k: int) -> int:
        hit = np.nonzero((self.cell_obstacle == o) & (self.cell_mode == k))[0]
        if hit.size == 0:
            raise KeyError((o, k))
        return int(hit[0])

    @classmethod
    def from_ovehicles(cls, ovehicles, device=None, with_yaws: bool = True) -> "CellBatch":
        """Pack reference ``OVehicle`` objects (attributes ``pred_positions``, ``pred_yaws``,
        ``n_states``, ``bbox``; midlevel/ovehicle.py:119-131) into one device batch."""
        device = _device(device)
        pos, yaw, offs, co, ck, ns, bb = [], [], [0], [], [], [], []
        T = None
        for o, ov in enumerate(ovehicles):
            ns.append(int(ov.n_states))
            bb.append(np.asarray(getattr(ov, "bbox", BBOX_DEFAULT), dtype=np.float64))
            pp, yy = ov.pred_positions, (ov.pred_yaws if with_yaws else None)
            for k in range(ns[-1]):
                p = pp[k]
                if type(p) is not np.ndarray or p.dtype != np.float64:
                    p = np.asarray(p, dtype=np.float64)
                shp = p.shape
                if len(shp) != 3 or shp[2] != 2:
                    raise ValueError("pred_positions[k] must be (N_k, T, 2)")
                if T is None:
                    T = shp[1]
                elif shp[1] != T:
                    raise ValueError("all cells must share the horizon T")
                pos.append(p)
                if with_yaws:
                    y = yy[k]
                    if type(y) is not np.ndarray or y.dtype != np.float64:
                        y = np.asarray(y, dtype=np.float64)
                    if y.shape != shp[:2]:
                        raise ValueError("pred_yaws[k] must be (N_k, T)")
                    yaw.append(y)
                offs.append(offs[-1] + shp[0])
                co.append(o)
                ck.append(k)
        if not pos:
            raise ValueError("no cells")
        offsets_host = np.asarray(offs, dtype=np.int64)
        bboxes = np.stack(bb)
        # the per-mode arrays are concatenated straight into a reusable PINNED buffer (positions, then yaws) and go to
        # the device in one asynchronous copy: no temporary, no pageable staging by the driver (2.6 + 1.3 MB at c2)
        N = int(offs[-1])
        n_pos, n_yaw = N * T * 2, (N * T if with_yaws else 0)
        stage = _pinned_stage(device, (n_pos + n_yaw) * 8).view(torch.float64)
        host = stage.numpy()
        np.concatenate(pos, axis=0, out=host[:n_pos].reshape(N, T, 2))
        if with_yaws:
            np.concatenate(yaw, axis=0, out=host[n_pos:n_pos + n_yaw].reshape(N, T))
        dev_buf = torch.empty(n_pos + n_yaw, dtype=torch.float64, device=device)
        dev_buf.copy_(stage[:n_pos + n_yaw], non_blocking=True)
        _pinned_stage_done(device)
        return cls(samples=dev_buf[:n_pos].view(N, T, 2),
                   yaws=dev_buf[n_pos:].view(N, T) if with_yaws else None,
                   offsets=_h2d(offsets_host, device), offsets_host=offsets_host,
                   cell_obstacle=np.asarray(co), cell_mode=np.asarray(ck), n_states=ns,
                   lw_cell=_h2d(bboxes[np.asarray(co)], device), bboxes=bboxes)

    @classmethod
    def uniform(cls, traj: torch.Tensor, cell_obstacle, cell_mode, n_states) -> "CellBatch":
        """Cells of equal size from a [C, n, T, 2] device tensor (e.g. a rollout)."""
        C, n, T, _ = traj.shape
        offsets_host = np.arange(C + 1, dtype=np.int64) * n
        return cls(samples=traj.reshape(C * n, T, 2), yaws=None,
                   offsets=torch.arange(C + 1, dtype=torch.int64, device=traj.device) * n,
                   offsets_host=offsets_host, cell_obstacle=np.asarray(cell_obstacle),
                   cell_mode=np.asarray(cell_mode), n_states=list(n_states))


@dataclass
class Moments:
    """Sample moments of one step in the save_moments layout, device resident."""
    mean: torch.Tensor       # [C, T, 2]
    cov: torch.Tensor        # [C, T, 2, 2]
    cross: torch.Tensor      # [C, T, T, 2, 2]  (tau < t filled)
    cell_obstacle: np.ndarray
    cell_mode: np.ndarray
    n_states: List[int]
    T: int

    def nested(self) -> dict:
        """The dict the reference pickles (v8ideal:2580-2610): nested lists sized
        (O, max K, T[, T-1]) with ``None`` in unused slots."""
        O, Kmax, T = len(self.n_states), max(self.n_states), self.T
        mean = self.mean.cpu().numpy()
        cov = self.cov.cpu().numpy()
        cross = self.cross.cpu().numpy()
        mean_l = np.empty((O, Kmax, T), dtype=object).tolist()
        cov_l = np.empty((O, Kmax, T), dtype=object).tolist()
        cross_l = np.empty((O, Kmax, T, max(T - 1, 0)), dtype=object).tolist()
        for c, (o, k) in enumerate(zip(self.cell_obstacle, self.cell_mode)):
            for t in range(T):
                mean_l[o][k][t] = mean[c, t].copy()
                cov_l[o][k][t] = cov[c, t].copy()
                for tau in range(t):
                    cross_l[o][k][t][tau] = cross[c, t, tau].copy()
        return {"mean_p0p1": mean_l, "cov_p0p1": cov_l, "cross_cov": cross_l}

    @classmethod
    def from_nested(cls, d: dict, device=None) -> "Moments":
        """Inverse of :meth:`nested` (reads a reference ``_moments`` pickle payload)."""
        device = _device(device)
        mean_l, cov_l, cross_l = d["mean_p0p1"], d["cov_p0p1"], d["cross_cov"]
        O = len(mean_l)
        co, ck, ns = [], [], []
        T = None
        for o in range(O):
            kk = [k for k in range(len(mean_l[o])) if mean_l[o][k] is not None and mean_l[o][k][0] is not None]
            ns.append(len(kk))
            for k in kk:
                co.append(o)
                ck.append(k)
                T = len(mean_l[o][k])
        C = len(co)
        mean = np.zeros((C, T, 2))
        cov = np.zeros((C, T, 2, 2))
        cross = np.zeros((C, T, T, 2, 2))
        for c, (o, k) in enumerate(zip(co, ck)):
            for t in range(T):
                mean[c, t] = mean_l[o][k][t]
                cov[c, t] = cov_l[o][k][t]
                for tau in range(t):
                    cross[c, t, tau] = cross_l[o][k][t][tau]
        return cls(_h2d(mean, device), _h2d(cov, device), _h2d(cross, device), np.asarray(co), np.asarray(ck), ns, T)


@dataclass
class HalfPlanes:
    """Rows of one builder call.  Minkowski: one row per (cell, pair p = t(t-1)/2 + tau);
    affine: one row per (cell, t).  ``side`` = +1 for ``n.x >= rhs`` and -1 for ``n.x <= rhs``."""
    n: np.ndarray            # [..., 2]
    d: np.ndarray            # [...]
    rhs: np.ndarray          # [...]  (= d for Minkowski; d +- offset for affine)
    idx: np.ndarray          # [...] best_idx of choose_closest_tangent
    side: np.ndarray         # [...]
    t: np.ndarray            # [...] timestep each row constrains
    status: np.ndarray       # [...] CCMPC_ST_*
    cell_obstacle: np.ndarray
    cell_mode: np.ndarray
    extra: Dict[str, np.ndarray] = field(default_factory=dict)
    dev: Optional[dict] = None   # device tensors of the same rows (nd, rhs, idx_side, t) for :meth:`block`
    T: int = 0                   # horizon of the rows (columns of the stacked state = nx * T)

    def block(self, nx: int = 4, nu: int = 2, Gamma=None, x_bar=None, u_bar=None, T: Optional[int] = None,
              to_host: bool = True):
        """All rows as ONE inequality block in the planner's stacked state x (X = x.reshape(T, nx), v8ideal:2894):
        ``A_x @ x <= b_x`` - what ``cvxpy_constraints`` hands over as one Python expression per row
        (v8ideal:927-936).  With the LTV model ``Gamma, x_bar, u_bar`` of v8ideal:2873-2889 the block is also
        given in the decision variable: ``A_u @ u <= b_u``.  Built on the device (ccmpc_halfplanes_to_block)."""
        if self.dev is None:
            raise RuntimeError("these rows were not produced by a builder (no device handles)")
        raise_for_status(self.status, what="block")
        T = int(self.T if T is None else T)
        d = self.dev
        g = xb = ub = None
        if Gamma is not None:
            dev = d["nd"].device
            g, xb, ub = (_h2d(np.asarray(a, dtype=np.float64), dev) for a in (Gamma, x_bar, u_bar))
        out = ops.halfplanes_to_block(d["nd"], d["rhs"], d["idx_side"], d["t"], T, nx=nx, nu=nu, gamma=g, x_bar=xb,
                                      u_bar=ub)
        return {k: v.cpu().numpy() for k, v in out.items()} if to_host else out

    def A_b(self):
        """(A, b) with A x <= b, the form v8ideal:933-939 saves."""
        A = -self.side[..., None] * self.n
        b = -self.side * self.rhs
        return A, b

    def cvxpy_constraints(self, temp_x):
        """Rows as cvxpy constraints on the planner's ``temp_x`` (v8ideal:927-936 / :1503-1515)."""
        import cvxpy as cp  # planner-side dependency; not needed by the kernels
        raise_for_status(self.status, what="cvxpy_constraints")
        rows = []
        flat_n = self.n.reshape(-1, 2)
        for r in range(flat_n.shape[0]):
            t = int(self.t.reshape(-1)[r])
            lhs = flat_n[r].T @ cp.vstack([temp_x[t][0], temp_x[t][1]])
            rhs = float(self.rhs.reshape(-1)[r])
            rows.append(lhs >= rhs if self.side.reshape(-1)[r] > 0 else lhs <= rhs)
        return rows


def _pair_t(T: int) -> np.ndarray:
    return np.asarray([t for t in range(T) for _ in range(t)], dtype=np.int32)


class MidlevelConstraints:
    """Owner of the device buffers and the per-step moment cache of one planner."""

    def __init__(self, prediction_horizon: int, ego_vehicle_id: int = 0, out_dir: Optional[str] = "out/data",
                 device=None, R: float = ops.R_EV_OV, target_p: float = ops.TARGET_P,
                 n_ideal_samples: int = N_IDEAL_SAMPLES, seed: int = 0, accel: int = 0,
                 frame_stride: int = 10, road_boundary_constraints: bool = False):
        if not 1 <= prediction_horizon <= ops.MAX_T:
            raise ValueError(f"prediction_horizon must be in 1..{ops.MAX_T}")
        self.accel = _accel_mode(accel)
        self.prediction_horizon = int(prediction_horizon)
        self.ego_vehicle_id = ego_vehicle_id
        self.out_dir = out_dir
        self.device = _device(device)
        self.R = float(R)
        self.target_p = float(target_p)
        self.n_ideal_samples = int(n_ideal_samples)
        self.seed = int(seed)
        self.frame_stride = int(frame_stride)          # predict_ideal reads frame - 10 (v8ideal:2626)
        # the reference's flag (v8ideal:797, :1395): with it set every row carries the road-boundary slack
        # S_big_repeated[0, t], a cvxpy variable that lives in the solver - the builders refuse instead of dropping it
        self.road_boundary_constraints = bool(road_boundary_constraints)
        self._moments: Dict[int, Moments] = {}
        self._calls = 0
        self.prob_lower_save = None
        self.last_ideal: Optional[CellBatch] = None

    # ------------------------------------------------------------------ sums
    def _sums(self, cells: CellBatch, ego=None, pivot=None):
        if pivot is None:
            pivot = ops.cell_pivot(cells.samples, cells.offsets)
        ego_t = None if ego is None else _h2d(np.asarray(ego, dtype=np.float64), self.device)
        sums = ops.sample_sums(cells.samples, cells.offsets, pivot, ego=ego_t, yaws=cells.yaws, R=self.R)
        return sums, pivot

    def _eps_cells(self, cells: CellBatch, eps_ura) -> torch.Tensor:
        e = np.asarray(eps_ura, dtype=np.float64)
        return _h2d(e[cells.cell_obstacle, cells.cell_mode], self.device)

    # ---------------------------------------------------------- save_moments
    def moments_of(self, cells: CellBatch, sums=None, pivot=None) -> Moments:
        if sums is None:
            sums, pivot = self._sums(cells)
        m = ops.moments_from_sums(sums, pivot, want_counts=False, want_yaw=False)
        return Moments(m["mean"], m["cov"], m["cross"], cells.cell_obstacle, cells.cell_mode, cells.n_states, cells.T)

    def save_moments(self, ovehicles, O, K, T, Tpred, ego_vehicle_id, params, _precomputed: Moments = None):
        """v8ideal:2575-2618.  Keeps the moments on the device for the next step's
        :meth:`predict_ideal` and (when ``out_dir`` is set) also writes the reference's
        ``agent{id}_frame{f}_moments`` pickle.  Returns the :class:`Moments`."""
        if _precomputed is None:
            cells = ovehicles if isinstance(ovehicles, CellBatch) else CellBatch.from_ovehicles(
                ovehicles, self.device, with_yaws=False)
            if cells.T != T:
                # the reference strides the vstack by T (:2588); only T' == T is meaningful
                raise ValueError(f"save_moments: cells have horizon {cells.T}, T = {T}")
            _precomputed = self.moments_of(cells)
        frame = int(getattr(params, "frame", 0))
        self._moments[frame] = _precomputed
        for old in [f for f in self._moments if f < frame - 4 * self.frame_stride]:
            del self._moments[old]
        if self.out_dir:
            path = os.path.join(self.out_dir, f"agent{ego_vehicle_id}_frame{frame}_moments")
            try:
                os.makedirs(self.out_dir, exist_ok=True)
                with open(path, "wb") as f:
                    pickle.dump(_precomputed.nested(), f)
            except OSError:
                pass    # the reference logs and continues (:2615-2618)
        return _precomputed

    def _load_moments(self, frame: int, ego_vehicle_id) -> Moments:
        if frame in self._moments:
            return self._moments[frame]
        if self.out_dir:
            path = os.path.join(self.out_dir, f"agent{ego_vehicle_id}_frame{frame}_moments")
            if os.path.isfile(path):
                with open(path, "rb") as f:
                    m = Moments.from_nested(pickle.load(f), self.device)
                self._moments[frame] = m
                return m
        raise InSimulationException(f"no moments of frame {frame} (predict_ideal needs the step {self.frame_stride} "
                                    f"frames earlier, v8ideal:2626)")

    # ------------------------------------------------------------ _cov side file
    def cov_payload(self, built, shrinking: bool, params=None, agent_params=None, apply_robust: bool = True) -> dict:
        """The ``data_save`` dict the planner pickles next to every solve (v8ideal:2979-3003), filled from the
        9-tuple a builder of this class returned.  The caller adds the solver's own entries
        (``solve_time``, ``process_time``, ``x_init``, ``cost``, ``goal``, ``U_star``, ``X_star``, ``timeout``,
        ``infeasible``; v8ideal:3033-3058, :3102-3103) before :meth:`save_data`."""
        _, _, _, _, OVconstraint, direct, ov_mean, ov_cov, meanNtangent = built
        d = {"direct": direct, "MeanCov": True, "OVconstraint": OVconstraint, "ovStateMean_tau_1": ov_mean,
             "ovStateCov_tau_1": ov_cov, "shrinking": shrinking}
        if shrinking and apply_robust:
            d["meanNtangent"] = meanNtangent                      # :2989-2991 (the Minkowski builder returns 0)
        d["params"] = params
        d["__params"] = agent_params
        return d

    def save_data(self, data_save: dict, params, ego_vehicle_id):
        """v8ideal:2559-2567: ``out/data/agent{id}_frame{frame}_cov`` (read back by :meth:`load_data` ten frames
        later and by ``dataForCost_ref.getData_singleEpisode`` / ``ccmpc_b200.cost``)."""
        if not self.out_dir:
            return None
        path = os.path.join(self.out_dir, f"agent{ego_vehicle_id}_frame{int(params.frame)}_cov")
        try:
            os.makedirs(self.out_dir, exist_ok=True)
            with open(path, "wb") as f:
                pickle.dump(data_save, f)
        except OSError:
            return None                                           # the reference logs and continues (:2566-2567)
        return path

    def load_data(self, params, ego_vehicle_id):
        """v8ideal:2547-2557: ``meanNtangent`` of the step ten frames earlier."""
        path = os.path.join(self.out_dir or "", f"agent{ego_vehicle_id}_frame{int(params.frame) - self.frame_stride}_cov")
        with open(path, "rb") as f:
            return pickle.load(f)["meanNtangent"]

    # --------------------------------------------------------- predict_ideal
    def predict_ideal_cells(self, n_states: Sequence[int], T: int, moments: Moments, x0=None, z=None,
                            n_samples: Optional[int] = None, seed: Optional[int] = None,
                            sample_offset: int = 0) -> CellBatch:
        """Device form of predict_ideal: one [C, n, T, 2] rollout for all (o, k)."""
        n = self.n_ideal_samples if n_samples is None else int(n_samples)
        if moments.T < T + 1:
            raise InSimulationException(f"moments cover {moments.T} steps; predict_ideal indexes [t+1] up to {T}")
        sel, co, ck = self._select_cells(n_states, moments)
        mean, cov, cross = self._gather(moments, sel)
        x0_t = None if x0 is None else _h2d(np.asarray(x0, dtype=np.float64).reshape(len(sel), 2), self.device)
        if seed is None:
            self._calls += 1
            seed = (self.seed << 20) + self._calls
        traj, _, status, x0_used = ops.mc_rollout(mean, cov, cross, x0_t, n, T, seed=seed, sample_offset=sample_offset,
                                                  z=z)
        cells = CellBatch.uniform(traj, co, ck, n_states)
        cells.status = status
        cells.x0 = x0_used
        return cells

    def predict_ideal(self, ovehicles, T, ego_vehicle_id, params, as_numpy: bool = False, **kw):
        """v8ideal:2620-2711.  ``{ov_idx: {latent_idx: (n_samples, T, 2)}}`` - device tensors
        (views of one buffer) unless ``as_numpy``.  A non-PD conditional covariance raises
        ``np.linalg.LinAlgError`` like ``np.linalg.cholesky`` (:2696)."""
        frame = int(getattr(params, "frame", 0)) - self.frame_stride
        moments = self._load_moments(frame, ego_vehicle_id)
        n_states = ovehicles.n_states if isinstance(ovehicles, CellBatch) else [int(ov.n_states) for ov in ovehicles]
        cells = self.predict_ideal_cells(n_states, T, moments, **kw)
        st = cells.status.cpu().numpy()
        if st.any():
            bad = int(np.nonzero(st)[0][0])
            raise np.linalg.LinAlgError("Singular matrix" if st[bad] == ops.ST_SINGULAR
                                        else "Matrix is not positive definite")
        self.last_ideal = cells
        n = cells.N // cells.C
        traj = cells.samples.reshape(cells.C, n, T, 2)
        out: Dict[int, Dict[int, object]] = {}
        for c, (o, k) in enumerate(zip(cells.cell_obstacle, cells.cell_mode)):
            out.setdefault(int(o), {})[int(k)] = traj[c].cpu().numpy() if as_numpy else traj[c]
        return out

    @staticmethod
    def _raise_for_rollout_status(st: np.ndarray) -> None:
        """What np.linalg.inv / cholesky raise inside predict_ideal (v8ideal:2683, :2699) for a degenerate cell."""
        if st.any():
            bad = int(np.nonzero(st)[0][0])
            raise np.linalg.LinAlgError("Singular matrix" if st[bad] == ops.ST_SINGULAR
                                        else "Matrix is not positive definite")

    def _predict_ideal_fused(self, cells: CellBatch, T: int, params, ego_xy, x0=None, n_samples=None, seed=None,
                             check: bool = True):
        """predict_ideal (v8ideal:2620-2711) + the sums the builder needs, in one kernel (ccmpc_mc_rollout_sums).
        -> (CellBatch of the ideal trajectories, packed sums, pivot).  Raises like :meth:`predict_ideal`;
        ``check=False`` leaves the per-cell status on the device (``work.status``) for a caller that reads it
        with its other results (one synchronisation per step instead of one in the middle of it)."""
        frame = int(getattr(params, "frame", 0)) - self.frame_stride
        moments = self._load_moments(frame, self.ego_vehicle_id)
        if moments.T < T + 1:
            raise InSimulationException(f"moments cover {moments.T} steps; predict_ideal indexes [t+1] up to {T}")
        n = self.n_ideal_samples if n_samples is None else int(n_samples)
        sel, co, ck = self._select_cells(cells.n_states, moments)
        mean, cov, cross = self._gather(moments, sel)
        pivot = mean[:, 1:T + 1].contiguous()              # x_{t+1} ~ mean[t+1]: a pivot every rank could compute
        if seed is None:
            self._calls += 1
            seed = (self.seed << 20) + self._calls
        ego_t = ego_xy if torch.is_tensor(ego_xy) else _h2d(np.asarray(ego_xy, dtype=np.float64)[:T, :2], self.device)
        x0_t = None if x0 is None else _h2d(np.asarray(x0, dtype=np.float64).reshape(len(sel), 2), self.device)
        traj, sums, status, x0_used = ops.mc_rollout_sums(mean, cov, cross, x0_t, n, T, pivot, ego_t, R=self.R,
                                                          seed=seed, workspace=self._fused_ws(len(sel), n, T))
        if check:
            self._raise_for_rollout_status(status.cpu().numpy())
        work = CellBatch.uniform(traj, co, ck, cells.n_states)
        work.status, work.x0 = status, x0_used
        self.last_ideal = work
        return work, sums, pivot

    # --------------------------------------------------------------- builders
    def minkowski(self, cells: CellBatch, ref_traj, eps_ura, ego=None, want_violations: bool = False,
                  fields=ops.MINK_FIELDS, sums=None, pivot=None):
        """Device core of v8ideal:881-947 on one CellBatch -> (out dict of device tensors, sums, pivot).
        ``sums`` / ``pivot``: the packed sums of the cells when something already produced them (the fused
        rollout)."""
        T = cells.T
        ref = _h2d(np.asarray(ref_traj, dtype=np.float64)[:T, :2], self.device)
        if sums is None:
            sums, pivot = self._sums(cells, ego=ego)
        out = ops.minkowski_alloc(cells.C, T, self.device, fields=fields)
        ops.minkowski_halfplanes(sums, pivot, ref, self._eps_cells(cells, eps_ura), self.prediction_horizon,
                                 R=self.R, target_p=self.target_p, accel=self.accel, out=out)
        if want_violations:
            out["viol"] = ops.violation_counts(cells.samples, cells.offsets, out["nd"], out["idx_side"])
        return out, sums, pivot

    def _l4(self, cells: CellBatch, sums, pivot):
        mom = ops.moments_from_sums(sums, pivot, want_cross=False, want_counts=True, want_yaw=True)
        At, b0 = ops.l4_outer(cells.samples, cells.yaws, cells.offsets, mom["yaw_mean"], lw_cell=cells.lw_cell)
        return At, b0, mom

    @staticmethod
    def _nested_tko(cells: CellBatch, arr: np.ndarray, T: int):
        """[C, T, ...] -> nested list [t][k][o] sized (T, max K, O) (v8ideal:631, :714-716)."""
        O, Kmax = cells.O, max(cells.n_states)
        if cells.C == O * Kmax and min(cells.n_states) == Kmax:
            # every obstacle has Kmax modes, cells obstacle-major: [C, T, ...] -> [T, K, O, ...] views, no Python
            # loop over the cells (100 x 8 assignments at c2)
            tko = np.moveaxis(arr.reshape((O, Kmax) + arr.shape[1:]), (0, 1, 2), (2, 1, 0))[:T]
            flat = list(np.ascontiguousarray(tko).reshape((T * Kmax * O,) + arr.shape[2:]))     # one pass: all views
            return [[flat[(t * Kmax + k) * O:(t * Kmax + k + 1) * O] for k in range(Kmax)] for t in range(T)]
        out = np.empty((T, Kmax, O), dtype=object).tolist()
        for c, (o, k) in enumerate(zip(cells.cell_obstacle, cells.cell_mode)):
            for t in range(T):
                out[t][k][o] = arr[c, t]
        return out

    @staticmethod
    def _nested_ok(cells: CellBatch, vec: np.ndarray):
        O, Kmax = cells.O, max(cells.n_states)
        if cells.C == O * Kmax and min(cells.n_states) == Kmax:      # full, obstacle-major: rows of the reshaped vector
            return [list(row) for row in np.asarray(vec).reshape(O, Kmax)]
        out = np.empty((O, Kmax), dtype=object).tolist()
        for c, (o, k) in enumerate(zip(cells.cell_obstacle, cells.cell_mode)):
            out[o][k] = vec[c]
        return out

    def _ov_state(self, cells: CellBatch, mom: dict):
        """ovStateMean_tau_1 / ovStateCov_tau_1 tuples (v8ideal:871-876); ``mom``: host arrays (_mean, ...)."""
        mean, cov, ym, yv = mom["_mean"], mom["_cov"], mom["_yaw_mean"], mom["_yaw_var"]
        f = lambda v: self._nested_ok(cells, v)
        ov_mean = (f(mean[:, 0, 0]), f(mean[:, 0, 1]), f(ym[:, 0]))
        ov_cov = (f(cov[:, 0, 0, 0]), f(cov[:, 0, 1, 1]), f(yv[:, 0]))
        # OV_injuction (v8ideal:834-844): per obstacle, decided by its LAST latent
        inj = [None] * cells.O
        for c, o in enumerate(cells.cell_obstacle):
            inj[o] = not (mean[c, 0, 0] >= 190 or mean[c, 0, 1] <= -80)
        return ov_mean, ov_cov, inj

    def _vertices(self, cells: CellBatch):
        """Lazy ``vertices[t][k][o]`` (N_k, 4, 2) device tensors (v8ideal:627-640)."""
        T = cells.T

        def make():
            out = np.empty((T, max(cells.n_states), cells.O), dtype=object).tolist()
            for c, (o, k) in enumerate(zip(cells.cell_obstacle, cells.cell_mode)):
                b, e = cells.offsets_host[c], cells.offsets_host[c + 1]
                for t in range(T):
                    out[t][k][o] = ops.vertices_of_bboxes(cells.samples[b:e, t].contiguous(),
                                                          cells.yaws[b:e, t].contiguous(), cells.bboxes[o])
            return out
        return make

    def compute_obstacle_constraints_GMM_Minkowski_idealprediction(
            self, params, ovehicles, Delta2=None, Omicron=None, temp_x=None, eps_ura=None, segments=None,
            Tsh=None, ref_traj=None, ego=None, want_violations=False, ideal_noise: Optional[dict] = None):
        """v8ideal:781-964.  Returns the reference's 9-tuple
        ``(constraints, vertices, A_union, b_union, OVconstraint, direct, ovStateMean_tau_1,
        ovStateCov_tau_1, 0)`` where ``constraints`` is a :class:`HalfPlanes` (or the list of
        cvxpy rows when ``temp_x`` is given) and ``vertices`` is a zero-argument callable that
        materialises the per-sample box corners on demand (only plots read them).

        ``ideal_noise`` = dict(x0 [C,2], z [C,n,Tsh,2] device tensor) replaces the Philox draws of the
        predict_ideal branch (deterministic parity mode).  A degenerate cell raises what the reference
        raises at the same point (:func:`raise_for_status`)."""
        cells = ovehicles if isinstance(ovehicles, CellBatch) else CellBatch.from_ovehicles(ovehicles, self.device)
        Tpred = self.prediction_horizon
        T = int(Tsh if Tsh is not None else Tpred)
        if eps_ura is None:
            eps_ura = eps_allocation(cells.n_states)
        if cells.T != Tpred:
            raise ValueError(f"cells carry horizon {cells.T}; the planner's prediction_horizon is {Tpred}")
        if self.road_boundary_constraints:
            raise NotImplementedError("road-boundary slack S_big (v8ideal:797-818) is not part of this path")
        # sums of the raw predictions: L4 polytopes, [tau+1|tau] bookkeeping, and the rows when T == Tpred
        if T < Tpred:
            if (ideal_noise or {}).get("z") is None:
                # Philox noise: ONE kernel rolls the ideal trajectories out, stores them and reduces them (:824-825
                # + the moment reads of :893-904 and :960)
                work, sums_w, pivot_w = self._predict_ideal_fused(cells, T, params, ref_traj if ego is None else ego,
                                                                 check=False, **(ideal_noise or {}))
                out, _, _ = self.minkowski(work, ref_traj, eps_ura, want_violations=want_violations, sums=sums_w,
                                           pivot=pivot_w)
            else:
                self.predict_ideal(cells, T, self.ego_vehicle_id, params, **ideal_noise)       # :824-825
                work = self.last_ideal
                out, sums_w, pivot_w = self.minkowski(work, ref_traj, eps_ura, ego=ego,
                                                      want_violations=want_violations)
            sums_c, pivot_c = self._sums(cells)
        else:
            work = cells
            out, sums_w, pivot_w = self.minkowski(cells, ref_traj, eps_ura, ego=ego, want_violations=want_violations)
            sums_c, pivot_c = sums_w, pivot_w
        At, b0, mom = self._l4(cells, sums_c, pivot_c)                               # :951-954
        moments = self.moments_of(work, sums_w, pivot_w)
        # ---- one synchronising copy of the small results
        allh = _to_host({**out, "_At": At, "_b0": b0, "_mean": mom["mean"], "_cov": mom["cov"],
                         "_yaw_mean": mom["yaw_mean"], "_yaw_var": mom["yaw_var"],
                         "_ideal_status": getattr(work, "status", None) if work is not cells else None})
        if "_ideal_status" in allh:
            self._raise_for_rollout_status(allh["_ideal_status"])                    # predict_ideal's own exceptions first
        host = {k: v for k, v in allh.items() if not k.startswith("_")}
        raise_for_status(host["status"], what="Minkowski builder")                   # :907, :924-926
        A_union = self._nested_tko(cells, allh["_At"], Tpred)
        b_union = self._nested_tko(cells, allh["_b0"], Tpred)
        ov_mean, ov_cov, inj = self._ov_state(cells, allh)
        OVconstraint = any(bool(v) for v in inj)                                     # :849-851
        direct = [None] * cells.O
        side = np.where(host["idx_side"] & 2, -1, 1).astype(np.int32)
        C, P = side.shape
        t_rows = np.broadcast_to(_pair_t(T), (C, P)).copy()
        hp = HalfPlanes(n=host["nd"][..., :2], d=host["nd"][..., 2], rhs=host["nd"][..., 2],
                        idx=(host["idx_side"] & 1).astype(np.int32), side=side,
                        t=t_rows, status=host["status"],
                        cell_obstacle=work.cell_obstacle, cell_mode=work.cell_mode,
                        extra={k: host[k] for k in host if k not in ("nd", "idx_side", "status")},
                        dev=dict(nd=out["nd"], rhs=out["nd"][..., 2], idx_side=out["idx_side"],
                                 t=_h2d(t_rows, self.device)), T=T)
        self.save_moments(None, cells.O, cells.n_states, T, Tpred, self.ego_vehicle_id, params,
                          _precomputed=moments)                                       # :960
        # prob_lower_save[t] is overwritten across (o, k): the last cell wins (:947)
        prob_lower_save = host["prob_lower"][-1].tolist() if "prob_lower" in host else None
        if T == Tpred:
            self.prob_lower_save = prob_lower_save                                    # :961-962
        hp.extra["prob_lower_save"] = np.asarray(prob_lower_save)
        constraints = hp.cvxpy_constraints(temp_x) if temp_x is not None else hp
        return (constraints, self._vertices(cells), A_union, b_union, OVconstraint, direct, ov_mean, ov_cov, 0)

    def compute_obstacle_constraints_GMM_affine(
            self, params, ovehicles, Delta2=None, Omicron=None, temp_x=None, eps_ura=None, segments=None,
            Tsh=None, ref_traj=None):
        """v8ideal:1378-1539 (receding horizon, nominal).  Same 9-tuple; rows are
        ``n.x >= d + offset`` / ``n.x <= d - offset`` (:1503-1515).  ``OVconstraint`` is always False:
        the reference hard-codes ``out_junction = False`` (:1424)."""
        cells = ovehicles if isinstance(ovehicles, CellBatch) else CellBatch.from_ovehicles(ovehicles, self.device)
        Tpred = self.prediction_horizon
        T = int(Tsh if Tsh is not None else Tpred)
        if eps_ura is None:
            eps_ura = eps_allocation(cells.n_states)
        if cells.T != Tpred:
            raise ValueError(f"cells carry horizon {cells.T}; the planner's prediction_horizon is {Tpred}")
        if self.road_boundary_constraints:
            # the reference adds the slack S_big_repeated[0, t] to every row here (:1505, :1511)
            raise NotImplementedError("road-boundary slack S_big (v8ideal:1395-1416) is not part of this path")
        ref_np = np.zeros((Tpred, 2))
        r = np.asarray(ref_traj, dtype=np.float64)[:, :2]
        ref_np[:min(T, r.shape[0])] = r[:T]
        sums, pivot = self._sums(cells)
        out = ops.affine_halfplanes(sums, pivot, _h2d(ref_np, self.device), self._eps_cells(cells, eps_ura), Tpred,
                                    R=self.R)
        At, b0, mom = self._l4(cells, sums, pivot)
        allh = _to_host({**out, "_At": At, "_b0": b0, "_mean": mom["mean"], "_cov": mom["cov"],
                         "_yaw_mean": mom["yaw_mean"], "_yaw_var": mom["yaw_var"]})
        host = {k: v[:, :T] for k, v in allh.items() if not k.startswith("_")}
        raise_for_status(host["status"], what="affine builder")                       # :1499-1503
        A_union = self._nested_tko(cells, allh["_At"], Tpred)
        b_union = self._nested_tko(cells, allh["_b0"], Tpred)
        ov_mean, ov_cov, _ = self._ov_state(cells, allh)
        side = np.where(host["idx_side"] & 2, -1, 1).astype(np.int32)
        C = side.shape[0]
        t_rows = np.broadcast_to(np.arange(T, dtype=np.int32), (C, T)).copy()
        dev_rows = {k: out[k][:, :T].contiguous() for k in ("nd", "offset_rhs", "idx_side")}
        hp = HalfPlanes(n=host["nd"][..., :2], d=host["nd"][..., 2], rhs=host["offset_rhs"][..., 1],
                        idx=(host["idx_side"] & 1).astype(np.int32), side=side,
                        t=t_rows, status=host["status"],
                        cell_obstacle=cells.cell_obstacle, cell_mode=cells.cell_mode,
                        extra=dict(offset=host["offset_rhs"][..., 0], mean=host["mean"], cov=host["cov"]),
                        dev=dict(nd=dev_rows["nd"], rhs=dev_rows["offset_rhs"][..., 1], idx_side=dev_rows["idx_side"],
                                 t=_h2d(t_rows, self.device)), T=T)
        constraints = hp.cvxpy_constraints(temp_x) if temp_x is not None else hp
        OVconstraint = False                                                          # out_junction = False (:1424)
        return (constraints, self._vertices(cells), A_union, b_union, OVconstraint, [None] * cells.O, ov_mean,
                ov_cov, 0)

    # ------------------------------------------------------ MC risk estimate
    def mc_collision_risk(self, moments: Moments, n_states, T: int, ego_xy, n_samples: Optional[int] = None,
                          seed: Optional[int] = None, sample_offset: Optional[int] = None, process_group=None,
                          want_moments: bool = True, x0=None, z=None, want_trajectories: bool = True):
        """Stage 1 of the path: rollout n samples per (o, k) cell from the moments, then ONE read of
        the trajectories -> collision counts (A9), soft cost and the shifted power sums; with a
        ``process_group`` the packed sums are all-reduced (sum) across sample shards.

        ``n_samples`` is the TOTAL per cell: with a ``process_group`` and no explicit ``sample_offset`` this rank
        rolls its ``sharding.sample_shard`` of them (Philox counter = global sample index, so the union of the
        shards does not depend on the world size); an explicit ``sample_offset`` means "roll exactly n_samples
        here, starting at that global index" (weak scaling).

        Returns dict(sums [C, L] device, counts [C, T+1] int64 device, pivot, cells)."""
        from . import sharding
        n = self.n_ideal_samples if n_samples is None else int(n_samples)
        if sample_offset is None:
            sample_offset = 0
            if process_group is not None:
                import torch.distributed as dist
                sample_offset, n = sharding.sample_shard(n, dist.get_rank(process_group),
                                                         dist.get_world_size(process_group))
        # common pivot on every rank: the moment means x_{t+1} ~ mean[t+1]  
        sel, co, ck = self._select_cells(n_states, moments)
        mean, cov, cross = self._gather(moments, sel)
        pivot = mean[:, 1:T + 1].contiguous()
        if z is None and 1 <= T <= ops.MAX_T and moments.T >= T + 1:
            # one kernel: rollout, trajectory store, counts and power sums (ccmpc_mc_rollout_sums)
            if seed is None:
                self._calls += 1
                seed = (self.seed << 20) + self._calls
            ego_t = _h2d(np.asarray(ego_xy, dtype=np.float64), self.device) if not torch.is_tensor(ego_xy) else ego_xy
            x0_t = None if x0 is None else _h2d(np.asarray(x0, dtype=np.float64).reshape(len(sel), 2), self.device)
            traj, sums, status, x0_used = ops.mc_rollout_sums(mean, cov, cross, x0_t, n, T, pivot, ego_t, R=self.R,
                                                              seed=seed, sample_offset=sample_offset,
                                                              materialize=want_trajectories,
                                                              workspace=self._fused_ws(len(sel), n, T))
            cells = CellBatch.uniform(traj, co, ck, n_states) if traj is not None else None
        else:
            cells = self.predict_ideal_cells(n_states, T, moments, n_samples=n, seed=seed,
                                             sample_offset=sample_offset, x0=x0, z=z)
            sums, _ = self._sums(cells, ego=ego_xy, pivot=pivot)
            status = cells.status
        if process_group is not None:
            sharding.allreduce_sums(sums, group=process_group)
        mom = ops.moments_from_sums(sums, pivot, want_cross=want_moments, want_counts=True, want_yaw=False)
        return dict(sums=sums, counts=mom["counts"], pivot=pivot, cells=cells, moments=mom, status=status)

    def _fused_ws(self, C: int, n: int, T: int) -> torch.Tensor:
        key = (C, n, T)
        if getattr(self, "_fused_ws_key", None) != key:
            self._fused_ws_buf = ops.mc_rollout_sums_workspace(C, n, T, self.device)
            self._fused_ws_key = key
        return self._fused_ws_buf

    def _select_cells(self, n_states: Sequence[int], moments: Moments):
        """Rows of ``moments`` for the cells (o, k < K_o), with predict_ideal's fallback to the last saved mode
        when K grew (v8ideal:2652-2656)."""
        key = tuple(int(k) for k in n_states)
        cache = getattr(moments, "_sel_cache", None)
        if cache is not None and cache[0] == key:
            return cache[1]
        sel, co, ck = [], [], []
        Kmax_saved = max(moments.n_states) if moments.n_states else 0
        for o, K_o in enumerate(n_states):
            for k in range(int(K_o)):
                data_idx = k if k < Kmax_saved else max(Kmax_saved - 1, 0)        # v8ideal:2652-2656
                hit = np.nonzero((moments.cell_obstacle == o) & (moments.cell_mode == data_idx))[0]
                if hit.size == 0:
                    raise InSimulationException(f"saved moments have no entry for obstacle {o} mode {data_idx}")
                sel.append(int(hit[0]))
                co.append(o)
                ck.append(k)
        moments._sel_cache = (key, (sel, co, ck))
        return sel, co, ck

    def _gather(self, moments: Moments, sel):
        """(mean, cov, cross) rows ``sel`` of the moments; the tensors themselves when that is every row in order."""
        if sel == list(range(moments.mean.shape[0])):
            return moments.mean, moments.cov, moments.cross
        idx = torch.as_tensor(sel, dtype=torch.int64, device=self.device)
        return (moments.mean.index_select(0, idx).contiguous(), moments.cov.index_select(0, idx).contiguous(),
                moments.cross.index_select(0, idx).contiguous())


class StepPlan:
    """Fixed-shape constraint build of one MPC step with every buffer pre-allocated and the whole
    launch sequence (H2D of the samples and the reference trajectory, yaws, one-pass sums + counts,
    moments, Minkowski half-planes, L4 polytopes, violation counts, D2H of the packed results)
    captured once in a CUDA graph: one graph launch + one stream synchronise per MPC step.

    This is the latency path for BASELINE.json's "MPC-step constraint build p50 ms"; the arithmetic is
    the same kernels :class:`MidlevelConstraints` calls (v8ideal:881-947, :627-640, :694-736)."""

    _FIELDS = (("nd", "f8", ("C", "P", 3)), ("ab", "f8", ("C", "P", 3)), ("prob_lower", "f8", ("C", "T")),
               ("mean", "f8", ("C", "T", 2)), ("cov", "f8", ("C", "T", 2, 2)), ("At", "f8", ("C", "T", 4, 2)),
               ("b0", "f8", ("C", "T", 4)), ("yaw_mean", "f8", ("C", "T")), ("yaw_var", "f8", ("C", "T")),
               ("counts", "i8", ("C", "T1")), ("viol", "i8", ("C", "P")), ("idx_side", "i1", ("C", "P")),
               ("status", "u1", ("C", "P")),
               # affine (receding-horizon) builder, v8ideal:1476-1515
               ("aff_nd", "f8", ("C", "T", 3)), ("aff_offset_rhs", "f8", ("C", "T", 2)),
               ("aff_cov", "f8", ("C", "T", 3)), ("aff_idx_side", "i1", ("C", "T")), ("aff_status", "u1", ("C", "T")))
    _INGEST_FIELDS = (("g_offsets", "i8", ("C1",)), ("g_cell_pmf", "f8", ("C",)), ("g_init_center", "f8", ("C", 2)),
                      ("g_n_states", "i4", ("V",)), ("g_cell_latent", "i4", ("C",)), ("g_status", "u1", ("V",)))
    _MINK_ONLY = ("nd", "ab", "prob_lower", "viol", "idx_side", "status")
    _AFF_ONLY = ("aff_nd", "aff_offset_rhs", "aff_cov", "aff_idx_side", "aff_status")
    _TORCH = {"f8": torch.float64, "i8": torch.int64, "i4": torch.int32, "i1": torch.int8, "u1": torch.uint8}
    _NP = {"f8": np.float64, "i8": np.int64, "i4": np.int32, "i1": np.int8, "u1": np.uint8}

    def __init__(self, offsets_host, cell_obstacle, cell_mode, n_states, pos_last_cell, T: int,
                 prediction_horizon: int, eps_ura=None, device=None, bbox=BBOX_DEFAULT, R: float = ops.R_EV_OV,
                 target_p: float = ops.TARGET_P, accel: int = 0, use_graph: bool = True,
                 builder: str = "minkowski", ingest: Optional[dict] = None):
        if builder not in ("minkowski", "affine", "both"):
            raise ValueError("builder must be 'minkowski' (shrinking horizon, v8ideal:881-947), "
                             "'affine' (receding horizon, v8ideal:1476-1515) or 'both'")
        self.builder = builder
        self.device = _device(device)
        self.offsets_host = np.asarray(offsets_host, dtype=np.int64)
        self.cell_obstacle, self.cell_mode = np.asarray(cell_obstacle), np.asarray(cell_mode)
        self.n_states = list(n_states)
        self.T, self.prediction_horizon = int(T), int(prediction_horizon)
        self.C = int(self.offsets_host.size - 1)
        self.N = int(self.offsets_host[-1])
        self.P = self.T * (self.T - 1) // 2
        self.R, self.target_p, self.accel = float(R), float(target_p), _accel_mode(accel)
        dev = self.device
        if eps_ura is None:
            eps_ura = eps_allocation(self.n_states)
        self.eps_cell = _h2d(np.asarray(eps_ura, dtype=np.float64)[self.cell_obstacle, self.cell_mode], dev)
        self.offsets = _h2d(self.offsets_host, dev)
        self.pos_last = _h2d(np.asarray(pos_last_cell, dtype=np.float64).reshape(self.C, 2), dev)
        self.lw_cell = _h2d(np.broadcast_to(np.asarray(bbox, dtype=np.float64), (self.C, 2)).copy(), dev)
        # ingest mode (from_predictions): raw float32 predictions + latents in, grouping done on the device
        self.ingest = ingest
        self.sampler = None
        if ingest is not None:
            V, Nv, S = int(ingest["V"]), int(ingest["N"]), int(ingest["S"])
            self.host_pred = torch.zeros((V, Nv, self.T, 2), dtype=torch.float32).pin_memory()
            self.host_z = torch.zeros((V, Nv), dtype=torch.int32).pin_memory()
            self.host_pmf = torch.zeros((V, S), dtype=torch.float64).pin_memory()
            self.pred = torch.empty((V, Nv, self.T, 2), dtype=torch.float32, device=dev)
            self.z = torch.empty((V, Nv), dtype=torch.int32, device=dev)
            self.pmf = torch.empty((V, S), dtype=torch.float64, device=dev)
            self.perm_dev = torch.empty((V, Nv), dtype=torch.int64, device=dev)
            self.pos_last_vehicle = self.pos_last.view(V, S, 2)[:, 0].contiguous()
            self.group_ws = torch.empty(int(ops._lib.load().ccmpc_group_workspace_bytes(V, Nv)), dtype=torch.uint8,
                                        device=dev)
            # sampler mode (from_gmm): the predictor's GMM2D parameters in, samples drawn on the device
            self.sampler = ingest.get("sampler")
            if self.sampler is not None:
                G = int(self.sampler["G"])
                f32 = torch.float32
                self.host_gmm = dict(mus=torch.zeros((V, S, self.T, G, 2), dtype=f32).pin_memory(),
                                     log_sigmas=torch.zeros((V, S, self.T, G, 2), dtype=f32).pin_memory(),
                                     corrs=torch.zeros((V, S, self.T, G), dtype=f32).pin_memory(),
                                     log_pis=torch.zeros((V, S, self.T, G), dtype=f32).pin_memory() if G > 1 else None,
                                     seed=torch.zeros(1, dtype=torch.int64).pin_memory(),
                                     x0=(torch.zeros((V, 4), dtype=f32).pin_memory()
                                         if self.sampler.get("dynamics") not in (None, "none") else None))
                self.gmm = {k: (None if v is None else torch.empty(v.shape, dtype=v.dtype, device=dev))
                            for k, v in self.host_gmm.items()}
        self.host_samples = (torch.empty((self.N, self.T, 2), dtype=torch.float64).pin_memory()
                             if ingest is None else None)
        self.host_ref = torch.empty((self.T, 2), dtype=torch.float64).pin_memory()
        self.samples = torch.empty((self.N, self.T, 2), dtype=torch.float64, device=dev)
        self.ref = torch.empty((self.T, 2), dtype=torch.float64, device=dev)
        self.yaws = torch.empty((self.N, self.T), dtype=torch.float64, device=dev)
        self.pivot = torch.empty((self.C, self.T, 2), dtype=torch.float64, device=dev)
        self.sums = torch.empty((self.C, ops.sums_len(self.T)), dtype=torch.float64, device=dev)
        self.ws = ops.sample_sums_workspace(self.N, self.C, self.T, dev)
        # ---- one output arena on the device, one pinned mirror on the host
        dims = {"C": self.C, "P": self.P, "T": self.T, "T1": self.T + 1, "C1": self.C + 1,
                "V": int(ingest["V"]) if ingest is not None else 0}
        off, self._layout = 0, {}
        for name, ty, shp in self._FIELDS + (self._INGEST_FIELDS if ingest is not None else ()):
            if (builder == "affine" and name in self._MINK_ONLY) or (builder == "minkowski" and name in self._AFF_ONLY):
                continue
            shape = tuple(dims.get(s, s) for s in shp)
            nbytes = int(np.prod(shape)) * np.dtype(self._NP[ty]).itemsize
            self._layout[name] = (off, nbytes, ty, shape)
            off = (off + nbytes + 15) // 16 * 16
        self.arena = torch.zeros(max(off, 16), dtype=torch.uint8, device=dev)
        self.host_arena = torch.zeros(max(off, 16), dtype=torch.uint8).pin_memory()
        self.out = {n: self.arena[o:o + b].view(self._TORCH[ty]).view(shape)
                    for n, (o, b, ty, shape) in self._layout.items()}
        hb = self.host_arena.numpy()
        self.host_out = {n: hb[o:o + b].view(self._NP[ty]).reshape(shape)
                         for n, (o, b, ty, shape) in self._layout.items()}
        self._mom = dict(mean=self.out["mean"], cov=self.out["cov"], cross=None, counts=self.out["counts"],
                         yaw_mean=self.out["yaw_mean"], yaw_var=self.out["yaw_var"])
        self._mink = ({k: self.out[k] for k in ("nd", "ab", "idx_side", "prob_lower", "status")}
                      if builder != "affine" else None)
        self._aff = (dict(nd=self.out["aff_nd"], offset_rhs=self.out["aff_offset_rhs"],
                          idx_side=self.out["aff_idx_side"], mean=self.out["mean"], cov=self.out["aff_cov"],
                          status=self.out["aff_status"]) if builder != "minkowski" else None)
        if ingest is not None:
            self.offsets = self.out["g_offsets"]          # written by the grouping kernel every step
            self._group_out = dict(samples=self.samples, offsets=self.out["g_offsets"], perm=self.perm_dev,
                                   n_states=self.out["g_n_states"], cell_latent=self.out["g_cell_latent"],
                                   cell_pmf=self.out["g_cell_pmf"], init_center=self.out["g_init_center"],
                                   status=self.out["g_status"], yaws=self.yaws)
        self._ego_is_ref = True
        self._graph_full = self._graph_dev = None
        self.launches_per_step = None
        if use_graph:
            self._capture()

    @classmethod
    def from_samples(cls, z, pos, pos_last, K: int, prediction_horizon: int, device=None, **kw) -> "StepPlan":
        """Bucket each obstacle's samples by latent z in sample order (v8ideal:489-494) on the host ONCE to fix
        the cell layout; ``host_samples`` is filled with the grouped tensor and ``perm`` keeps the mapping."""
        z = np.asarray(z)
        pos = np.asarray(pos, dtype=np.float64)
        O, N, T = pos.shape[0], pos.shape[1], pos.shape[2]
        perm, offs, co, ck, ns = [], [0], [], [], []
        for o in range(O):
            order = np.argsort(z[o], kind="stable")
            zs = z[o][order]
            ks = np.unique(zs)
            ns.append(int(ks.size))
            for j, k in enumerate(ks):
                sel = order[zs == k]
                perm.append(o * N + sel)
                offs.append(offs[-1] + sel.size)
                co.append(o)
                ck.append(j)
        plan = cls(np.asarray(offs), co, ck, ns, np.asarray(pos_last)[np.asarray(co)], T, prediction_horizon,
                   device=device, **kw)
        plan.perm = np.concatenate(perm)
        plan.host_samples.copy_(torch.from_numpy(pos.reshape(O * N, T, 2)[plan.perm]))
        plan.cell_latent = np.asarray([np.unique(z[o])[k] for o, k in zip(co, ck)])
        return plan

    @classmethod
    def from_predictions(cls, V: int, N: int, T: int, pos_last, prediction_horizon: int, S: int = 25,
                         minpos=(0.0, 0.0), filter_pmf: float = 0.1, device=None, sampler: Optional[dict] = None,
                         **kw) -> "StepPlan":
        """Ingest mode: the planner hands over what ``generate_vehicle_latents`` returns (prediction.py:103-104) -
        ``predictions (V, N, T, 2)`` float32, ``z (V, N)``, ``latent_probs (V, S)`` - and the bucket-by-latent /
        rare-latent fold of v8ideal:486-494 + ovehicle.py:55-113 runs on the device inside the same CUDA graph.
        Fixed cell layout: cell = v * S + j, cells j >= K_v empty (their rows are NaN, ``g_n_states`` says K_v).
        Fill ``host_pred`` / ``host_z`` / ``host_pmf`` (pinned) or pass them to :meth:`run_predictions`."""
        C = V * S
        offs = np.zeros(C + 1, dtype=np.int64)
        offs[-1] = V * N
        co = np.repeat(np.arange(V), S)
        ck = np.tile(np.arange(S), V)
        kw.setdefault("eps_ura", np.full((V, S), EPS_TOTAL / V))            # eps_ura[o, k] = eps / O (v8ideal:2920-2926)
        plan = cls(offs, co, ck, [S] * V, np.asarray(pos_last, dtype=np.float64).reshape(V, 2)[co], T,
                   prediction_horizon, device=device,
                   ingest=dict(V=V, N=N, S=S, minpos=tuple(float(m) for m in minpos), filter_pmf=float(filter_pmf),
                               sampler=sampler),
                   **kw)
        return plan

    @classmethod
    def from_gmm(cls, V: int, N: int, T: int, pos_last, prediction_horizon: int, Z: int = 25, G: int = 1,
                 seed: int = 0, dynamics: Optional[str] = None, dt: float = 0.5, **kw) -> "StepPlan":
        """Sampler mode: the planner hands over the predictor's OUTPUT DISTRIBUTION - latent pmf (V, Z) and GMM2D
        parameters per (vehicle, latent, timestep, component) - instead of N sampled trajectories per vehicle.
        ``model.latent.sample_p`` + ``GMM2D.rsample`` (prediction.py:81-86), the D2H copy of the samples
        (:88-93) and the bucket-by-latent loop (v8ideal:486-494) all run on the device inside one CUDA graph;
        the step's H2D traffic is the parameters (KBs) rather than the sample tensor (MBs).
        ``dynamics``: None when the GMM2D is over positions; "unicycle" / "single_integrator" when it is over
        controls and each sample is integrated from the vehicle state x0 = (x, y, phi, v) with step ``dt`` -
        ``dynamic.integrate_samples``, which is what p_y_xz hands back as `predictions` (prediction.py:85-86)."""
        if dynamics not in ops.DYNAMICS:
            raise ValueError(f"from_gmm: unknown dynamics {dynamics!r}")
        return cls.from_predictions(V, N, T, pos_last, prediction_horizon, S=Z,
                                    sampler=dict(G=int(G), seed=int(seed), dynamics=dynamics, dt=float(dt)), **kw)

    def run_gmm(self, latent_probs=None, mus=None, log_sigmas=None, corrs=None, log_pis=None, ref_traj=None,
                step_seed: Optional[int] = None, x0=None) -> dict:
        """Sampler-mode step: GMM2D parameters in -> host rows out (one graph launch + one sync).  ``step_seed``
        is added to the plan's seed on the device, so every replay of the captured graph draws fresh samples."""
        if self.sampler is None:
            raise RuntimeError("run_gmm needs a plan built by StepPlan.from_gmm")
        for name, val in (("mus", mus), ("log_sigmas", log_sigmas), ("corrs", corrs), ("log_pis", log_pis), ("x0", x0)):
            if val is not None:
                if self.host_gmm[name] is None:
                    raise ValueError(f"{name} given but the plan was built " +
                                     ("without dynamics" if name == "x0" else "with G = 1"))
                self.host_gmm[name].copy_(torch.as_tensor(np.asarray(val, dtype=np.float32)).reshape(self.host_gmm[name].shape))
        if latent_probs is not None:
            self.host_pmf.copy_(torch.as_tensor(np.asarray(latent_probs, dtype=np.float64)).reshape(self.host_pmf.shape))
        if step_seed is not None:
            self.host_gmm["seed"][0] = int(step_seed)
        return self.run(ref_traj=ref_traj)

    def run_predictions(self, predictions=None, z=None, latent_probs=None, ref_traj=None) -> dict:
        """Ingest-mode step: host predictions in -> host rows out (one graph launch + one sync)."""
        if self.ingest is None:
            raise RuntimeError("run_predictions needs a plan built by StepPlan.from_predictions")
        if predictions is not None:
            self.host_pred.copy_(torch.as_tensor(np.asarray(predictions, dtype=np.float32)).reshape(self.host_pred.shape))
        if z is not None:
            self.host_z.copy_(torch.as_tensor(np.asarray(z).astype(np.int32)).reshape(self.host_z.shape))
        if latent_probs is not None:
            self.host_pmf.copy_(torch.as_tensor(np.asarray(latent_probs, dtype=np.float64)).reshape(self.host_pmf.shape))
        return self.run(ref_traj=ref_traj)

    # -------------------------------------------------------------- launches
    def _enqueue_device(self, yaws_done: bool = False):
        if not yaws_done:                   # ingest mode: the grouping kernel already wrote the yaws of its rows
            ops.yaws_from_positions(self.samples, self.offsets, self.pos_last, out=self.yaws)
        ops.cell_pivot(self.samples, self.offsets, out=self.pivot)
        ops.sample_sums(self.samples, self.offsets, self.pivot, ego=self.ref, yaws=self.yaws, R=self.R,
                        workspace=self.ws, out=self.sums)
        ops.moments_from_sums(self.sums, self.pivot, out=self._mom)
        if self._mink is not None:
            ops.minkowski_halfplanes(self.sums, self.pivot, self.ref, self.eps_cell, self.prediction_horizon,
                                     R=self.R, target_p=self.target_p, accel=self.accel, out=self._mink)
        if self._aff is not None:
            ops.affine_halfplanes(self.sums, self.pivot, self.ref, self.eps_cell, self.prediction_horizon, R=self.R,
                                  out=self._aff)
        ops.l4_outer(self.samples, self.yaws, self.offsets, self.out["yaw_mean"], lw_cell=self.lw_cell,
                     out=(self.out["At"], self.out["b0"]))
        if self._mink is not None:
            ops.violation_counts(self.samples, self.offsets, self.out["nd"], self.out["idx_side"],
                                 out=self.out["viol"])

    def _enqueue_ingest(self):
        """Device grouping of the raw predictions (v8ideal:486-494 + ovehicle.py:55-113) into the CSR layout."""
        ops.group_by_mode(self.pred, self.z, self.pmf, minpos=self.ingest["minpos"],
                          filter_pmf=self.ingest.get("filter_pmf", 0.1), workspace=self.group_ws, out=self._group_out,
                          pos_last=self.pos_last_vehicle)

    def _enqueue_full(self):
        if self.sampler is not None:
            # stage 1 on the device: a few KB of GMM2D parameters in, samples never cross PCIe (prediction.py:81-93)
            for k, h in self.host_gmm.items():
                if h is not None:
                    self.gmm[k].copy_(h, non_blocking=True)
            self.pmf.copy_(self.host_pmf, non_blocking=True)
            self.ref.copy_(self.host_ref, non_blocking=True)
            ops.gmm2d_sample(self.pmf, self.gmm["mus"], self.gmm["log_sigmas"], self.gmm["corrs"], self.pred.shape[1],
                             log_pis=self.gmm["log_pis"], seed=int(self.sampler.get("seed", 0)),
                             seed_dev=self.gmm["seed"], out=(self.z, self.pred),
                             dynamics=self.sampler.get("dynamics"), x0=self.gmm["x0"], dt=self.sampler.get("dt", 0.5))
            self._enqueue_ingest()
        elif self.ingest is not None:
            self.pred.copy_(self.host_pred, non_blocking=True)
            self.z.copy_(self.host_z, non_blocking=True)
            self.pmf.copy_(self.host_pmf, non_blocking=True)
            self.ref.copy_(self.host_ref, non_blocking=True)
            self._enqueue_ingest()
        else:
            self.samples.copy_(self.host_samples, non_blocking=True)
            self.ref.copy_(self.host_ref, non_blocking=True)
        self._enqueue_device(yaws_done=self.ingest is not None)
        self.host_arena.copy_(self.arena, non_blocking=True)

    def _capture(self):
        from . import _lib
        lib = _lib.load()
        if self.ingest is None:
            self.samples.copy_(self.host_samples)
        self.ref.copy_(self.host_ref)
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                      # warm-up outside capture
            l0 = lib.ccmpc_launch_count()
            self._enqueue_full()
            self.launches_per_step = int(lib.ccmpc_launch_count() - l0)
        side.synchronize()
        self._graph_full, self._graph_dev = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph_full, stream=side):
            self._enqueue_full()
        with torch.cuda.graph(self._graph_dev, stream=side):
            self._enqueue_device()
        torch.cuda.current_stream().wait_stream(side)

    # ------------------------------------------------------------------ runs
    def run_device(self):
        """Samples and reference already in ``self.samples`` / ``self.ref``; results stay in ``self.out``."""
        if self._graph_dev is not None:
            self._graph_dev.replay()
        else:
            self._enqueue_device()

    def run(self, pos_host=None, ref_traj=None) -> dict:
        """Host in -> host out.  ``pos_host`` (N, T, 2) grouped samples (``None``: reuse ``host_samples``),
        ``ref_traj`` (>=T, >=2).  Returns NumPy views of the pinned result buffer (valid until the next run)."""
        if pos_host is not None and pos_host is not self.host_samples:
            self.host_samples.copy_(torch.as_tensor(pos_host).reshape(self.N, self.T, 2))
        if ref_traj is not None:
            self.host_ref.copy_(torch.from_numpy(np.ascontiguousarray(np.asarray(ref_traj, dtype=np.float64)[:self.T, :2])))
        if self._graph_full is not None:
            self._graph_full.replay()
        else:
            self._enqueue_full()
        torch.cuda.current_stream().synchronize()
        return self.host_out
