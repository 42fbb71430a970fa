"""Twin of the reference's ``utility.npu`` functions on the hot path.

python-utility/utility/utility/npu.py:308-350 ``vertices_of_bboxes`` and
collect/in_simulation/midlevel/util.py:171-200 ``compute_L4_outerapproximation``
(same names, NumPy in / NumPy out; device tensors in -> device tensors out).
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops


def _dev():
    if not torch.cuda.is_available():
        raise RuntimeError("ccmpc_b200 needs a CUDA device (there is no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def vertices_of_bboxes(centers, theta, lw):
    """Vertices of N boxes, shape (N, 4, 2) (npu.py:308-350).

    ``centers`` (N, 2); ``theta`` (N,) or a scalar shared by all boxes; ``lw`` (2,) or (N, 2)
    (a per-box (N, 2) ``lw`` is handled by grouping equal rows)."""
    on_device = isinstance(centers, torch.Tensor)
    dev = centers.device if on_device else _dev()
    c = centers if on_device else torch.from_numpy(np.ascontiguousarray(centers, dtype=np.float64)).to(dev)
    if c.ndim != 2 or c.shape[1] != 2:
        raise ValueError("centers must be (N, 2)")
    N = c.shape[0]
    if isinstance(theta, torch.Tensor):
        th = theta.reshape(-1)
    elif np.ndim(theta) == 0:
        th = torch.full((N,), float(theta), dtype=torch.float64, device=dev)
    else:
        th = torch.from_numpy(np.ascontiguousarray(theta, dtype=np.float64).ravel()).to(dev)
    if th.numel() != N:
        raise ValueError("theta must be a scalar or have N entries")
    lw_np = np.asarray(lw.cpu() if isinstance(lw, torch.Tensor) else lw, dtype=np.float64)
    if lw_np.ndim == 1:
        out = ops.vertices_of_bboxes(c.contiguous(), th.contiguous(), lw_np)
    else:
        if lw_np.shape != (N, 2):
            raise ValueError("lw must be (2,) or (N, 2)")
        out = torch.empty((N, 4, 2), dtype=torch.float64, device=dev)
        uniq, inv = np.unique(lw_np, axis=0, return_inverse=True)
        for j in range(uniq.shape[0]):
            sel = torch.from_numpy(np.nonzero(inv.ravel() == j)[0]).to(dev)
            out[sel] = ops.vertices_of_bboxes(c[sel].contiguous(), th[sel].contiguous(), uniq[j])
    return out if on_device else out.cpu().numpy()


def compute_L4_outerapproximation(theta, vertices, lw=None):
    """4-face outer approximation of a set of boxes (midlevel/util.py:171-200).

    ``vertices`` (N, 4, 2) corners (as ``vertices_of_bboxes`` returns).  The kernel works
    from centres, yaws and box size, which are recovered exactly from the corners:
    centre = mean of the 4 corners, yaw = atan2 of the (v2 - v1) edge, (l, w) = edge lengths.
    Returns (At (4, 2), b0 (4,))."""
    on_device = isinstance(vertices, torch.Tensor)
    dev = vertices.device if on_device else _dev()
    v = vertices if on_device else torch.from_numpy(np.ascontiguousarray(vertices, dtype=np.float64)).to(dev)
    if v.ndim != 3 or tuple(v.shape[1:]) != (4, 2):
        raise ValueError("vertices must be (N, 4, 2)")
    # the projection max is taken over the supplied corners themselves: treat each corner as a
    # zero-size box so the result does not depend on re-deriving (l, w, yaw) from rounded corners
    pts = v.reshape(-1, 1, 2).contiguous()
    n = pts.shape[0]
    yaws = torch.zeros((n, 1), dtype=torch.float64, device=dev)
    offsets = torch.tensor([0, n], dtype=torch.int64, device=dev)
    ym = torch.full((1, 1), float(theta), dtype=torch.float64, device=dev)
    At, b0 = ops.l4_outer(pts, yaws, offsets, ym, lw=(0.0, 0.0))
    At, b0 = At.reshape(4, 2), b0.reshape(4)
    return (At, b0) if on_device else (At.cpu().numpy(), b0.cpu().numpy())
