"""ctypes binding of libccmpc.so (the C ABI declared in include/ccmpc.h).

Loading fails loudly when the CUDA library is missing: the product path has no
CPU, NumPy or PyTorch fallback.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_double, c_int, c_int64, c_size_t, c_uint64, c_void_p

from . import _build

_P = c_void_p

# name -> (restype, argtypes); mirrors include/ccmpc.h one to one
SIGNATURES = {
    "ccmpc_version": (c_int, []),
    "ccmpc_last_error_string": (c_char_p, []),
    "ccmpc_launch_count": (c_uint64, []),
    "ccmpc_source_hash": (c_uint64, []),
    "ccmpc_mvoe_fit": (c_int, [_P, _P, c_double, c_int64, c_double, c_int, c_int, _P, _P, _P, _P, _P]),
    "ccmpc_mvoe_halfplane_sweep": (c_int, [_P, _P, _P, c_int64, c_int, c_int, c_int, c_double, c_double, c_double,
                                            c_int, c_int, _P, _P, _P, _P, _P]),
    "ccmpc_tangent": (c_int, [_P, _P, _P, _P, _P, _P, c_int64, _P, _P, _P, _P, _P]),
    "ccmpc_scale_lower": (c_int, [_P, _P, _P, _P, _P, c_double, c_int64, _P, _P, _P]),
    "ccmpc_sums_len": (c_int64, [c_int]),
    "ccmpc_sample_sums_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int]),
    "ccmpc_sample_sums": (c_int, [_P, _P, c_int64, c_int64, c_int, _P, _P, c_int64, c_double, _P, _P, _P, c_size_t, _P]),
    "ccmpc_sample_sums_static_max_bytes": (c_size_t, []),
    "ccmpc_sample_sums_static": (c_int, [_P, _P, c_int64, c_int64, c_int, _P, _P, c_int64, c_double, _P, _P, _P]),
    "ccmpc_moments_from_sums": (c_int, [_P, _P, c_int64, c_int, _P, _P, _P, _P, _P, _P, _P]),
    "ccmpc_cell_pivot": (c_int, [_P, _P, c_int64, c_int, _P, _P]),
    "ccmpc_minkowski_halfplanes": (c_int, [_P, _P, _P, c_int64, _P, c_int64, c_int, c_int, c_double, c_double,
                                            c_double, c_int, c_int] + [_P] * 12 + [_P]),
    "ccmpc_affine_halfplanes": (c_int, [_P, _P, _P, c_int64, _P, c_int64, c_int, c_int, c_double] + [_P] * 6 + [_P]),
    "ccmpc_violation_counts": (c_int, [_P, _P, c_int64, c_int64, c_int, _P, _P, _P, _P]),
    "ccmpc_vertices_of_bboxes": (c_int, [_P, _P, c_int64, c_double, c_double, c_int64, _P, _P]),
    "ccmpc_l4_outer": (c_int, [_P, _P, _P, c_int64, c_int64, c_int, _P, _P, c_double, c_double, _P, _P, _P]),
    "ccmpc_mc_rollout": (c_int, [_P, _P, _P, c_int, _P, c_int64, c_int64, c_int, c_uint64, c_uint64, _P, _P, _P, _P, _P,
                                  _P]),
    "ccmpc_yaws_from_positions": (c_int, [_P, _P, c_int64, c_int64, c_int, _P, _P, _P]),
    "ccmpc_group_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "ccmpc_group_by_mode": (c_int, [_P, _P, _P, _P, c_int64, c_int64, c_int, c_int, c_double, c_double, c_double,
                                     _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, c_size_t, _P]),
    "ccmpc_episode_cost": (c_int, [_P, _P, _P, c_int64, c_int, c_double, c_double, _P, _P, _P]),
    "ccmpc_halfplanes_to_block": (c_int, [_P, c_int, _P, c_int, _P, _P, c_int64, c_int, c_int, c_int, _P, c_int64, _P, _P,
                                           _P, _P, _P, _P, _P]),
    "ccmpc_philox4x32": (c_int, [_P, _P, c_int64, _P, _P]),
    "ccmpc_mc_risk_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int]),
    "ccmpc_mc_risk": (c_int, [_P, _P, _P, c_int, _P, c_int64, c_int64, c_int, c_uint64, c_uint64, _P, c_int64, c_double,
                               _P, _P, _P, _P, c_size_t, _P]),
    "ccmpc_mc_rollout_sums_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int]),
    "ccmpc_mc_rollout_sums": (c_int, [_P, _P, _P, c_int, _P, c_int64, c_int64, c_int, c_uint64, c_uint64, _P, _P, c_int64,
                                       c_double, _P, _P, _P, _P, _P, c_size_t, _P]),
    "ccmpc_gmm2d_sample": (c_int, [_P, _P, _P, _P, _P, c_int64, c_int64, c_int, c_int, c_int, c_uint64, _P,
                                    c_uint64, _P, _P, _P, _P, _P, _P]),
    "ccmpc_gmm2d_sample_dynamics": (c_int, [_P, _P, _P, _P, _P, c_int, _P, c_double, c_int64, c_int64, c_int, c_int,
                                             c_int, c_uint64, _P, c_uint64, _P, _P, _P, _P, _P, _P]),
}


class CcmpcError(RuntimeError):
    """A libccmpc entry point returned a negative status."""


_lib = None


def library_path() -> str:
    # CCMPC_LIB: an alternative build of the SAME sources (tools/: compile-time experiment variants)
    return os.environ.get("CCMPC_LIB") or _build.LIB


def load(build_if_missing: bool = True):
    """Load libccmpc.so (building it with nvcc when absent).  Raises if impossible."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.isfile(path):
        if not build_if_missing:
            raise CcmpcError(f"{path} not found (run `python -m ccmpc_b200._build`); there is no fallback path")
        _build.build_library()
    elif path == _build.LIB and _build.is_stale():
        # a .so older than csrc/ or include/ccmpc.h would be called with the NEW argument lists below:
        # rebuild when a compiler is here, refuse otherwise (no silent ABI mismatch)
        if _build.find_nvcc() is None:
            raise CcmpcError(f"{path} is older than its sources and nvcc is not available to rebuild it")
        _build.build_library(force=True)
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError if the symbol is missing: fail loudly
        fn.restype = res
        fn.argtypes = args
    if lib.ccmpc_version() <= 0:
        raise CcmpcError("libccmpc.so reports a bad version")
    if int(lib.ccmpc_source_hash()) != _build.source_hash():
        raise CcmpcError(f"{path} was built from other sources than the ones next to it "
                         f"(run `python -m ccmpc_b200._build --force`)")
    _lib = lib
    return lib


def last_error() -> str:
    return load().ccmpc_last_error_string().decode("utf-8", "replace")


def check(rc: int, what: str):
    if rc != 0:
        raise CcmpcError(f"{what} failed (rc={rc}): {last_error()}")
