"""Import the reference's own importable modules BY FILE PATH.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Only usable where
``/root/reference`` exists (the build container); it does not exist on the GPU
box, so nothing that runs there may depend on this module - the committed
``tests/golden/ref_*.npz`` vectors carry the reference's outputs instead.

Importable (numpy + scipy only):
  collect/in_simulation/midlevel/v8ideal/makeconstraint.py
  python-utility/utility/utility/npu.py
Not importable here (carla / cvxpy / docplex / matplotlib / control / openpyxl):
  v8ideal/__init__.py, midlevel/util.py, midlevel/ovehicle.py, dataForCost_ref.py
"""
from __future__ import annotations

import importlib.util
import os

REFERENCE_ROOT = os.environ.get("CCMPC_REFERENCE_ROOT", "/root/reference")

_MAKECONSTRAINT = "collect/in_simulation/midlevel/v8ideal/makeconstraint.py"
_NPU = "python-utility/utility/utility/npu.py"


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, _MAKECONSTRAINT))


def _load(name: str, rel: str):
    path = os.path.join(REFERENCE_ROOT, rel)
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_cache = {}


def makeconstraint():
    """The reference's makeconstraint module (functions :7-303)."""
    if "mc" not in _cache:
        _cache["mc"] = _load("_ref_makeconstraint", _MAKECONSTRAINT)
    return _cache["mc"]


def npu():
    """The reference's utility.npu module (vertices_of_bboxes :308-350).

    npu.py does ``from . import UtilityException`` (:8), so it is imported as a
    member of its own package with python-utility/utility on sys.path (the
    reference's py38.env.sh does the same through PYTHONPATH)."""
    if "npu" not in _cache:
        import sys
        import warnings
        root = os.path.join(REFERENCE_ROOT, "python-utility", "utility")
        sys.path.insert(0, root)
        try:
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                import utility.npu as ref_npu  # noqa: the reference's package
        finally:
            sys.path.remove(root)
        _cache["npu"] = ref_npu
    return _cache["npu"]
