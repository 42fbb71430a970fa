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


# --------------------------------------------------------------------------
# Executing single functions / methods of modules that cannot be IMPORTED here
# --------------------------------------------------------------------------
# v8ideal/__init__.py, midlevel/util.py, midlevel/ovehicle.py and
# dataForCost_ref.py import carla / cvxpy / docplex / matplotlib / control /
# openpyxl at module level, but the bodies of the functions on the hot path use
# only numpy / scipy / pickle / os / logging / copy.  `extract` parses the
# reference file, takes the named FunctionDef nodes (nothing else of the file
# is compiled or executed, nothing is copied into this repo) and executes them
# in a caller-supplied namespace.  Methods are re-wrapped in a ClassDef of the
# reference's class name so that private-name mangling (`self.__params`,
# `self.__compute_vertices`) resolves exactly as in the reference.
import ast

V8IDEAL = "collect/in_simulation/midlevel/v8ideal/__init__.py"
MIDLEVEL_UTIL = "collect/in_simulation/midlevel/util.py"
OVEHICLE = "collect/in_simulation/midlevel/ovehicle.py"
DATA_FOR_COST = "dataForCost_ref.py"


def _find_defs(tree, names, class_name=None):
    body = tree.body
    if class_name is not None:
        cls = [n for n in body if isinstance(n, ast.ClassDef) and n.name == class_name]
        if not cls:
            raise KeyError(class_name)
        body = cls[0].body
    found = {n.name: n for n in body if isinstance(n, ast.FunctionDef) and n.name in names}
    missing = [n for n in names if n not in found]
    if missing:
        raise KeyError(f"{missing} not found")
    return [found[n] for n in names]


def extract(rel: str, names, namespace: dict, class_name: str = None):
    """Compile ONLY the named functions of reference file ``rel`` and exec them in ``namespace``.

    With ``class_name`` the functions are that class's methods; a class of the same name holding only
    those methods is created in ``namespace`` (decorators such as ``@classmethod`` are kept,
    ``@profile``-style project decorators must be present in the namespace).  Returns the namespace."""
    path = os.path.join(REFERENCE_ROOT, rel)
    import warnings
    with open(path, "r", encoding="utf-8") as f, warnings.catch_warnings():
        warnings.simplefilter("ignore", SyntaxWarning)       # plot labels elsewhere in the file
        tree = ast.parse(f.read(), filename=path)
    defs = _find_defs(tree, list(names), class_name)
    if class_name is None:
        mod = ast.Module(body=defs, type_ignores=[])
    else:
        cls = ast.ClassDef(name=class_name, bases=[ast.Name(id="object", ctx=ast.Load())], keywords=[],
                           body=defs, decorator_list=[])
        if "type_params" in ast.ClassDef._fields:
            cls.type_params = []
        mod = ast.Module(body=[cls], type_ignores=[])
    ast.fix_missing_locations(mod)
    exec(compile(mod, path, "exec"), namespace)      # line numbers stay the reference's own
    return namespace
