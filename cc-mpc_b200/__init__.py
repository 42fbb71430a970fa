"""ccmpc_b200 - B200-native prediction -> chance-constraint hot path of CC-MPC.

The product is ``libccmpc.so`` (hand-written sm_100a CUDA behind the C ABI of
``include/ccmpc.h``).  This package is the thin Python host side the reference
planner calls:

  ccmpc_b200.makeconstraint   drop-in twin of v8ideal/makeconstraint.py
  ccmpc_b200.npu              vertices_of_bboxes (utility/npu.py:308-350)
  ccmpc_b200.midlevel         MidlevelConstraints: the hot methods of v8ideal.MidlevelAgent
  ccmpc_b200.cost             cost_final / getData_singleEpisode (dataForCost_ref.py)
  ccmpc_b200.ops              batched tensor twins over the C ABI
  ccmpc_b200.sharding         sample / scenario sharding over GPUs + the one all-reduce

There is no CPU, NumPy or PyTorch fallback: importing is cheap, but every
compute call needs the CUDA library and a CUDA device and fails loudly otherwise.
"""
from . import _build, _lib  # noqa: F401

__version__ = "0.2.0"


def build(force: bool = False) -> str:
    """Compile libccmpc.so in-tree (nvcc, sm_100a)."""
    return _build.build_library(force=force)


def load():
    """Load libccmpc.so; raises when it is missing and cannot be built."""
    return _lib.load()
