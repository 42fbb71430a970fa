"""Seeded synthetic inputs moved to ``ccmpc_b200.synth`` (they are generators, not reference arithmetic, and the
product arm of bench.py must not import anything from ``oracle/``); this shim keeps ``oracle.synth`` importable
for the tests and tools that grew up with it."""
from ccmpc_b200.synth import *  # noqa: F401,F403
from ccmpc_b200.synth import AR1, DT, GOAL, MINPOS  # noqa: F401
