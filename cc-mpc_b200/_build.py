"""Build libccmpc.so (hand-written sm_100a CUDA) in-tree with nvcc.

The library is the product: there is no CPU or PyTorch fallback.  It is built
next to this file so that it travels with the repo snapshot to the GPU box.
"""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libccmpc.so")
HASH_FILE = os.path.join(HERE, "libccmpc.srchash")
SOURCES = ["abi.cu", "mvoe.cu", "stats.cu", "rollout.cu", "l4.cu", "cost.cu", "group.cu", "gmm.cu", "risk.cu", "fused.cu", "block.cu"]
HEADERS = ["common.cuh", "philox.cuh", "rollout.cuh", "sums_common.cuh", "tma.cuh", os.path.join("..", "..", "include", "ccmpc.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
]


def _sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.isfile(os.path.join(CSRC, s))]


def source_hash() -> int:
    """64-bit digest of every source the library is built from (csrc/*.cu, *.cuh, include/ccmpc.h).  It is
    compiled into the library (``ccmpc_source_hash()``) and written next to it, so a stale .so is recognised by
    content, not by file times (the .so travels to the GPU box in a snapshot whose mtimes mean nothing)."""
    import hashlib
    h = hashlib.sha256()
    for d in sorted(_sources() + [os.path.normpath(os.path.join(CSRC, x)) for x in HEADERS]):
        h.update(os.path.basename(d).encode())
        with open(d, "rb") as f:
            h.update(f.read())
    return int.from_bytes(h.digest()[:8], "little")


def is_stale() -> bool:
    if not os.path.isfile(LIB) or not os.path.isfile(HASH_FILE):
        return True
    with open(HASH_FILE) as f:
        return f.read().strip() != f"{source_hash():016x}"


def find_nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.isfile(cand):
            return cand
    return None


def _file_digest(paths) -> str:
    import hashlib
    h = hashlib.sha256()
    for d in paths:
        with open(d, "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu into libccmpc.so for sm_100a.  Returns the path.

    Each translation unit is compiled on its own (in parallel) into ``build/<name>.<digest>.o`` - the digest
    covers the file and every header - and the objects are linked with ``nvcc -shared``; only changed units
    are recompiled."""
    if not force and not is_stale():
        return LIB
    nvcc = find_nvcc()
    if nvcc is None:
        raise RuntimeError("nvcc not found: cannot build libccmpc.so (no CPU fallback exists)")
    from concurrent.futures import ThreadPoolExecutor
    digest = source_hash()
    obj_dir = os.path.join(HERE, "build")
    os.makedirs(obj_dir, exist_ok=True)
    headers = [os.path.normpath(os.path.join(CSRC, x)) for x in HEADERS]
    extra_env = os.environ.get("CCMPC_EXTRA_NVCC_FLAGS", "").split()     # e.g. -DCCMPC_WS_EXPERIMENTS (tools only)
    flags = [f for f in NVCC_FLAGS if f != "-shared"] + extra_env

    def compile_one(src):
        name = os.path.basename(src)
        extra = [f"-DCCMPC_SRC_HASH=0x{digest:016x}ULL"] if name == "abi.cu" else []
        tag = _file_digest([src] + headers) + (f"{digest:016x}" if extra else "") + \
            ("x" + format(abs(hash(tuple(extra_env))) % (1 << 32), "08x") if extra_env else "")
        obj = os.path.join(obj_dir, f"{name}.{tag}.o")
        if os.path.isfile(obj) and not (force and verbose):
            return obj, ""
        for old in os.listdir(obj_dir):
            if old.startswith(name + ".") and old.endswith(".o"):
                os.remove(os.path.join(obj_dir, old))
        cmd = [nvcc] + flags + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        res = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
        return obj, res.stderr

    with ThreadPoolExecutor(max_workers=min(len(_sources()), os.cpu_count() or 4)) as ex:
        results = list(ex.map(compile_one, _sources()))
    if verbose:
        print("".join(log for _, log in results))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + [o for o, _ in results]
    res = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    with open(HASH_FILE, "w") as f:
        f.write(f"{digest:016x}\n")
    return LIB


if __name__ == "__main__":
    import sys
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
