"""CPU-only checks: the C-ABI library loads and exports every symbol include/ccmpc.h declares
(no compute calls), the host logic (sharding, eps allocation), and the no-fallback guarantee."""
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "ccmpc.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ccmpc_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from ccmpc_b200 import _lib
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ccmpc.h but not exported by libccmpc.so"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    assert sorted(_lib.SIGNATURES) == names, "ctypes table and header disagree"
    assert lib.ccmpc_version() == 200
    assert lib.ccmpc_sums_len(8) == 1 + 16 + 16 * 17 // 2 + 8 + 1 + 8 + 8 + 8 + 8
    assert lib.ccmpc_sums_len(0) == -1 and lib.ccmpc_sums_len(17) == -1
    assert lib.ccmpc_sample_sums_workspace_bytes(5000 * 4, 100, 8) > 0


def test_library_is_sm100a_only():
    import subprocess
    from ccmpc_b200 import _build
    out = subprocess.run(["cuobjdump", "-lelf", _build.LIB], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_no_cpu_fallback():
    from ccmpc_b200 import ops
    with pytest.raises(TypeError):
        ops.mvoe_fit(torch.ones((4, 3), dtype=torch.float64))
    with pytest.raises(TypeError):
        ops.sample_sums(torch.zeros((4, 2, 2), dtype=torch.float64), torch.tensor([0, 4]), torch.zeros((1, 2, 2)))
    if not torch.cuda.is_available():
        from ccmpc_b200 import makeconstraint
        with pytest.raises(RuntimeError):
            makeconstraint.compute_mvoe(np.eye(2), np.eye(2))


def test_sums_layout_mirror():
    from ccmpc_b200 import _lib, ops
    for T in range(1, 17):
        assert ops.sums_layout(T)["len"] == _lib.load().ccmpc_sums_len(T)


def test_eps_allocation_matches_oracle():
    from ccmpc_b200 import midlevel
    from oracle import restate
    for ks in ([3], [1, 4, 2], [25] * 4):
        np.testing.assert_array_equal(midlevel.eps_allocation(ks), restate.eps_allocation(ks))


def test_shard_range_partitions():
    from ccmpc_b200 import sharding
    for total in (0, 1, 7, 4096, 1_000_000):
        for world in (1, 2, 3, 8):
            r = [sharding.shard_range(total, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == total
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    from ccmpc_b200 import ops, sharding
    from oracle import restate, synth
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        T, n = 4, 1001
        sc = synth.make_scene(5, 1, 2, T)
        _, pos = synth.sample_scene(sc, n, 5)
        ego = sc.ref_traj + np.array([0.0, 8.0])       # close enough to collide sometimes
        lo, n_loc = sharding.sample_shard(n)
        L = ops.sums_layout(T)
        sums = torch.zeros((1, L["len"]), dtype=torch.float64)
        part = pos[0, lo:lo + n_loc]
        coll, any_, pen, d2s = restate.collision_counts(part, ego, R=30.0)
        sums[0, 0] = n_loc
        sums[0, L["coll"]:L["coll"] + T] = torch.from_numpy(coll.astype(np.float64))
        sums[0, L["any"]] = float(any_)
        sums[0, L["pen"]:L["pen"] + T] = torch.from_numpy(pen)
        sharding.allreduce_sums(sums)
        counts = sharding.counts_from_sums(sums, T)
        wc, wa, wp, _ = restate.collision_counts(pos[0], ego, R=30.0)
        ok = (int(sums[0, 0]) == n and np.array_equal(counts[0, :T].numpy(), wc) and int(counts[0, T]) == wa
              and np.allclose(sums[0, L["pen"]:L["pen"] + T].numpy(), wp, rtol=1e-12) and wc.sum() > 0)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_count_allreduce_world2_gloo():
    """N > 1 host path on CPU: per-shard counts all-reduced over gloo equal the unsharded counts exactly."""
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]


def test_cov_side_file_roundtrip(tmp_path):
    """MidlevelConstraints.cov_payload / save_data write what dataForCost_ref.get_data (:49-68) and
    montecarlo_cost_ref.py:11 (file-name regex) read; load_data returns meanNtangent of frame - 10 (v8ideal:2547-2557)."""
    import types
    from ccmpc_b200 import cost, midlevel
    mc = midlevel.MidlevelConstraints.__new__(midlevel.MidlevelConstraints)      # no device needed for the file helpers
    mc.out_dir, mc.frame_stride = str(tmp_path), 10
    built = ("rows", None, None, None, True, [None], ([[1.0]], [[2.0]], [[0.1]]), ([[0.5]], [[0.6]], [[0.01]]), 0)
    for i, frame in enumerate((100, 110, 120)):
        d = mc.cov_payload(built, shrinking=True, params={"frame": frame})
        assert d["MeanCov"] is True and d["meanNtangent"] == 0 and d["OVconstraint"] is True
        d.update(solve_time=0.1 * (i + 1), process_time=0.2, x_init=np.zeros(4), cost=1.0, goal=np.zeros(2),
                 U_star=np.full((8, 2), 0.1 * i), X_star=np.full((8, 4), float(i)), timeout=False, infeasible=False)
        assert mc.save_data(d, types.SimpleNamespace(frame=frame), 7).endswith(f"agent7_frame{frame}_cov")
    assert mc.load_data(types.SimpleNamespace(frame=110), 7) == 0
    assert cost.scan_episodes(str(tmp_path)) == {7: (100, 120)}
    ep = cost.getData_singleEpisode(7, 100, 3, root=str(tmp_path) + "/")
    avrP, avrS, avrSOV, worst, X_star, U_star, goal, x_init, timeout = ep.get_data()
    assert abs(avrS - 0.2) < 1e-12 and abs(worst - 0.3) < 1e-12 and timeout == [False] * 3
    assert np.allclose(X_star[2], 2.0) and np.allclose(U_star[1], 0.1)


def test_error_convention_argument_checks():
    """SURVEY 8(b) error convention of the C ABI: 0 = ok, negative code + ccmpc_last_error_string() otherwise.
    Argument validation returns before any CUDA call, so this runs without a GPU (no compute is issued: every
    call below is rejected, the pointers are never dereferenced)."""
    import ctypes
    from ccmpc_b200 import _lib
    lib = _lib.load()
    buf = (ctypes.c_double * 64)()
    p = ctypes.cast(buf, ctypes.c_void_p).value & ~15
    rc = lib.ccmpc_mvoe_fit(None, None, 0.0, -1, 1e-8, 1000, 0, None, None, None, None, None)
    assert rc == -1 and "ccmpc_mvoe_fit" in _lib.last_error()
    rc = lib.ccmpc_mvoe_fit(None, None, 0.0, 4, 1e-8, 1000, 0, None, None, None, None, None)
    assert rc == -1 and "null" in _lib.last_error()
    rc = lib.ccmpc_gmm2d_sample_dynamics(None, None, None, None, None, 7, None, 0.5, 1, 1, 1, 1, 1, 0, None, 0, None,
                                         None, None, None, None, None)
    assert rc == -1 and "dynamics" in _lib.last_error()
    rc = lib.ccmpc_gmm2d_sample(p, None, p, p, p, 1, 1, 1, 1, 2, 0, None, 0, None, None, None, None, p, None)
    assert rc == -1 and "log_pis" in _lib.last_error()                       # G > 1 without mixture weights
    rc = lib.ccmpc_sample_sums(p, p, 10, 1, 17, p, None, 0, 3.4, None, p, None, 0, None)
    assert rc == -1 and "T" in _lib.last_error()                             # horizon beyond CCMPC_MAX_T
    rc = lib.ccmpc_sample_sums(p, p, 10, 1, 8, p, None, 0, 3.4, None, p, None, 0, None)
    assert rc == -3 and "workspace" in _lib.last_error()                     # CCMPC_E_WORKSPACE
    with pytest.raises(_lib.CcmpcError, match="workspace"):
        _lib.check(rc, "ccmpc_sample_sums")
    # 16-byte alignment of vector-accessed pointers is checked before anything is launched (a misaligned address
    # would be a sticky CUDA fault)
    rc = lib.ccmpc_mvoe_fit(p + 8, None, 11.56, 4, 1e-8, 1000, 0, p, p, None, None, None)
    assert rc == -1 and "aligned" in _lib.last_error()
    rc = lib.ccmpc_mc_rollout(p, p, p, 9, None, 1, 4, 8, 0, 0, None, p + 8, None, None, None, None)
    assert rc == -1 and "aligned" in _lib.last_error()
    # empty problems are accepted without touching the device
    assert lib.ccmpc_gmm2d_sample(None, None, None, None, None, 0, 0, 1, 1, 1, 0, None, 0, None, None, None, None,
                                  None, None) == 0
