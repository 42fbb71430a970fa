"""ccmpc_sample_sums_static: the allocation-free variant of ccmpc_sample_sums (scratch = a static device array inside
libccmpc.so; include/ccmpc.h).  No counterpart in the reference.

Called through the C ABI (ctypes) and checked against the existing entry point on two calls in a row that share the
scratch: a large call first, then a smaller one with other data, so stale contents of the array would show.
Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def lib(dev):
    from ccmpc_b200 import _lib
    return _lib.load(build_if_missing=False)


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _case(seed, counts, T, dev, with_yaws):
    """Ragged cells (one of them empty) of trajectories around an ego plan, so that the A9 counts are not all zero."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    C, N = len(counts), int(sum(counts))
    offsets = torch.tensor(np.concatenate([[0], np.cumsum(counts)]), dtype=torch.int64)
    ego = torch.cumsum(torch.full((T, 2), 0.7, dtype=torch.float64), 0) + 100.0
    samples = ego[None] + 3.0 * torch.randn((N, T, 2), generator=g, dtype=torch.float64)
    pivot = ego[None].repeat(C, 1, 1) + 0.25
    yaws = torch.randn((N, T), generator=g, dtype=torch.float64) if with_yaws else None
    to = lambda x: None if x is None else x.contiguous().to(dev)
    return dict(samples=to(samples), offsets=to(offsets), pivot=to(pivot), ego=to(ego), yaws=to(yaws), N=N, C=C, T=T)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _existing(lib, k, R):
    from ccmpc_b200 import _lib
    nbytes = lib.ccmpc_sample_sums_workspace_bytes(k["N"], k["C"], k["T"])
    ws = torch.empty(nbytes, dtype=torch.uint8, device=k["samples"].device)
    out = torch.full((k["C"], lib.ccmpc_sums_len(k["T"])), float("nan"), dtype=torch.float64, device=ws.device)
    rc = lib.ccmpc_sample_sums(_p(k["samples"]), _p(k["offsets"]), k["N"], k["C"], k["T"], _p(k["pivot"]), _p(k["ego"]),
                               0, R, _p(k["yaws"]), _p(out), _p(ws), nbytes, _stream())
    _lib.check(rc, "ccmpc_sample_sums")
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _check(got, want, T):
    from ccmpc_b200 import ops
    L = ops.sums_layout(T)
    assert np.isfinite(got).all()
    np.testing.assert_array_equal(got[:, 0], want[:, 0])                                      # n
    np.testing.assert_array_equal(got[:, L["coll"]:L["pen"]], want[:, L["coll"]:L["pen"]])    # coll[T], any: exact
    np.testing.assert_allclose(got, want, rtol=1e-11, atol=1e-9)                              # the bar of test_gpu_parity


def test_static_scratch_variant_matches_sample_sums_on_two_calls_in_a_row(dev, lib):
    from ccmpc_b200 import _lib
    R = 3.4
    first = _case(1, [5000, 0, 3001, 7, 4096, 2500], 16, dev, with_yaws=True)
    second = _case(2, [130, 1, 999], 8, dev, with_yaws=False)
    want = [_existing(lib, k, R) for k in (first, second)]
    assert want[0][:, 1 + 32 + 528:1 + 32 + 528 + 16].sum() > 0        # the predicate fires somewhere
    outs = [torch.full((k["C"], lib.ccmpc_sums_len(k["T"])), float("nan"), dtype=torch.float64, device=dev)
            for k in (first, second, first)]
    torch.cuda.synchronize()
    free0 = None
    # first, second, first again: the second call starts from the first one's scratch contents and so on
    for i, (k, out) in enumerate(zip((first, second, first), outs)):
        if i == 1:                                 # the module is loaded and the symbol resolved by now
            free0 = torch.cuda.mem_get_info()[0]
        rc = lib.ccmpc_sample_sums_static(_p(k["samples"]), _p(k["offsets"]), k["N"], k["C"], k["T"], _p(k["pivot"]),
                                          _p(k["ego"]), 0, R, _p(k["yaws"]), _p(out), _stream())
        _lib.check(rc, "ccmpc_sample_sums_static")
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0   # nothing was allocated on the device by the later calls
    _check(outs[0].cpu().numpy(), want[0], 16)
    _check(outs[1].cpu().numpy(), want[1], 8)
    _check(outs[2].cpu().numpy(), want[0], 16)


def test_static_scratch_variant_rejects_sizes_beyond_its_array(dev, lib):
    from ccmpc_b200 import _lib
    cap = lib.ccmpc_sample_sums_static_max_bytes()
    assert cap == 4 << 20
    assert lib.ccmpc_sample_sums_workspace_bytes(8_000_000, 32, 16) <= cap        # BASELINE config c3 fits
    C, T = 900, 16
    assert lib.ccmpc_sample_sums_workspace_bytes(1000, C, T) > cap
    d = torch.zeros(16, dtype=torch.float64, device=dev)
    rc = lib.ccmpc_sample_sums_static(_p(d), _p(d), 1000, C, T, _p(d), None, 0, 3.4, None, _p(d), _stream())
    assert rc == -3                                                               # CCMPC_E_WORKSPACE, nothing launched
    assert b"static" in lib.ccmpc_last_error_string() and b"4194304" in lib.ccmpc_last_error_string()
    with pytest.raises(Exception):
        _lib.check(rc, "ccmpc_sample_sums_static")
    torch.cuda.synchronize()
