"""CUDA path (through the C ABI of libccmpc.so) against the goldens produced by the
reference's own functions and against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): counts, best_idx and side bit-exact; matrices,
half-plane coefficients and costs within 1e-5 relative - the kernels are fp64, so the
tests hold them to far tighter tolerances, written next to each check."""
import numpy as np
import pytest
import torch

from _util import relerr

pytestmark = pytest.mark.gpu

RTOL = 1e-5          # the contract
TIGHT = 1e-9         # what fp64 kernels actually deliver on well-conditioned inputs


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from ccmpc_b200 import _lib
    _lib.load(build_if_missing=False)          # the CUDA library must be there: no fallback
    return torch.device("cuda:0")


def sym3(S):
    S = np.asarray(S)
    return np.stack([S[..., 0, 0], S[..., 0, 1], S[..., 1, 1]], axis=-1)


def full(q3):
    q3 = np.asarray(q3)
    out = np.empty(q3.shape[:-1] + (2, 2))
    out[..., 0, 0] = q3[..., 0]; out[..., 0, 1] = q3[..., 1]
    out[..., 1, 0] = q3[..., 1]; out[..., 1, 1] = q3[..., 2]
    return out


def T_(a, dev, dtype=torch.float64):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dtype).to(dev)


# ----------------------------------------------------------------------------- K3
def test_mvoe_fit_matches_reference_golden(dev, golden):
    from ccmpc_b200 import ops
    from oracle import restate
    g = golden("ref_mvoe")
    beta, q, iters, status = ops.mvoe_fit(T_(sym3(g["Sigma1"]), dev), T_(sym3(g["Sigma2"]), dev))
    assert not status.any()
    np.testing.assert_allclose(beta.cpu().numpy(), g["beta"], rtol=1e-7)     # both stop at |delta| < 1e-8
    assert relerr(full(q.cpu().numpy()), g["Q"]) < 1e-8
    for i in range(len(g["beta"])):
        assert np.max(np.abs(full(q.cpu().numpy()[i]) - g["Q"][i])) <= 1e-7 * np.max(np.abs(g["Q"][i]))
    # same iterate sequence => same iteration count as the restated loop
    it_ref = np.array([restate.mvoe(a, b, return_iters=True)[2] for a, b in zip(g["Sigma1"], g["Sigma2"])])
    it = iters.cpu().numpy()
    assert np.mean(it == it_ref) > 0.97 and np.max(np.abs(it - it_ref)) <= 1
    # accelerated start converges to the same fixed point
    beta_a, q_a, it_a, st_a = ops.mvoe_fit(T_(sym3(g["Sigma1"]), dev), T_(sym3(g["Sigma2"]), dev), accel=True)
    assert not st_a.any()
    np.testing.assert_allclose(beta_a.cpu().numpy(), g["beta"], rtol=1e-6)
    assert relerr(full(q_a.cpu().numpy()), g["Q"]) < 1e-7
    assert it_a.float().mean() < iters.float().mean()
    # direct solve (accel = 2): the root of the stationarity cubic, accepted under the reference's stopping rule
    beta_d, q_d, it_d, st_d = ops.mvoe_fit(T_(sym3(g["Sigma1"]), dev), T_(sym3(g["Sigma2"]), dev), accel=2)
    assert not st_d.any()
    np.testing.assert_allclose(beta_d.cpu().numpy(), g["beta"], rtol=1e-7)   # north_star: 1e-5 relative
    assert relerr(full(q_d.cpu().numpy()), g["Q"]) < 1e-7
    assert int(it_d.max()) == 1, "every golden pair is solved without falling back to the iteration"


def test_mvoe_fit_footprint_form_and_edge_cases(dev):
    from ccmpc_b200 import ops
    from oracle import restate
    rng = np.random.default_rng(11)
    from oracle.synth import random_spd, sym_to_full
    s1 = random_spd(rng, (1000,))
    R2 = 3.4 ** 2
    beta, q, _, status = ops.mvoe_fit(T_(s1, dev), None, footprint_r2=R2)
    assert not status.any()
    for i in range(0, 1000, 37):
        b, Q = restate.mvoe(sym_to_full(s1[i]), np.eye(2) * R2)
        assert abs(beta[i].item() - b) < 1e-7
        assert relerr(full(q[i].cpu().numpy()), Q) < 1e-8
    # KATs: radius 1 (+) radius 2 = radius 3
    beta, q, _, _ = ops.mvoe_fit(T_([[1.0, 0.0, 1.0]], dev), T_([[4.0, 0.0, 4.0]], dev))
    assert abs(beta.item() - 0.5) < 1e-8 and relerr(q.cpu().numpy(), [[9.0, 0.0, 9.0]]) < 1e-8
    # empty batch, singular Sigma1, maxiter exhaustion
    b0, q0, _, _ = ops.mvoe_fit(torch.empty((0, 3), dtype=torch.float64, device=dev), None)
    assert b0.numel() == 0 and q0.shape == (0, 3)
    _, _, _, st = ops.mvoe_fit(T_([[1.0, 1.0, 1.0]], dev), T_([[1.0, 0.0, 1.0]], dev))
    assert st.item() == ops.ST_SINGULAR
    _, _, it, st = ops.mvoe_fit(T_([[1.0, 0.2, 2.0]], dev), T_([[30.0, 0.0, 0.5]], dev), maxiter=2)
    assert st.item() == ops.ST_NOT_CONVERGED and it.item() == 2


@pytest.mark.parametrize("accel", [0, 1, 2])
def test_mvoe_halfplane_sweep_matches_oracle(dev, accel):
    """c4 kernel (v8ideal:915-939 per cell) against the restated compute_mvoe + choose_closest_tangent, on a small
    (S, O, T, K) block that is NOT a multiple of the 256-fit tile, with eigenvalue ratios up to 1e8 (drives the
    direct solve into its fall-back), a singular covariance and a cell whose slope is undefined."""
    from ccmpc_b200 import ops
    from oracle import restate, synth
    S, O, T, K = 3, 2, 5, 7
    rng = np.random.default_rng(40 + accel)
    cov = synth.random_spd(rng, (S, O, T, K))
    wide = synth.random_spd(rng, (O, T, K), smin=1e-3, smax=30.0)
    cov[1] = wide
    cov[1, :, :, 0, 0] *= 1e4                       # sigma_x >> sigma_y: lam ratio up to ~1e8
    cov[1, :, :, 0, 1] *= 1e2
    mu = np.stack([rng.uniform(150, 200, (S, O, T, K)), rng.uniform(-100, -60, (S, O, T, K))], -1)
    ref = np.broadcast_to(synth.reference_trajectory(T), (S, T, 2)).copy()
    ref[..., 1] += 30.0
    cov[2, 1, 4, 6] = [1.0, 1.0, 1.0]               # singular Sigma1
    mu[2, 0, 3, 2, 1] = ref[2, 3, 1]                # ref_y == mu_y: slope -dx/0 (v8ideal:923)
    chi = -2.0 * np.log(0.05 / O / T)
    R2 = 3.4 ** 2
    q, nd, idx_side, status = ops.mvoe_halfplane_sweep(T_(cov, dev), T_(mu, dev), T_(ref, dev), chi, accel=accel)
    q, nd, idx_side, status = (a.cpu().numpy() for a in (q, nd, idx_side, status))
    assert status[2, 1, 4, 6] == ops.ST_SINGULAR and status[2, 0, 3, 2] != 0
    status[2, 1, 4, 6] = status[2, 0, 3, 2] = 0
    assert not status.any()
    for s, o, t, k in np.ndindex(S, O, T, K):
        if (s, o, t, k) in ((2, 1, 4, 6), (2, 0, 3, 2)):
            continue
        b_ref, Q = restate.mvoe(synth.sym_to_full(cov[s, o, t, k]) * chi, np.eye(2) * R2)
        # the reference stops at an ABSOLUTE |delta beta| < 1e-8, so Q = (1 + 1/beta) Sigma1 + ... is only
        # defined to ~1e-8 / beta relative by its own rule; the reference sequence (accel = 0) lands on the same iterate
        assert relerr(full(q[s, o, t, k]), Q) < (1e-7 if accel == 0 else max(1e-7, 3e-8 / min(b_ref, 1.0) ** 2))
        m = -(ref[s, t, 0] - mu[s, o, t, k, 0]) / (ref[s, t, 1] - mu[s, o, t, k, 1])
        n, d, idx = restate.closest_tangent(mu[s, o, t, k], Q, 1, m, ref[s, t])
        assert np.allclose(nd[s, o, t, k, :2], n, rtol=1e-12) and abs(nd[s, o, t, k, 2] - d) <= 1e-7 * abs(d)
        assert (idx_side[s, o, t, k] & 1) == idx                                       # bit-exact best_idx
        assert bool(idx_side[s, o, t, k] & 2) == (not (n @ mu[s, o, t, k] <= d))       # side (v8ideal:926)


# ----------------------------------------------------------------------------- K4
def test_tangent_matches_reference_golden(dev, golden):
    from ccmpc_b200 import ops
    g = golden("ref_tangent")
    args = [T_(g["mu"], dev), T_(sym3(g["Sigma"]), dev), T_(g["c"], dev), T_(g["m"], dev), T_(g["a"], dev)]
    d12, dist, d_star, idx = ops.tangent(*args)
    np.testing.assert_array_equal(idx.cpu().numpy(), g["idx"])                # bit-exact index
    np.testing.assert_allclose(d12.cpu().numpy()[:, 0], g["d1"], rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(d12.cpu().numpy()[:, 1], g["d2"], rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(dist.cpu().numpy(), g["dist"], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(d_star.cpu().numpy(), g["d_star"], rtol=1e-12, atol=1e-12)
    for j in (0, 1):
        ci = torch.full((len(g["c"]),), j, dtype=torch.int32, device=dev)
        _, _, ds, ix = ops.tangent(*args, const_idx=ci)
        np.testing.assert_allclose(ds.cpu().numpy(), g["d_forced"][:, j], rtol=1e-12, atol=1e-12)
        assert (ix.cpu().numpy() == j).all()
    # no tangent: n^T Sigma n <= 0
    _, _, _, ix = ops.tangent(T_([[0.0, 0.0]], dev), T_([[0.0, 0.0, 0.0]], dev), T_([1.0], dev), T_([0.3], dev),
                              T_([[1.0, 1.0]], dev))
    assert ix.item() == -1


def test_scale_lower_matches_reference_golden(dev, golden):
    from ccmpc_b200 import ops
    g = golden("ref_scale_lower")
    scale, lower = ops.scale_lower(T_(sym3(g["cov_infer"]), dev), T_(sym3(g["cov_mu"]), dev), T_(sym3(g["cov_t"]), dev),
                                   gamma=T_(g["Gamma"], dev), eps=T_(g["eps"], dev))
    np.testing.assert_allclose(scale.cpu().numpy(), g["scale"], rtol=1e-11)
    np.testing.assert_allclose(lower.cpu().numpy(), g["lower"], rtol=1e-9, atol=1e-14)


# ------------------------------------------------------ scalar drop-in twins
def test_makeconstraint_scalar_api(dev, golden):
    from ccmpc_b200 import makeconstraint as mc
    b, Q = mc.compute_mvoe(np.array([[2, .5], [.5, 1]]), np.array([[1, -.2], [-.2, 3]]))
    assert b == pytest.approx(0.8915954797056089, rel=1e-7)
    np.testing.assert_allclose(Q, [[6.134765219261006, 0.6824733389477273], [0.6824733389477273, 7.796371308894525]],
                               rtol=1e-7)
    with pytest.raises(np.linalg.LinAlgError):
        mc.compute_mvoe(np.zeros((2, 2)), np.eye(2))
    n, d, idx = mc.choose_closest_tangent([1, 2], [[2, .5], [.5, 1]], 1, 0.5, [10, 3])
    np.testing.assert_array_equal(n, [-0.5, 1.0])
    assert d == pytest.approx(0.5) and idx == 1
    n, d, idx = mc.choose_closest_tangent([1, 2], [[2, .5], [.5, 1]], 1, 0.5, [10, 30])
    assert d == pytest.approx(2.5) and idx == 0
    assert mc.choose_closest_tangent([0, 0], np.zeros((2, 2)), 1.0, 0.3, [1, 1]) == (None, None, None, None)
    assert mc.tangent_lines_of_slope_m([0, 0], np.zeros((2, 2)), 1.0, 0.3) == []
    (n1, d1), (n2, d2) = mc.tangent_lines_of_slope_m([1, 2], [[2, .5], [.5, 1]], 1, 0.5)
    assert d1 == pytest.approx(2.5) and d2 == pytest.approx(0.5)
    assert mc.distance_point_to_line(np.array([-0.5, 1.0]), 0.5, [10, 3]) == pytest.approx(
        abs(-5 + 3 - 0.5) / np.sqrt(1.25), rel=1e-12)
    ci, cm = np.array([[.5, .1], [.1, .4]]), np.array([[.3, .05], [.05, .2]])
    assert mc.compute_lower_bound(ci, cm, ci + cm, 0.05 / 8) == pytest.approx(0.29104264152591025, rel=1e-9)
    assert mc.compute_scale(ci, cm, ci + cm, 2.497705474412374) == pytest.approx(3.3529447079923145, rel=1e-11)
    g = golden("ref_predict_moments")
    for i in range(0, 64, 7):
        a, b_, c = mc.predict_moments(list(g["rows"][i]))
        assert relerr(a, g["cov_infer"][i]) < TIGHT and relerr(b_, g["cov_mu"][i]) < TIGHT
        assert relerr(c, g["cov_t"][i]) < TIGHT
    with pytest.raises(np.linalg.LinAlgError):
        x = np.arange(10.0)
        mc.predict_moments([x, x * 2, np.ones(10), np.ones(10)])       # Sigma_tau = 0


# ----------------------------------------------------- per-step pipeline goldens
def _pipeline(dev, g):
    from ccmpc_b200 import ops
    samples, offsets = T_(g["samples"], dev), T_(g["offsets"], dev, torch.int64)
    pivot = ops.cell_pivot(samples, offsets)
    sums = ops.sample_sums(samples, offsets, pivot)
    eps = T_(g["eps_ura"][g["cell_obstacle"], g["cell_mode"]], dev)
    return samples, offsets, pivot, sums, eps


@pytest.mark.parametrize("accel", [0, 2])
def test_minkowski_halfplanes_match_reference_golden(dev, golden, accel):
    """ref_pipeline.npz was produced by the reference's predict_moments / compute_mvoe /
    choose_closest_tangent / compute_lower_bound inside the restated loop of v8ideal:881-947.
    accel = 2: direct MVOE solve + MUFU/Newton reciprocals, same tolerances."""
    from ccmpc_b200 import ops
    g = golden("ref_pipeline")
    T = int(g["T"])
    samples, offsets, pivot, sums, eps = _pipeline(dev, g)
    out = ops.minkowski_halfplanes(sums, pivot, T_(g["ref_traj"], dev), eps, prediction_horizon=T, accel=accel)
    h = {k: v.cpu().numpy() for k, v in out.items()}
    assert not h["status"].any()
    np.testing.assert_array_equal(h["idx_side"] & 1, g["mk_idx"])                          # bit-exact
    np.testing.assert_array_equal(np.where(h["idx_side"] & 2, -1, 1), g["mk_side"])        # bit-exact
    assert relerr(h["nd"][..., :2], g["mk_n"]) < TIGHT and relerr(h["nd"][..., 2], g["mk_d"]) < TIGHT
    assert relerr(h["ab"][..., :2], g["mk_A"]) < TIGHT and relerr(h["ab"][..., 2], g["mk_b"]) < TIGHT
    for i in np.ndindex(g["mk_d"].shape):                                                   # per-matrix 1e-5 bar
        assert relerr(full(h["q"][i]), g["mk_Q"][i]) < 1e-7
        assert relerr(full(h["q1"][i]), g["mk_Q1"][i]) < 1e-7
        assert relerr(full(h["cov_infer"][i]), g["mk_cov_infer"][i]) < TIGHT
        assert relerr(full(h["cov_mu"][i]), g["mk_cov_mu"][i]) < TIGHT
    np.testing.assert_allclose(h["betas"][..., 0], g["mk_beta1"], rtol=1e-6)
    np.testing.assert_allclose(h["betas"][..., 1], g["mk_beta2"], rtol=1e-6)
    np.testing.assert_allclose(h["prob_pair"], g["mk_prob_pair"], rtol=1e-7, atol=1e-12)
    np.testing.assert_allclose(h["prob_lower"], g["mk_prob_lower"], rtol=1e-7, atol=1e-12)
    assert relerr(h["mean"], g["mk_mean"]) < 1e-13


def test_affine_halfplanes_match_reference_golden(dev, golden):
    from ccmpc_b200 import ops
    g = golden("ref_pipeline")
    T = int(g["T"])
    samples, offsets, pivot, sums, eps = _pipeline(dev, g)
    out = ops.affine_halfplanes(sums, pivot, T_(g["ref_traj"], dev), eps, Tpred=T)
    h = {k: v.cpu().numpy() for k, v in out.items()}
    assert not h["status"].any()
    np.testing.assert_array_equal(h["idx_side"] & 1, g["af_idx"])
    np.testing.assert_array_equal(np.where(h["idx_side"] & 2, -1, 1), g["af_side"])
    assert relerr(h["nd"][..., :2], g["af_n"]) < TIGHT and relerr(h["nd"][..., 2], g["af_d"]) < TIGHT
    np.testing.assert_allclose(h["offset_rhs"][..., 0], g["af_offset"], rtol=1e-8)
    assert relerr(h["offset_rhs"][..., 1], g["af_rhs"]) < TIGHT
    assert relerr(h["mean"], g["af_mean"]) < 1e-13
    assert relerr(full(h["cov"]), g["af_cov"]) < TIGHT


# --------------------------------------------------------- K2 + K6 sums / counts
def _ragged_cells(rng, T, sizes, spread=3.0, centre=(170.0, -80.0)):
    chunks, offs = [], [0]
    for n in sizes:
        mu = np.asarray(centre) + rng.normal(size=(T, 2)) * 4
        L = rng.normal(size=(T, 2, 2)) * spread / 3
        x = mu + np.einsum("tij,ntj->nti", L, rng.normal(size=(n, T, 2)))
        chunks.append((x - 100).astype(np.float32).astype(np.float64) + 100)
        offs.append(offs[-1] + n)
    flat = np.concatenate(chunks) if chunks else np.zeros((0, T, 2))
    return flat, np.asarray(offs, np.int64)


@pytest.mark.parametrize("T,sizes", [(8, [200, 1, 0, 5000, 37, 4097]), (16, [1800, 1792, 1793, 3, 20000]),
                                     (1, [10, 300]), (3, [2, 999, 0, 0, 64]), (5, [12345]), (7, [0, 0, 17])])
def test_sample_sums_counts_bit_exact_and_moments(dev, T, sizes):
    from ccmpc_b200 import ops
    from oracle import restate
    rng = np.random.default_rng(100 + T)
    flat, offs = _ragged_cells(rng, T, sizes)
    ego = np.asarray([170.0, -80.0]) + rng.normal(size=(T, 2)) * 2
    yaws = rng.uniform(-np.pi, np.pi, size=flat.shape[:2])
    samples, offsets = T_(flat, dev), T_(offs, dev, torch.int64)
    pivot = ops.cell_pivot(samples, offsets)
    sums = ops.sample_sums(samples, offsets, pivot, ego=T_(ego, dev), yaws=T_(yaws, dev))
    mom = ops.moments_from_sums(sums, pivot)
    L = ops.sums_layout(T)
    assert sums.shape[1] == L["len"] == ops.sums_len(T)
    s = sums.cpu().numpy()
    counts = mom["counts"].cpu().numpy()
    for c, n in enumerate(sizes):
        pos = flat[offs[c]:offs[c + 1]]
        assert s[c, 0] == n
        coll, any_, pen, d2s = restate.collision_counts(pos, ego)
        np.testing.assert_array_equal(counts[c, :T], coll)                     # bit-exact integer counts
        assert counts[c, T] == any_
        np.testing.assert_array_equal(s[c, L["coll"]:L["coll"] + T], coll.astype(np.float64))
        np.testing.assert_allclose(s[c, L["pen"]:L["pen"] + T], pen, rtol=1e-12, atol=1e-9)
        np.testing.assert_allclose(s[c, L["d2s"]:L["d2s"] + T], d2s, rtol=1e-12)
        if n >= 2:
            mean, cov, cross = restate.cell_moments(pos)
            assert relerr(mom["mean"][c].cpu().numpy(), mean) < 1e-13
            for t in range(T):
                tol = TIGHT if n > 2 * T else 1e-6
                assert relerr(mom["cov"][c, t].cpu().numpy(), cov[t]) < tol
                for tau in range(t):
                    assert np.max(np.abs(mom["cross"][c, t, tau].cpu().numpy() - cross[t, tau])) < tol * max(
                        np.max(np.abs(cov[t])), np.max(np.abs(cov[tau])))
            yw = yaws[offs[c]:offs[c + 1]]
            np.testing.assert_allclose(mom["yaw_mean"][c].cpu().numpy(), yw.mean(axis=0), rtol=1e-12, atol=1e-13)
            np.testing.assert_allclose(mom["yaw_var"][c].cpu().numpy(), yw.var(axis=0, ddof=1), rtol=1e-9)


def test_sample_sums_per_cell_ego_and_linearity_over_shards(dev):
    """Counts and power sums of the whole = sum over sample shards with a common pivot
    (what the multi-GPU all-reduce relies on)."""
    from ccmpc_b200 import ops
    rng = np.random.default_rng(7)
    T = 8
    flat, offs = _ragged_cells(rng, T, [3000, 5000])
    ego = np.asarray([170.0, -80.0]) + rng.normal(size=(2, T, 2)) * 2
    pivot = T_(np.stack([flat[0], flat[3000]]), dev)
    whole = ops.sample_sums(T_(flat, dev), T_(offs, dev, torch.int64), pivot, ego=T_(ego, dev))
    parts = []
    for lo, hi in ((0, 0.3), (0.3, 1.0)):
        sel = [np.arange(offs[c] + int(lo * (offs[c + 1] - offs[c])), offs[c] + int(hi * (offs[c + 1] - offs[c])))
               for c in range(2)]
        sub = np.concatenate([flat[s] for s in sel])
        so = np.asarray([0, len(sel[0]), len(sel[0]) + len(sel[1])], np.int64)
        parts.append(ops.sample_sums(T_(sub, dev), T_(so, dev, torch.int64), pivot, ego=T_(ego, dev)))
    tot = (parts[0] + parts[1]).cpu().numpy()
    w = whole.cpu().numpy()
    L = ops.sums_layout(T)
    np.testing.assert_array_equal(tot[:, L["coll"]:L["any"] + 1], w[:, L["coll"]:L["any"] + 1])   # exact integers
    assert tot[:, 0].tolist() == [3000, 5000]
    np.testing.assert_allclose(tot, w, rtol=1e-11, atol=1e-7)


def test_violation_counts_bit_exact(dev, golden):
    from ccmpc_b200 import ops
    from oracle import restate
    g = golden("ref_pipeline")
    T = int(g["T"])
    samples, offsets, pivot, sums, eps = _pipeline(dev, g)
    out = ops.minkowski_halfplanes(sums, pivot, T_(g["ref_traj"], dev), eps, prediction_horizon=T)
    viol = ops.violation_counts(samples, offsets, out["nd"], out["idx_side"]).cpu().numpy()
    nd = out["nd"].cpu().numpy()
    side = np.where(out["idx_side"].cpu().numpy() & 2, -1, 1)
    t_of = [t for t in range(T) for _ in range(t)]
    for c in range(len(g["offsets"]) - 1):
        pos = g["samples"][g["offsets"][c]:g["offsets"][c + 1]]
        ref = restate.violation_counts(pos, nd[c, :, :2], nd[c, :, 2], side[c], t_of)
        np.testing.assert_array_equal(viol[c], ref)
    # shifting every line by +-6 sigma flips the count between 0 and N_k: a sanity check the
    # predicate is not vacuous
    assert viol.max() <= np.diff(g["offsets"]).max()


# ------------------------------------------------------------------------- K5
def test_vertices_match_reference_golden(dev, golden):
    from ccmpc_b200 import npu
    g = golden("ref_vertices")
    v = npu.vertices_of_bboxes(g["centers"], g["theta"], g["lw"])
    np.testing.assert_allclose(v, g["vertices"], rtol=0, atol=2e-13 * 200)       # coordinates ~ 2e2
    vs = npu.vertices_of_bboxes(g["centers"][:7], 0.3, [4.5, 2.5])      # as oracle/make_golden.py:114
    np.testing.assert_allclose(vs, g["vertices_scalar_theta"], rtol=0, atol=4e-11)
    per_box = npu.vertices_of_bboxes(g["centers"][:4], g["theta"][:4], np.array([[3.7, 1.79], [4.0, 2.0]] * 2))
    np.testing.assert_allclose(per_box[0], g["vertices"][0], atol=4e-11)


@pytest.mark.parametrize("T,sizes", [(8, [200, 1, 5000, 0, 4097]), (16, [33, 9000])])
def test_l4_outer_matches_oracle(dev, T, sizes):
    from ccmpc_b200 import ops, npu
    from oracle import restate
    rng = np.random.default_rng(5)
    flat, offs = _ragged_cells(rng, T, sizes)
    yaws = rng.uniform(-0.3, 0.3, size=flat.shape[:2]) + rng.uniform(-np.pi, np.pi)
    samples, offsets, yw = T_(flat, dev), T_(offs, dev, torch.int64), T_(yaws, dev)
    pivot = ops.cell_pivot(samples, offsets)
    sums = ops.sample_sums(samples, offsets, pivot, yaws=yw)
    mom = ops.moments_from_sums(sums, pivot, want_cross=False)
    At, b0 = ops.l4_outer(samples, yw, offsets, mom["yaw_mean"], lw=restate.BBOX_DEFAULT)
    At, b0 = At.cpu().numpy(), b0.cpu().numpy()
    for c, n in enumerate(sizes):
        if n == 0:
            assert np.isnan(b0[c]).all()
            continue
        A_ref, b_ref, th = restate.l4_cell(flat[offs[c]:offs[c + 1]], yaws[offs[c]:offs[c + 1]])
        assert relerr(At[c], A_ref) < 1e-11
        assert relerr(b0[c], b_ref) < 1e-11
    # scalar twin on explicit corners
    c = 0
    verts = restate.box_vertices(flat[:sizes[0], 2], yaws[:sizes[0], 2], np.asarray(restate.BBOX_DEFAULT))
    A1, b1 = npu.compute_L4_outerapproximation(0.37, verts)
    A2, b2 = restate.l4_outer(0.37, verts)
    assert relerr(A1, A2) < 1e-14 and relerr(b1, b2) < 1e-13


# ------------------------------------------------------------------------- K1
PHILOX_KAT = [  # Random123 kat_vectors, philox4x32-10
    ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


def test_philox_known_answers(dev):
    from ccmpc_b200 import ops
    ctr = torch.tensor(np.array([k[0] for k in PHILOX_KAT], dtype=np.uint32).view(np.int32), device=dev)
    key = torch.tensor(np.array([k[1] for k in PHILOX_KAT], dtype=np.uint32).view(np.int32), device=dev)
    out = ops.philox4x32(ctr, key).cpu().numpy().view(np.uint32)
    np.testing.assert_array_equal(out, np.array([k[2] for k in PHILOX_KAT], dtype=np.uint32))


def _scene_moments(seed, O, K, T):
    from oracle import synth
    sc = synth.make_scene(seed, O, K, T)
    mean, cov, cross = sc.moments()
    C = O * K
    return sc, mean.reshape(C, T, 2), cov.reshape(C, T, 2, 2), cross.reshape(C, T, T, 2, 2)


def test_rollout_with_supplied_noise_matches_oracle(dev):
    from ccmpc_b200 import ops
    from oracle import restate
    sc, mean, cov, cross = _scene_moments(3, 2, 3, 9)
    C, Tsh, n = 6, 8, 777
    rng = np.random.default_rng(0)
    Z = rng.standard_normal((C, n, Tsh, 2))
    x0 = mean[:, 0] + rng.normal(size=(C, 2)) * 0.1
    traj, _, status, x0_used = ops.mc_rollout(T_(mean, dev), T_(cov, dev), T_(cross, dev), T_(x0, dev), n, Tsh,
                                              z=T_(Z, dev))
    assert not status.any()
    np.testing.assert_array_equal(x0_used.cpu().numpy(), x0)
    traj = traj.cpu().numpy()
    for c in range(C):
        ref = restate.ideal_rollout(mean[c], cov[c], cross[c], Tsh, Z[c], x0[c])
        assert relerr(traj[c], ref) < 1e-13


def test_rollout_philox_statistics_and_shard_invariance(dev):
    """Sampling parity is statistical (the reference draws from MT19937): the rollout's sample
    moments must match the conditional-Gaussian chain's analytic moments within CLT bounds, and
    the stream must not depend on how samples are sharded."""
    from ccmpc_b200 import ops
    sc, mean, cov, cross = _scene_moments(3, 1, 2, 9)
    C, Tsh, n = 2, 8, 400_000
    args = (T_(mean, dev), T_(cov, dev), T_(cross, dev))
    traj, z, status, x0 = ops.mc_rollout(*args, None, n, Tsh, seed=42, want_z=True)
    assert not status.any()
    z = z.cpu().numpy()
    assert abs(z.mean()) < 5 / np.sqrt(z.size) and abs(z.var() - 1) < 5 * np.sqrt(2 / z.size)
    assert abs(np.mean(z ** 4) - 3) < 0.05 and np.abs(z).max() < 7
    assert abs(np.corrcoef(z[0, :, 0, 0], z[0, :, 1, 0])[0, 1]) < 0.01
    # same seed, two shards with sample_offset == one launch
    a, _, _, _ = ops.mc_rollout(*args, x0, n // 2, Tsh, seed=42)
    b, _, _, _ = ops.mc_rollout(*args, x0, n - n // 2, Tsh, seed=42, sample_offset=n // 2)
    assert torch.equal(torch.cat([a, b], dim=1), traj)
    # chain moments: x_{t+1} | x_t Gaussian with the reference's A_t, L_t starting from the point x0
    x0 = x0.cpu().numpy()
    tr = traj.cpu().numpy()
    from oracle import restate
    for c in range(C):
        A, L = restate.rollout_params(mean[c], cov[c], cross[c], Tsh)
        m, S = x0[c].copy(), np.zeros((2, 2))
        for t in range(Tsh):
            m = mean[c, t + 1] + A[t] @ (m - mean[c, t])
            S = A[t] @ S @ A[t].T + L[t] @ L[t].T
            xs = tr[c, :, t]
            se = np.sqrt(np.diag(S) / n)
            assert np.all(np.abs(xs.mean(axis=0) - m) < 5 * se + 1e-12)
            assert relerr(np.cov(xs.T), S) < 0.02
    # non-PD conditional covariance is reported, not silently sampled
    bad = cov.copy(); bad[0, 3] = -bad[0, 3]
    _, _, st, _ = ops.mc_rollout(T_(mean, dev), T_(bad, dev), T_(cross, dev), None, 16, Tsh)
    assert st[0].item() in (ops.ST_SINGULAR, ops.ST_NO_TANGENT)


# ------------------------------------------------------------------------ A12
def test_episode_cost_matches_oracle(dev):
    from ccmpc_b200 import cost
    from oracle import restate
    rng = np.random.default_rng(9)
    for H in (2, 3, 14, 50):
        X = [np.array([160 + rng.normal(), -80 + rng.normal(), 0.1, 5.0]) for _ in range(H)]
        U = [rng.normal(size=2) for _ in range(H)]
        c, cu = cost.cost_final(X, U)
        rc, rcu = restate.episode_cost(X, U)
        assert c == pytest.approx(rc, rel=1e-13) and cu == pytest.approx(rcu, rel=1e-13)
    with pytest.raises(UnboundLocalError):
        cost.cost_final([np.zeros(4)], [np.zeros(2)])


# --------------------------------------------------------- one MPC step as a CUDA graph
@pytest.mark.parametrize("builder", ["minkowski", "affine", "both"])
def test_step_plan_graph_matches_oracle_and_eager(dev, builder):
    """StepPlan (H2D -> yaws -> sums/counts -> moments -> half-planes -> L4 -> D2H in one CUDA graph) against
    the oracle loops (v8ideal:881-947 / :1476-1515 / :694-736) and against its own eager launch sequence,
    replayed on two different sample sets (the graph must not bake inputs in)."""
    from ccmpc_b200 import midlevel
    from oracle import restate, synth
    O, K, T, N = 2, 3, 8, 700
    sc = synth.make_scene(21, O, K, T)
    z, pos = synth.sample_scene(sc, N, 21)
    plan = midlevel.StepPlan.from_samples(z, pos, sc.pos_last, K, prediction_horizon=T, device=dev, builder=builder)
    eager = midlevel.StepPlan.from_samples(z, pos, sc.pos_last, K, prediction_horizon=T, device=dev,
                                           builder=builder, use_graph=False)
    eps = restate.eps_allocation(plan.n_states)
    for rep in range(2):
        hs = plan.host_samples.numpy().copy()
        if rep == 1:                                  # second replay: shifted samples, shifted reference
            hs = (hs + np.asarray([0.37, -0.21])).copy()
        ref_traj = sc.ref_traj + (0.0 if rep == 0 else 0.15)
        res = {k: v.copy() for k, v in plan.run(pos_host=hs, ref_traj=ref_traj).items()}
        res_e = eager.run(pos_host=hs, ref_traj=ref_traj)
        for k in res:
            np.testing.assert_array_equal(res[k], res_e[k], err_msg=f"graph != eager for {k}")
        for c in range(plan.C):
            cell = hs[plan.offsets_host[c]:plan.offsets_host[c + 1]]
            e = eps[plan.cell_obstacle[c], plan.cell_mode[c]]
            if builder in ("minkowski", "both"):
                ref = restate.minkowski_cell(cell, ref_traj, e, T)
                np.testing.assert_array_equal(res["idx_side"][c] & 1, ref["idx"])
                np.testing.assert_array_equal(np.where(res["idx_side"][c] & 2, -1, 1), ref["side"])
                np.testing.assert_allclose(res["nd"][c][:, 2], ref["d"], rtol=1e-7)
                np.testing.assert_allclose(res["ab"][c][:, :2], ref["A"], rtol=1e-7)
                viol = restate.violation_counts(cell, res["nd"][c][:, :2], res["nd"][c][:, 2],
                                                np.where(res["idx_side"][c] & 2, -1, 1),
                                                [t for t in range(T) for _ in range(t)])
                np.testing.assert_array_equal(res["viol"][c], viol)
            if builder in ("affine", "both"):
                ref = restate.affine_cell(cell, ref_traj, e, T)
                np.testing.assert_array_equal(res["aff_idx_side"][c] & 1, ref["idx"])
                np.testing.assert_array_equal(np.where(res["aff_idx_side"][c] & 2, -1, 1), ref["side"])
                np.testing.assert_allclose(res["aff_nd"][c][:, 2], ref["d"], rtol=1e-7)
                np.testing.assert_allclose(res["aff_offset_rhs"][c][:, 1], ref["rhs"], rtol=1e-7)
            coll, any_, _, _ = restate.collision_counts(cell, ref_traj)
            np.testing.assert_array_equal(res["counts"][c][:T], coll)
            assert res["counts"][c][T] == any_
            yaws = restate.yaws_from_positions(cell, sc.pos_last[plan.cell_obstacle[c]])
            At, b0, _ = restate.l4_cell(cell, yaws)
            np.testing.assert_allclose(res["At"][c], At, atol=1e-10)
            np.testing.assert_allclose(res["b0"][c], b0, rtol=1e-10)


# --------------------------------------------------------- A2: mode grouping on the device
def _group_inputs(rng, V, N, T, S, sparse):
    """Per-vehicle latent probabilities with a few dominant modes (ovehicle.py:58 keeps pmf > 0.1)."""
    pmf = np.zeros((V, S))
    z = np.zeros((V, N), np.int32)
    for v in range(V):
        w = rng.dirichlet(np.full(S, 0.08 if sparse else 1.0))
        pmf[v] = w
        z[v] = rng.choice(S, size=N, p=w)
    centres = rng.uniform(-30, 30, size=(V, S, 1, 2))
    step = rng.normal(size=(V, S, 1, 2)) * 2
    pred = np.zeros((V, N, T, 2), np.float32)
    for v in range(V):
        base = centres[v][z[v]] + np.arange(1, T + 1)[None, :, None] * step[v][z[v]]
        pred[v] = (base + rng.normal(size=(N, T, 2)) * 0.7).astype(np.float32)
    return pred, z, pmf


@pytest.mark.parametrize("V,N,T,S,sparse,f32", [(3, 5000, 8, 25, True, True), (2, 1237, 16, 25, True, False),
                                                (1, 40, 3, 7, False, True), (2, 3000, 8, 25, False, True),
                                                (2, 5, 4, 6, False, True), (1, 20001, 16, 25, True, True)])
def test_group_by_mode_matches_oracle(dev, V, N, T, S, sparse, f32):
    """ccmpc_group_by_mode against the restated bucket + fold of v8ideal:486-494 / ovehicle.py:55-113:
    identical row order (perm), cell sizes, retained latents; centres and pmf to rounding."""
    from ccmpc_b200 import ops
    from oracle import restate
    rng = np.random.default_rng(1000 + N)
    pred, z, pmf = _group_inputs(rng, V, N, T, S, sparse)
    minpos = np.asarray([151.25, -93.5])
    pred64 = pred.astype(np.float64) + minpos                    # v8ideal:486
    src = T_(pred, dev, torch.float32) if f32 else T_(pred64, dev)
    out = ops.group_by_mode(src, T_(z, dev, torch.int32), T_(pmf, dev), minpos=minpos if f32 else (0, 0))
    h = {k: v.cpu().numpy() for k, v in out.items()}
    for v in range(V):
        keep = np.argwhere(pmf[v] > 0.1).ravel()
        assert h["n_states"][v] == keep.size
        if keep.size == 0:
            assert h["status"][v] == 2
            continue
        pos_l, yaw_l, centers, pmf_o, mem_l = restate.group_by_mode(pred64[v], z[v], pmf[v], np.zeros(2), T)
        for j in range(S):
            c = v * S + j
            b, e = h["offsets"][c], h["offsets"][c + 1]
            if j >= keep.size:
                assert b == e and h["cell_latent"][c] == -1
                continue
            assert h["cell_latent"][c] == keep[j]
            np.testing.assert_array_equal(h["perm"][v][b - v * N:e - v * N], mem_l[j])       # same row order
            np.testing.assert_array_equal(h["samples"][b:e], pos_l[j])                         # bit-exact rows
            np.testing.assert_allclose(h["init_center"][c], centers[j], rtol=1e-13)
            np.testing.assert_allclose(h["cell_pmf"][c], pmf_o[j], rtol=1e-15)
    assert h["offsets"][-1] == V * N


def test_step_plan_ingest_matches_host_grouped_plan(dev):
    """StepPlan.from_predictions (float32 predictions + z + latent pmf in, grouping on the device inside the
    graph) gives, for every retained (vehicle, mode), exactly the rows of a plan fed with the oracle's
    host-side grouping (v8ideal:486-494 + ovehicle.py:55-113)."""
    from ccmpc_b200 import midlevel
    from oracle import restate, synth
    rng = np.random.default_rng(77)
    V, N, T, S = 3, 2500, 8, 25
    pred, z, pmf = _group_inputs(rng, V, N, T, S, sparse=True)
    minpos = np.asarray([160.0, -90.0])
    pos_last = minpos + rng.normal(size=(V, 2)) * 3
    ref_traj = synth.reference_trajectory(T)
    plan = midlevel.StepPlan.from_predictions(V, N, T, pos_last, prediction_horizon=T, S=S, minpos=minpos,
                                              device=dev, builder="both")
    for rep in range(2):                                  # replay the graph on fresh inputs
        if rep == 1:
            pred, z, pmf = _group_inputs(rng, V, N, T, S, sparse=True)
        res = plan.run_predictions(pred, z, pmf, ref_traj=ref_traj)
        pred64 = pred.astype(np.float64) + minpos
        for v in range(V):
            keep = np.argwhere(pmf[v] > 0.1).ravel()
            assert res["g_n_states"][v] == keep.size
            if keep.size == 0:
                continue
            pos_l, _, _, pmf_o, _ = restate.group_by_mode(pred64[v], z[v], pmf[v], pos_last[v], T)
            offs = np.concatenate([[0], np.cumsum([p.shape[0] for p in pos_l])])
            href = midlevel.StepPlan(offs, [0] * keep.size, list(range(keep.size)), [keep.size],
                                     np.tile(pos_last[v], (keep.size, 1)), T, T, device=dev, builder="both",
                                     eps_ura=np.full((1, keep.size), 0.05 / V), use_graph=False)
            want = href.run(pos_host=np.concatenate(pos_l), ref_traj=ref_traj)
            for j in range(keep.size):
                c = v * S + j
                assert res["g_cell_latent"][c] == keep[j]
                np.testing.assert_allclose(res["g_cell_pmf"][c], pmf_o[j], rtol=1e-15)
                for k in ("nd", "ab", "idx_side", "status", "prob_lower", "mean", "cov", "At", "b0", "yaw_mean",
                          "counts", "viol", "aff_nd", "aff_offset_rhs", "aff_idx_side"):
                    np.testing.assert_array_equal(res[k][c], want[k][j], err_msg=f"{k} differs (v={v}, j={j})")


# --------------------------------------------------------- stage 1: GMM2D sampling on the device
@pytest.mark.parametrize("V,N,Z,T,G", [(3, 4001, 25, 8, 1), (2, 2500, 6, 7, 3), (1, 1, 1, 1, 1)])
def test_gmm2d_sample_supplied_noise_matches_oracle(dev, V, N, Z, T, G):
    """Deterministic mode: same uniforms / normals in, same latents and (float32) positions out as the oracle's
    restatement of the published GMM2D.rsample (prediction.py:81-86; submodule absent => parity unpinned)."""
    from ccmpc_b200 import ops
    from oracle import restate
    rng = np.random.default_rng(100 + G)
    pmf = rng.dirichlet(np.full(Z, 0.3), size=V)
    if Z > 1:
        pmf[:, Z // 2] = 0.0                                       # an impossible latent is never drawn
    mus = rng.uniform(-50, 50, (V, Z, T, G, 2)).astype(np.float32)
    ls = rng.uniform(-2, 1, (V, Z, T, G, 2)).astype(np.float32)
    rho = rng.uniform(-0.95, 0.95, (V, Z, T, G)).astype(np.float32)
    lp = rng.normal(size=(V, Z, T, G)).astype(np.float32) if G > 1 else None
    u_z = rng.random((V, N), dtype=np.float32)
    u_c = rng.random((V, N, T), dtype=np.float32)
    eps = rng.standard_normal((V, N, T, 2)).astype(np.float32)
    f = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    z, pred = ops.gmm2d_sample(T_(pmf, dev), f(mus), f(ls), f(rho), N, log_pis=f(lp), noise=(f(u_z), f(u_c), f(eps)))
    z_ref, pred_ref = restate.gmm2d_sample(pmf, lp, mus, ls, rho, u_z, u_c, eps)
    np.testing.assert_array_equal(z.cpu().numpy(), z_ref)
    if Z > 1:
        assert not (z_ref == Z // 2).any()
    got = pred.cpu().numpy()
    bad = np.abs(got - pred_ref) > 1e-5 * (1.0 + np.abs(pred_ref))
    # a component boundary can fall on different sides by one float32 ulp of exp(): at most a couple of draws
    assert bad.any(axis=-1).sum() <= (2 if G > 1 else 0)


def test_gmm2d_sample_philox_statistics_and_shard_invariance(dev):
    """Philox mode: latent frequencies follow the pmf, per-(latent, timestep) sample moments follow the GMM
    parameters within CLT bounds, the stream is independent of sharding, and seed_dev shifts the key."""
    from ccmpc_b200 import ops
    from oracle import synth
    O, K, T, Z, N = 2, 4, 8, 25, 400_000
    sc = synth.make_scene(31, O, K, T, flat=False)
    pmf, mus, ls, rho = synth.gmm2d_params(sc, Z=Z, slots=[3, 7, 11, 20])
    f = lambda a: torch.from_numpy(a).to(dev)
    args = (T_(pmf, dev), f(mus), f(ls), f(rho))
    z, pred = ops.gmm2d_sample(*args, N, seed=5)
    zc, pc = z.cpu().numpy(), pred.cpu().numpy().astype(np.float64)
    for o in range(O):
        freq = np.bincount(zc[o], minlength=Z) / N
        assert np.all(np.abs(freq - pmf[o]) < 5 * np.sqrt(pmf[o] * (1 - pmf[o]) / N) + 1e-12)
        for j, slot in enumerate([3, 7, 11, 20]):
            sel = pc[o][zc[o] == slot]
            if len(sel) < 5000:
                continue
            for t in (0, T - 1):
                S = sc.cov()[o, j, t]
                m = sc.mean[o, j, t] - synth.MINPOS
                assert np.all(np.abs(sel[:, t].mean(axis=0) - m) < 5 * np.sqrt(np.diag(S) / len(sel)) + 1e-4)
                assert relerr(np.cov(sel[:, t].T), S) < 0.05
            # independent draws per timestep (GMM2D.rsample draws eps per (sample, t))
            assert abs(np.corrcoef(sel[:, 0, 0], sel[:, 1, 0])[0, 1]) < 0.02
    # shards
    za, pa = ops.gmm2d_sample(*args, N // 2, seed=5)
    zb, pb = ops.gmm2d_sample(*args, N - N // 2, seed=5, sample_offset=N // 2)
    assert torch.equal(torch.cat([za, zb], dim=1), z) and torch.equal(torch.cat([pa, pb], dim=1), pred)
    # the device seed word is added to the key
    sd = torch.tensor([2], dtype=torch.int64, device=dev)
    z2, p2 = ops.gmm2d_sample(*args, N, seed=3, seed_dev=sd)
    assert torch.equal(z2, z) and torch.equal(p2, pred)
    z3, _ = ops.gmm2d_sample(*args, N, seed=6)
    assert not torch.equal(z3, z)
    # the sampler's output is the ingest format of group_by_mode
    g = ops.group_by_mode(pred, z, T_(pmf, dev), minpos=tuple(synth.MINPOS))
    assert int(g["offsets"][-1]) == O * N and not g["status"].any()


def test_step_plan_sampler_mode_matches_ingest_plan(dev):
    """StepPlan.from_gmm (GMM2D parameters in, sampling + grouping + constraint build in one CUDA graph) gives
    exactly the rows of StepPlan.from_predictions fed with the samples the sampler draws for the same key, and
    a new step_seed gives new draws without re-capturing the graph."""
    from ccmpc_b200 import midlevel, ops
    from oracle import synth
    O, K, T, Z, N = 3, 3, 8, 25, 3000
    sc = synth.make_scene(55, O, K, T)
    pmf, mus, ls, rho = synth.gmm2d_params(sc, Z=Z, slots=[2, 9, 17])
    plan = midlevel.StepPlan.from_gmm(O, N, T, sc.pos_last, prediction_horizon=T, Z=Z, minpos=tuple(synth.MINPOS),
                                      device=dev, builder="both", seed=100)
    ing = midlevel.StepPlan.from_predictions(O, N, T, sc.pos_last, prediction_horizon=T, S=Z,
                                             minpos=tuple(synth.MINPOS), device=dev, builder="both")
    f = lambda a: torch.from_numpy(a).to(dev)
    prev = None
    for step in (0, 1):
        res = {k: v.copy() for k, v in plan.run_gmm(pmf, mus, ls, rho, ref_traj=sc.ref_traj, step_seed=step).items()}
        z, pred = ops.gmm2d_sample(T_(pmf, dev), f(mus), f(ls), f(rho), N, seed=100 + step)
        want = ing.run_predictions(pred.cpu().numpy(), z.cpu().numpy(), pmf, ref_traj=sc.ref_traj)
        assert list(res["g_n_states"]) == [K] * O
        for k in res:
            np.testing.assert_array_equal(res[k], want[k], err_msg=f"{k} differs at step {step}")
        if prev is not None:
            assert not np.array_equal(prev["counts"], res["counts"]) or not np.array_equal(prev["mean"], res["mean"])
        prev = res
    # the Monte Carlo means converge on the GMM means (3000 samples / 3 modes)
    for o in range(O):
        for j in range(K):
            c = o * Z + j
            se = 5 * np.sqrt(np.diagonal(sc.cov()[o, j], axis1=-2, axis2=-1) / (N / K / 2))
            assert np.all(np.abs(res["mean"][c] - sc.mean[o, j]) < se + 1e-3)


# --------------------------------------------------------- fused rollout + evaluation (no trajectory tensor)
@pytest.mark.parametrize("Tsh,n", [(16, 100_003), (8, 5000), (5, 777), (16, 1)])
def test_mc_risk_equals_rollout_then_sample_sums(dev, Tsh, n):
    """ccmpc_mc_risk draws the SAME samples as ccmpc_mc_rollout (same Philox counters) and applies the same A9
    predicate as ccmpc_sample_sums: collision counts bit-exact, soft-cost sums to round-off, and shards add up."""
    from ccmpc_b200 import ops
    from oracle import synth
    sc, mean, cov, cross = _scene_moments(9, 2, 3, Tsh + 1)
    C = 6
    rng = np.random.default_rng(1)
    # ego plans close to each cell's mean path so that hits are common
    ego = mean[:, 1:Tsh + 1] + rng.normal(size=(C, Tsh, 2)) * 1.5
    args = (T_(mean, dev), T_(cov, dev), T_(cross, dev))
    traj, _, st, x0 = ops.mc_rollout(*args, None, n, Tsh, seed=77)
    offsets = torch.arange(C + 1, dtype=torch.int64, device=dev) * n
    pivot = T_(mean[:, 1:Tsh + 1], dev)
    sums = ops.sample_sums(traj.view(C * n, Tsh, 2), offsets, pivot, ego=T_(ego, dev))
    risk, st2, x0b = ops.mc_risk(*args, None, n, Tsh, T_(ego, dev), seed=77)
    assert torch.equal(st, st2) and torch.equal(x0, x0b)
    L, R = ops.sums_layout(Tsh), ops.risk_layout(Tsh)
    s, r = sums.cpu().numpy(), risk.cpu().numpy()
    np.testing.assert_array_equal(r[:, R["n"]], s[:, L["n"]])
    np.testing.assert_array_equal(r[:, R["coll"]:R["coll"] + Tsh], s[:, L["coll"]:L["coll"] + Tsh])
    np.testing.assert_array_equal(r[:, R["any"]], s[:, L["any"]])
    assert r[:, R["coll"]:R["coll"] + Tsh].sum() > 0 or n == 1
    np.testing.assert_allclose(r[:, R["pen"]:R["pen"] + Tsh], s[:, L["pen"]:L["pen"] + Tsh], rtol=1e-11, atol=1e-9)
    np.testing.assert_allclose(r[:, R["d2s"]:R["d2s"] + Tsh], s[:, L["d2s"]:L["d2s"] + Tsh], rtol=1e-11)
    # one shared ego plan [Tsh,2] and two sample shards
    if n > 1:
        e1 = T_(ego[0], dev)
        whole, _, _ = ops.mc_risk(*args, x0, n, Tsh, e1, seed=77)
        a, _, _ = ops.mc_risk(*args, x0, n // 3, Tsh, e1, seed=77)
        b, _, _ = ops.mc_risk(*args, x0, n - n // 3, Tsh, e1, seed=77, sample_offset=n // 3)
        w, ab = whole.cpu().numpy(), (a + b).cpu().numpy()
        np.testing.assert_array_equal(ab[:, :R["pen"]], w[:, :R["pen"]])
        np.testing.assert_allclose(ab, w, rtol=1e-11, atol=1e-9)
    # empty batch
    z0, _, _ = ops.mc_risk(*args, None, 0, Tsh, T_(ego, dev))
    assert not z0.any()


def test_group_by_mode_fused_yaws_equal_yaws_from_positions(dev):
    """The yaws written while the grouped rows are gathered are bit-identical to a separate
    ccmpc_yaws_from_positions pass over the grouped tensor (ovehicle.py:72-76)."""
    from ccmpc_b200 import ops
    rng = np.random.default_rng(5)
    V, N, T, S = 3, 4000, 8, 25
    pred, z, pmf = _group_inputs(rng, V, N, T, S, sparse=True)
    minpos = np.asarray([151.25, -93.5])
    pos_last = minpos + rng.normal(size=(V, 2)) * 3
    out = ops.group_by_mode(T_(pred, dev, torch.float32), T_(z, dev, torch.int32), T_(pmf, dev), minpos=minpos,
                            pos_last=T_(pos_last, dev))
    pos_last_cell = T_(np.repeat(pos_last, S, axis=0), dev)
    want = ops.yaws_from_positions(out["samples"], out["offsets"], pos_last_cell)
    assert torch.equal(out["yaws"], want)


# --------------------------------------------------------- fused rollout + store + sums
@pytest.mark.parametrize("Tsh,n", [(16, 100_003), (8, 5000), (16, 130), (8, 1), (7, 4099), (13, 1000), (1, 100), (9, 33)])
def test_mc_rollout_sums_equals_rollout_then_sample_sums(dev, Tsh, n):
    """ccmpc_mc_rollout_sums = ccmpc_mc_rollout followed by ccmpc_sample_sums: the trajectories are bit-identical,
    the counts bit-exact, the power sums equal to fp64 round-off (different partition of the same additions)."""
    from ccmpc_b200 import ops
    sc, mean, cov, cross = _scene_moments(9, 2, 3, Tsh + 1)
    C = 6
    rng = np.random.default_rng(2)
    ego = mean[:, 1:Tsh + 1] + rng.normal(size=(C, Tsh, 2)) * 1.5
    args = (T_(mean, dev), T_(cov, dev), T_(cross, dev))
    pivot = T_(mean[:, 1:Tsh + 1], dev)
    traj, _, st, x0 = ops.mc_rollout(*args, None, n, Tsh, seed=31)
    offsets = torch.arange(C + 1, dtype=torch.int64, device=dev) * n
    want = ops.sample_sums(traj.view(C * n, Tsh, 2), offsets, pivot, ego=T_(ego, dev))
    traj2, sums, st2, x0b = ops.mc_rollout_sums(*args, None, n, Tsh, pivot, T_(ego, dev), seed=31)
    assert torch.equal(traj2, traj) and torch.equal(st, st2) and torch.equal(x0, x0b)
    L = ops.sums_layout(Tsh)
    s, w = sums.cpu().numpy(), want.cpu().numpy()
    np.testing.assert_array_equal(s[:, 0], w[:, 0])
    np.testing.assert_array_equal(s[:, L["coll"]:L["pen"]], w[:, L["coll"]:L["pen"]])          # coll[T], any
    scale = np.maximum(np.abs(w), 1.0)
    assert np.max(np.abs(s - w)[:, :L["yaw"]] / scale[:, :L["yaw"]]) < 1e-11
    assert not s[:, L["yaw"]:].any()
    # without materialising, with one shared ego plan
    _, sums_b, _, _ = ops.mc_rollout_sums(*args, x0, n, Tsh, pivot, T_(ego[0], dev), seed=31, materialize=False)
    want_b = ops.sample_sums(traj.view(C * n, Tsh, 2), offsets, pivot, ego=T_(ego[0], dev))
    np.testing.assert_array_equal(sums_b.cpu().numpy()[:, L["coll"]:L["pen"]], want_b.cpu().numpy()[:, L["coll"]:L["pen"]])
    # moments downstream agree
    ma = ops.moments_from_sums(sums, pivot, want_yaw=False)
    mb = ops.moments_from_sums(want, pivot, want_yaw=False)
    if n > 1:
        assert relerr(ma["cov"].cpu().numpy(), mb["cov"].cpu().numpy()) < 1e-10
        assert relerr(ma["mean"].cpu().numpy(), mb["mean"].cpu().numpy()) < 1e-13


@pytest.mark.parametrize("accel", [0, 2])
def test_minkowski_halfplanes_batched_cells_are_independent(dev, golden, accel):
    """More cells than persistent blocks (the kernel then walks several cells per block through its double-buffered
    sums): every copy of a golden cell, wherever it sits in the batch and whichever eps its neighbours carry,
    gives bit-identical rows to the small launch checked against the reference above."""
    from ccmpc_b200 import ops
    g = golden("ref_pipeline")
    T = int(g["T"])
    samples, offsets, pivot, sums, eps = _pipeline(dev, g)
    C0 = sums.shape[0]
    ref = T_(g["ref_traj"], dev)
    small = ops.minkowski_halfplanes(sums, pivot, ref, eps, prediction_horizon=T, accel=accel)
    reps = 3001 // C0 + 1                                   # > 148 x 5 blocks, not a multiple of the grid
    C = reps * C0
    eps_b = eps.repeat(reps)
    big = ops.minkowski_halfplanes(sums.repeat(reps, 1), pivot.repeat(reps, 1, 1), ref, eps_b, prediction_horizon=T,
                                   accel=accel)
    for k, v in small.items():
        got = big[k].reshape((reps, C0) + tuple(v.shape[1:]))
        assert torch.equal(got, v.unsqueeze(0).expand_as(got)) or \
            bool(((got == v.unsqueeze(0)) | (got.isnan() & v.unsqueeze(0).isnan())).all()), k
    # a different eps on every other cell must only change those cells
    eps_c = eps_b.clone()
    eps_c[1::2] *= 0.5
    mixed = ops.minkowski_halfplanes(sums.repeat(reps, 1), pivot.repeat(reps, 1, 1), ref, eps_c, prediction_horizon=T,
                                     accel=accel)
    assert torch.equal(mixed["nd"][0::2], big["nd"][0::2]) and not torch.equal(mixed["nd"][1::2], big["nd"][1::2])
    half = ops.minkowski_halfplanes(sums.repeat(reps, 1)[1::2].contiguous(), pivot.repeat(reps, 1, 1)[1::2].contiguous(),
                                    ref, eps_c[1::2].contiguous(), prediction_horizon=T, accel=accel)
    assert torch.equal(mixed["nd"][1::2], half["nd"]) and torch.equal(mixed["prob_lower"][1::2], half["prob_lower"])


@pytest.mark.gpu
@pytest.mark.parametrize("dyn", ["unicycle", "single_integrator"])
@pytest.mark.parametrize("V,N,Z,T,G", [(3, 700, 25, 16, 1), (2, 333, 7, 9, 3), (1, 1, 1, 1, 1), (2, 129, 4, 2, 1)])
def test_gmm2d_sample_dynamics_supplied_noise_matches_oracle(dev, dyn, V, N, Z, T, G):
    """Control-space GMM2D + dynamic.integrate_samples on the device (prediction.py:85-86; published Trajectron++
    unicycle / single integrator, parity unpinned): same uniforms / normals in -> the oracle's latents bit-exact
    and its float32 positions within float32 recursion round-off (the reference arithmetic IS float32; sinf/cosf
    and FMA contraction differ from NumPy by an ulp per step, and (sin a - sin b)/dphi amplifies that by up to
    1/dphi = 100 -> 1e-4 m per step at 10 m/s)."""
    from ccmpc_b200 import ops
    from oracle import restate, synth
    pmf, lp, mus, ls, rho, x0 = synth.unicycle_control_params(V, Z, T, seed=200 + G, G=G)
    rng = np.random.default_rng(300 + T)
    u_z = rng.random((V, N), dtype=np.float32)
    u_c = rng.random((V, N, T), dtype=np.float32)
    eps = rng.standard_normal((V, N, T, 2)).astype(np.float32)
    f = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    z, pred = ops.gmm2d_sample(T_(pmf, dev), f(mus), f(ls), f(rho), N, log_pis=f(lp), noise=(f(u_z), f(u_c), f(eps)),
                               dynamics=dyn, x0=f(x0), dt=0.5)
    z_ref, u_ref = restate.gmm2d_sample(pmf, lp, mus, ls, rho, u_z, u_c, eps)
    pos_ref = restate.integrate_samples(u_ref, x0, 0.5, dyn)
    np.testing.assert_array_equal(z.cpu().numpy(), z_ref)
    err = np.abs(pred.cpu().numpy().astype(np.float64) - pos_ref).max(axis=(2, 3))
    # a G > 1 component boundary one float32 ulp away can flip a draw: at most a couple of samples
    assert (err > 2e-3).sum() <= (2 if G > 1 else 0), err.max()


@pytest.mark.gpu
def test_gmm2d_sample_dynamics_philox_consistent_with_control_sampler(dev):
    """Philox mode: the dynamics entry point draws the SAME controls as ccmpc_gmm2d_sample for the same key, so
    integrating that entry point's controls with the oracle reproduces its positions; latents are bit-identical,
    the stream is shard invariant, dynamics = none delegates, and the output feeds group_by_mode."""
    from ccmpc_b200 import ops
    from oracle import restate, synth
    V, Z, T, N = 4, 25, 16, 50_000
    pmf, lp, mus, ls, rho, x0 = synth.unicycle_control_params(V, Z, T, seed=9, G=2)
    f = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    args = (T_(pmf, dev), f(mus), f(ls), f(rho))
    zc, u = ops.gmm2d_sample(*args, N, log_pis=f(lp), seed=11)
    z0, u0 = ops.gmm2d_sample(*args, N, log_pis=f(lp), seed=11, dynamics="none")
    assert torch.equal(z0, zc) and torch.equal(u0, u)
    for dyn in ("unicycle", "single_integrator"):
        z, pos = ops.gmm2d_sample(*args, N, log_pis=f(lp), seed=11, dynamics=dyn, x0=f(x0), dt=0.5)
        assert torch.equal(z, zc)
        want = restate.integrate_samples(u.cpu().numpy(), x0, 0.5, dyn)
        assert np.abs(pos.cpu().numpy().astype(np.float64) - want).max() < 2e-3
        za, pa = ops.gmm2d_sample(*args, 20_001, log_pis=f(lp), seed=11, dynamics=dyn, x0=f(x0))
        zb, pb = ops.gmm2d_sample(*args, N - 20_001, log_pis=f(lp), seed=11, sample_offset=20_001, dynamics=dyn, x0=f(x0))
        assert torch.equal(torch.cat([za, zb], dim=1), z) and torch.equal(torch.cat([pa, pb], dim=1), pos)
        sd = torch.tensor([4], dtype=torch.int64, device=dev)
        z2, p2 = ops.gmm2d_sample(*args, N, log_pis=f(lp), seed=7, seed_dev=sd, dynamics=dyn, x0=f(x0))
        assert torch.equal(z2, z) and torch.equal(p2, pos)
    # unicycle samples: speeds stay physical and the grouped rows are complete
    z, pos = ops.gmm2d_sample(*args, N, log_pis=f(lp), seed=11, dynamics="unicycle", x0=f(x0))
    step = (pos[:, :, 1:] - pos[:, :, :-1]).norm(dim=-1) / 0.5
    assert float(step.max()) < 10 + 1.5 * 8 + 3 and torch.isfinite(pos).all()
    g = ops.group_by_mode(pos, z, T_(pmf, dev), minpos=tuple(synth.MINPOS))
    assert int(g["offsets"][-1]) == V * N and not g["status"].any()
    with pytest.raises(ValueError):
        ops.gmm2d_sample(*args, N, log_pis=f(lp), dynamics="unicycle")
    with pytest.raises(ValueError):
        ops.gmm2d_sample(*args, N, log_pis=f(lp), dynamics="bicycle", x0=f(x0))


@pytest.mark.gpu
def test_step_plan_sampler_mode_with_unicycle_dynamics(dev):
    """StepPlan.from_gmm(dynamics="unicycle"): control-space GMM2D + unicycle integration + grouping + constraint
    build in one CUDA graph gives exactly the rows of the ingest plan fed with the positions the dynamics sampler
    draws for the same key; a new x0 / step_seed changes the rows without re-capturing."""
    from ccmpc_b200 import midlevel, ops
    from oracle import synth
    V, Z, T, N = 3, 25, 8, 4000
    pmf, lp, mus, ls, rho, x0 = synth.unicycle_control_params(V, Z, T, seed=77)
    pos_last = x0[:, :2].astype(np.float64) + synth.MINPOS
    ref = synth.reference_trajectory(T)
    kw = dict(minpos=tuple(synth.MINPOS), device=dev, builder="affine")
    plan = midlevel.StepPlan.from_gmm(V, N, T, pos_last, prediction_horizon=T, Z=Z, seed=40, dynamics="unicycle",
                                      dt=0.5, **kw)
    ing = midlevel.StepPlan.from_predictions(V, N, T, pos_last, prediction_horizon=T, S=Z, **kw)
    f = lambda a: torch.from_numpy(a).to(dev)
    prev = None
    for step, x0s in ((0, x0), (1, x0 + np.float32([0.5, -0.25, 0.1, 0.3]))):
        res = {k: v.copy() for k, v in plan.run_gmm(pmf, mus, ls, rho, ref_traj=ref, step_seed=step, x0=x0s).items()}
        z, pos = ops.gmm2d_sample(T_(pmf, dev), f(mus), f(ls), f(rho), N, seed=40 + step, dynamics="unicycle",
                                  x0=f(np.ascontiguousarray(x0s)), dt=0.5)
        want = ing.run_predictions(pos.cpu().numpy(), z.cpu().numpy(), pmf, ref_traj=ref)
        for k in res:
            np.testing.assert_array_equal(res[k], want[k], err_msg=f"{k} differs at step {step}")
        assert int(res["g_offsets"][-1]) == V * N and not res["g_status"].any()
        if prev is not None:
            assert not np.array_equal(prev["mean"], res["mean"])
        prev = res
    with pytest.raises(ValueError):
        midlevel.StepPlan.from_gmm(V, N, T, pos_last, prediction_horizon=T, Z=Z).run_gmm(pmf, mus, ls, rho, ref_traj=ref, x0=x0)


@pytest.mark.parametrize("Tsh,origin", [(16, "first"), (8, "first"), (16, "mid"), (11, "negative")])
def test_mc_rollout_sums_counts_when_the_ego_plan_is_near_the_origin(dev, Tsh, origin):
    """The A9 screen of the fused kernel works on the hi words of the coordinates, per sign of the window around the
    ego: put the ego plan on the origin (windows that hold zero), across it, and in the all-negative quadrant; the
    counts must stay bit-exact against rollout -> sample_sums (whose screen is built differently)."""
    from ccmpc_b200 import ops
    sc, mean, cov, cross = _scene_moments(4, 2, 2, Tsh + 1)
    C, n = 4, 20_011
    rng = np.random.default_rng(8)
    ego = mean[:, 1:Tsh + 1] + rng.normal(size=(C, Tsh, 2)) * 1.0
    at = {"first": ego[0, 0], "mid": ego[1, Tsh // 2], "negative": ego.max(axis=(0, 1)) + 500.0}[origin]
    mean, ego = mean - at, ego - at
    args = (T_(mean, dev), T_(cov, dev), T_(cross, dev))
    pivot = T_(mean[:, 1:Tsh + 1], dev)
    traj, _, _, _ = ops.mc_rollout(*args, None, n, Tsh, seed=5)
    offsets = torch.arange(C + 1, dtype=torch.int64, device=dev) * n
    want = ops.sample_sums(traj.view(C * n, Tsh, 2), offsets, pivot, ego=T_(ego, dev)).cpu().numpy()
    traj2, sums, _, _ = ops.mc_rollout_sums(*args, None, n, Tsh, pivot, T_(ego, dev), seed=5)
    assert torch.equal(traj2, traj)
    L = ops.sums_layout(Tsh)
    s = sums.cpu().numpy()
    assert want[:, L["coll"]:L["coll"] + Tsh].sum() > 0
    np.testing.assert_array_equal(s[:, L["coll"]:L["pen"]], want[:, L["coll"]:L["pen"]])
    np.testing.assert_allclose(s[:, L["pen"]:L["d2s"]], want[:, L["pen"]:L["d2s"]], rtol=1e-11, atol=1e-9)
    # the same predicate in plain torch fp64 (no FMA): d2 = dx*dx + dy*dy < R*R
    d = T_(ego, dev)[:, None] - traj
    hit = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) < ops.R_EV_OV * ops.R_EV_OV
    np.testing.assert_array_equal(s[:, L["coll"]:L["coll"] + Tsh], hit.sum(dim=1).cpu().numpy())


@pytest.mark.parametrize("Tsh", [16, 8])
def test_mc_rollout_sums_is_independent_of_the_world_size(dev, Tsh):
    """Eight sample shards (sharding.sample_shard, Philox counter = global sample index) rolled out and reduced
    separately, packed sums added: counts bit-exact and trajectories identical to the one-shard launch, power sums
    equal to round-off - what the NCCL all-reduce of the packed sums relies on at 2, 4 and 8 GPUs.  Also: more
    cells than SMs and a cell boundary inside a CTA's share of the tile list (the kernel's cross-cell split)."""
    from ccmpc_b200 import ops, sharding
    O, K = 3, 2
    sc, mean, cov, cross = _scene_moments(12, O, K, Tsh + 1)
    C, n = O * K, 9973                                    # prime: ragged shards, partial last tiles
    rng = np.random.default_rng(4)
    ego = mean[:, 1:Tsh + 1] + rng.normal(size=(C, Tsh, 2)) * 1.5
    args = (T_(mean, dev), T_(cov, dev), T_(cross, dev))
    pivot, ego_t = T_(mean[:, 1:Tsh + 1], dev), T_(ego, dev)
    traj, whole, st, x0 = ops.mc_rollout_sums(*args, None, n, Tsh, pivot, ego_t, seed=5)
    assert not st.any()
    L = ops.sums_layout(Tsh)
    for world in (2, 8):
        total = torch.zeros_like(whole)
        parts = []
        for rank in range(world):
            lo, nl = sharding.sample_shard(n, rank, world)
            t_r, s_r, _, x0_r = ops.mc_rollout_sums(*args, None, nl, Tsh, pivot, ego_t, seed=5, sample_offset=lo)
            assert torch.equal(x0_r, x0)                  # the one x0 draw per cell does not depend on the shard
            total += s_r
            parts.append(t_r)
        assert torch.equal(torch.cat(parts, dim=1), traj)
        w, s = whole.cpu().numpy(), total.cpu().numpy()
        np.testing.assert_array_equal(s[:, 0], w[:, 0])
        np.testing.assert_array_equal(s[:, L["coll"]:L["pen"]], w[:, L["coll"]:L["pen"]])
        np.testing.assert_allclose(s[:, :L["coll"]], w[:, :L["coll"]], rtol=1e-10, atol=1e-7)
        np.testing.assert_allclose(s[:, L["pen"]:L["d2s"]], w[:, L["pen"]:L["d2s"]], rtol=1e-10, atol=1e-9)
    # many small cells: several cells per CTA
    Cm = 200
    idx = np.arange(Cm) % C
    big = tuple(T_(a[idx], dev) for a in (mean, cov, cross))
    _, sums_m, st_m, _ = ops.mc_rollout_sums(*big, T_(x0.cpu().numpy()[idx], dev), 70, Tsh, T_(mean[idx, 1:Tsh + 1], dev),
                                             T_(ego[idx], dev), seed=5, materialize=False)
    _, sums_1, _, _ = ops.mc_rollout_sums(*args, x0, 70, Tsh, pivot, ego_t, seed=5, materialize=False)
    # cell index enters the Philox counter, so copies of a cell at other indices draw other samples: compare the
    # cells that sit at their own index, and the sample counts of all
    a, b = sums_m[:C].cpu().numpy(), sums_1.cpu().numpy()
    np.testing.assert_array_equal(a[:, L["coll"]:L["pen"]], b[:, L["coll"]:L["pen"]])     # counts: exact
    np.testing.assert_allclose(a, b, rtol=1e-10, atol=1e-8)                                # sums: other partition
    assert (sums_m[:, 0] == 70).all() and not st_m.any()
