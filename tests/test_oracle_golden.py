"""The NumPy restatement (oracle/restate.py) against the goldens produced by the
REFERENCE'S OWN functions (oracle/make_golden.py -> tests/golden/ref_*.npz).
CPU only.  This is what pins the oracle."""
import numpy as np
import pytest
import scipy.stats

from oracle import restate


def test_mvoe_matches_reference(golden):
    g = golden("ref_mvoe")
    for s1, s2, beta, Q in zip(g["Sigma1"], g["Sigma2"], g["beta"], g["Q"]):
        b, q = restate.mvoe(s1, s2)
        assert b == pytest.approx(beta, rel=1e-13, abs=0)
        np.testing.assert_allclose(q, Q, rtol=1e-13, atol=0)


def test_mvoe_known_answers():
    # SURVEY §8c KATs (computed with the reference file): radius 1 (+) radius 2 = radius 3
    b, q = restate.mvoe(np.eye(2), 4 * np.eye(2))
    assert b == pytest.approx(0.5, abs=1e-9)
    np.testing.assert_allclose(q, 9 * np.eye(2), atol=1e-7)
    b, q = restate.mvoe([[2, .5], [.5, 1]], [[1, -.2], [-.2, 3]])
    assert b == pytest.approx(0.8915954797056089, rel=1e-12)
    np.testing.assert_allclose(q, [[6.134765219261006, 0.6824733389477273],
                                   [0.6824733389477273, 7.796371308894525]], rtol=1e-12)


def test_pair_moments_matches_reference(golden):
    g = golden("ref_predict_moments")
    for rows, ci, cm, ct in zip(g["rows"], g["cov_infer"], g["cov_mu"], g["cov_t"]):
        a, b, c = restate.pair_moments([rows[0], rows[1], rows[2], rows[3]])
        np.testing.assert_array_equal(a, ci)
        np.testing.assert_array_equal(b, cm)
        np.testing.assert_array_equal(c, ct)


def test_tangent_matches_reference(golden):
    g = golden("ref_tangent")
    for i in range(len(g["c"])):
        mu, S, c, m, a = g["mu"][i], g["Sigma"][i], g["c"][i], g["m"][i], g["a"][i]
        (n1, d1), (n2, d2) = restate.tangents_of_slope(mu, S, c, m)
        np.testing.assert_array_equal(n1, g["n"][i])
        assert d1 == g["d1"][i] and d2 == g["d2"][i]
        assert restate.point_line_distance(n1, d1, a) == g["dist"][i][0]
        assert restate.point_line_distance(n2, d2, a) == g["dist"][i][1]
        n, d, idx = restate.closest_tangent(mu, S, c, m, a)
        assert idx == g["idx"][i] and d == g["d_star"][i]
        for j in (0, 1):
            assert restate.closest_tangent(mu, S, c, m, a, const_idx=j)[1] == g["d_forced"][i][j]
    # KATs of SURVEY §8c
    assert g["d_star"][0] == pytest.approx(0.5) and g["idx"][0] == 1
    assert g["d_star"][1] == pytest.approx(2.5) and g["idx"][1] == 0


def test_tangent_degenerate_returns_four_nones():
    out = restate.closest_tangent([0, 0], np.zeros((2, 2)), 1.0, 0.3, [1, 1])
    assert out == (None, None, None, None)


def test_scale_and_lower_bound_match_reference(golden):
    g = golden("ref_scale_lower")
    for i in range(len(g["eps"])):
        s = restate.scale_factor(g["cov_infer"][i], g["cov_mu"][i], g["cov_t"][i], g["Gamma"][i])
        l = restate.lower_bound(g["cov_infer"][i], g["cov_mu"][i], g["cov_t"][i], g["eps"][i])
        assert s == g["scale"][i]
        assert l == g["lower"][i]
    assert g["lower"][0] == pytest.approx(0.29104264152591025, rel=1e-12)
    assert g["scale"][0] == pytest.approx(3.3529447079923145, rel=1e-12)


def test_vertices_match_reference(golden):
    g = golden("ref_vertices")
    v = restate.box_vertices(g["centers"], g["theta"], g["lw"])
    np.testing.assert_allclose(v, g["vertices"], rtol=0, atol=1e-12)
    v = restate.box_vertices(g["centers"][:7], 0.3, [4.5, 2.5])
    np.testing.assert_allclose(v, g["vertices_scalar_theta"], rtol=0, atol=1e-12)


def test_closed_form_quantiles_match_scipy(golden):
    g = golden("ref_scalars")
    for e, ref in zip(g["eps"], g["chi2_ppf_1m_eps"]):
        assert restate.chi2_ppf_df2(1 - e) == pytest.approx(ref, rel=1e-11)
    assert restate.chi2_ppf_df2(0.9999) == pytest.approx(float(g["chi2_ppf_target"]), rel=1e-11)
    assert float(g["chi2_ppf_target"]) == pytest.approx(18.420680743952584, rel=1e-14)
    for x, ref in zip(g["chi2_cdf_x"], g["chi2_cdf"]):
        assert restate.chi2_cdf_df2(x) == pytest.approx(ref, rel=1e-13)


def test_pipeline_loop_matches_reference_functions(golden):
    """restate.minkowski_cell / affine_cell with the restated functions equal
    the same loops driven by the reference's functions."""
    g = golden("ref_pipeline")
    T = int(g["T"])
    offs = g["offsets"]
    for c in range(len(offs) - 1):
        cell = g["samples"][offs[c]:offs[c + 1]]
        eps = g["eps_ura"][g["cell_obstacle"][c], g["cell_mode"][c]]
        r = restate.minkowski_cell(cell, g["ref_traj"], eps, T)
        for k in ["n", "d", "A", "b", "Q", "Q1", "cov_infer", "cov_mu", "prob_pair", "prob_lower", "mean"]:
            np.testing.assert_allclose(r[k], g["mk_" + k][c], rtol=1e-12, atol=1e-12, err_msg=k)
        np.testing.assert_array_equal(r["idx"], g["mk_idx"][c])
        np.testing.assert_array_equal(r["side"], g["mk_side"][c])
        r = restate.affine_cell(cell, g["ref_traj"], eps, T)
        for k in ["n", "d", "offset", "rhs", "mean", "cov"]:
            np.testing.assert_allclose(r[k], g["af_" + k][c], rtol=1e-12, atol=1e-12, err_msg=k)
        np.testing.assert_array_equal(r["idx"], g["af_idx"][c])
        np.testing.assert_array_equal(r["side"], g["af_side"][c])


def test_mvoe_iteration_counts_are_moderate():
    rng = np.random.default_rng(0)
    from oracle import synth
    s1 = synth.sym_to_full(synth.random_spd(rng, (300,)))
    s2 = synth.sym_to_full(synth.random_spd(rng, (300,)))
    its = [restate.mvoe(a, b, return_iters=True)[2] for a, b in zip(s1, s2)]
    assert max(its) < 60 and np.median(its) <= 14


def test_rollout_and_moments_roundtrip():
    """ideal_rollout from closed-form moments reproduces them statistically and
    cell_moments equals np.cov blockwise."""
    from oracle import synth
    scene = synth.make_scene(3, 1, 1, 6)
    mean, cov, cross = scene.moments()
    rng = np.random.default_rng(1)
    n = 200_000
    Z = rng.standard_normal((n, 5, 2))
    x0 = mean[0, 0, 0] + scene.chol[0, 0, 0] @ np.array([0.3, -0.2])
    traj = restate.ideal_rollout(mean[0, 0], cov[0, 0], cross[0, 0], 5, Z, x0)
    m, c, x = restate.cell_moments(traj)
    # conditional on x0: mean_1 = mu_1 + A (x0 - mu_0); cov_1 = cond_cov_0 (exact law)
    A, L = restate.rollout_params(mean[0, 0], cov[0, 0], cross[0, 0], 5)
    np.testing.assert_allclose(m[0], mean[0, 0, 1] + A[0] @ (x0 - mean[0, 0, 0]), atol=5 * np.sqrt(cov[0, 0, 1].max() / n))
    np.testing.assert_allclose(c[0], L[0] @ L[0].T, rtol=2e-2, atol=1e-4)
    full = np.cov(traj.reshape(n, -1).T)
    np.testing.assert_allclose(c[3], full[6:8, 6:8], rtol=1e-12)
    np.testing.assert_allclose(x[3, 1], full[6:8, 2:4], rtol=1e-10, atol=1e-12)


def test_collision_counts_small():
    pos = np.array([[[0.0, 0.0], [10.0, 0.0]], [[1.0, 1.0], [3.0, 0.0]], [[5.0, 5.0], [6.0, 0.1]]])
    ego = np.array([[0.0, 0.0], [6.0, 0.0]])
    coll, any_, pen, d2s = restate.collision_counts(pos, ego, R=3.4)
    assert coll.tolist() == [2, 2] and any_ == 3
    assert pen[0] == pytest.approx(3.4 ** 2 + (3.4 ** 2 - 2.0))
    assert d2s[0] == pytest.approx(0 + 2 + 50)


def test_episode_cost_quirk():
    X = [np.array([160.0, -80.0, 0, 0]), np.array([166.0, -81.0, 0, 0])]
    U = [np.array([1.0, 0.1]), np.array([0.5, -0.2]), np.array([0.2, 0.0])]
    cost, costU = restate.episode_cost(X, U)
    R1 = np.array([[.5, .2], [.2, 1.]]); R2 = np.array([[.5, .1], [.1, 2.]])
    du = U[1] - U[0]
    exp_u = U[1] @ R1 @ U[1] + du @ R2 @ du      # only tau = H-2 survives the first loop
    assert costU == pytest.approx(exp_u)
    assert cost == pytest.approx(6 * ((166 - 167.174698) ** 2 + (-81 + 81.759842) ** 2) + exp_u)


def test_oracle_gmm2d_sample_restatement_is_a_gmm():
    """The GMM2D sampling restatement (prediction.py:81-86; Trajectron++ submodule absent => parity unpinned):
    latent frequencies follow the pmf, a never-weighted latent is never drawn, and per-latent sample moments
    equal mu and L L^T of the published parameterisation."""
    from oracle import restate
    rng = np.random.default_rng(0)
    V, Z, T, G, N = 1, 5, 3, 2, 200_000
    pmf = np.array([[0.5, 0.0, 0.3, 0.2, 0.0]])
    mus = rng.uniform(-5, 5, (V, Z, T, G, 2)).astype(np.float32)
    ls = rng.uniform(-1, 0.5, (V, Z, T, G, 2)).astype(np.float32)
    rho = rng.uniform(-0.9, 0.9, (V, Z, T, G)).astype(np.float32)
    lp = np.log(np.array([0.25, 0.75], np.float32)) * np.ones((V, Z, T, G), np.float32) + 3.0   # unnormalised
    z, pred = restate.gmm2d_sample(pmf, lp, mus, ls, rho, rng.random((V, N), dtype=np.float32),
                                   rng.random((V, N, T), dtype=np.float32),
                                   rng.standard_normal((V, N, T, 2)).astype(np.float32))
    freq = np.bincount(z[0], minlength=Z) / N
    assert np.all(np.abs(freq - pmf[0]) < 0.005) and freq[1] == 0 and freq[4] == 0
    sel = pred[0][z[0] == 0][:, 1].astype(np.float64)                 # latent 0, timestep 1: 2-component mixture
    w = np.array([0.25, 0.75])
    m = (w[:, None] * mus[0, 0, 1]).sum(axis=0)
    assert np.all(np.abs(sel.mean(axis=0) - m) < 0.02)
    S = np.zeros((2, 2))
    for g in range(G):
        sx, sy = np.exp(ls[0, 0, 1, g].astype(np.float64))
        r = float(rho[0, 0, 1, g])
        L = np.array([[sx, 0.0], [r * sy, sy * np.sqrt(1 - r * r)]])
        d = mus[0, 0, 1, g].astype(np.float64) - m
        S += w[g] * (L @ L.T + np.outer(d, d))
    assert np.abs(np.cov(sel.T) - S).max() < 0.03 * np.abs(S).max() + 0.02


def test_oracle_integrate_samples_solves_the_unicycle_ode():
    """The integrate_samples restatement (prediction.py:85-86; Trajectron++ submodule absent => parity unpinned):
    the published unicycle step is the EXACT solution of x' = v cos(phi), y' = v sin(phi), phi' = dphi, v' = a
    under piecewise-constant controls, so it must agree with a fine RK4 integration of that ODE in float64;
    the small-angle branch (|dphi| <= 1e-2) is its dphi -> 0 limit; the single integrator is a cumulative sum."""
    from oracle import restate
    rng = np.random.default_rng(7)
    V, N, T, dt = 3, 40, 12, 0.5
    u = np.stack([rng.uniform(-0.4, 0.4, (V, N, T)), rng.uniform(-2, 2, (V, N, T))], axis=-1).astype(np.float32)
    u[:, ::4, :, 0] = rng.uniform(-9e-3, 9e-3, u[:, ::4, :, 0].shape)           # straight segments
    x0 = np.stack([rng.uniform(0, 50, V), rng.uniform(0, 40, V), rng.uniform(-3, 3, V), rng.uniform(3, 10, V)],
                  axis=1).astype(np.float32)
    got = restate.integrate_samples(u, x0, dt, "unicycle").astype(np.float64)

    def rhs(s, c):
        return np.stack([s[..., 3] * np.cos(s[..., 2]), s[..., 3] * np.sin(s[..., 2]), c[..., 0], c[..., 1]], axis=-1)
    s = np.broadcast_to(x0[:, None, :].astype(np.float64), (V, N, 4)).copy()
    sub, want = 64, np.zeros((V, N, T, 2))
    for t in range(T):
        c = u[:, :, t].astype(np.float64).copy()
        small = np.abs(c[..., 0]) <= 1e-2
        c[..., 0] = np.where(small, 0.0, c[..., 0])          # the published small-angle branch holds the heading
        h = dt / sub
        for _ in range(sub):
            k1 = rhs(s, c); k2 = rhs(s + 0.5 * h * k1, c); k3 = rhs(s + 0.5 * h * k2, c); k4 = rhs(s + h * k3, c)
            s = s + h / 6 * (k1 + 2 * k2 + 2 * k3 + k4)
        want[:, :, t] = s[..., :2]
    # float32 recursion over 12 steps (incl. the (sin a - sin b)/dphi cancellation) against the float64 ODE
    assert np.abs(got - want).max() < 5e-3
    si = restate.integrate_samples(u, x0, dt, "single_integrator").astype(np.float64)
    want_si = x0[:, None, None, :2].astype(np.float64) + np.cumsum(u.astype(np.float64), axis=2) * dt
    assert np.abs(si - want_si).max() < 1e-4
