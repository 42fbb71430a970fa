"""The drop-in planner class ``ccmpc_b200.midlevel.MidlevelConstraints`` on a B200, method by method against
what THE REFERENCE'S OWN MidlevelAgent methods produced (tests/golden/ref_planner_*.npz, made by
oracle/make_golden_planner.py from the unmodified reference code) and, at sizes without a golden, against the
restated drivers of oracle/restate_planner.py (themselves pinned to the same goldens on the CPU).

Bars: best_idx / side / t bit-exact; rows, polytopes, moments within 1e-7 relative (north_star: 1e-5)."""
import os
import pickle
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROW_RTOL = 1e-7


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from ccmpc_b200 import _lib
    _lib.load(build_if_missing=False)
    return torch.device("cuda:0")


class FakeOVehicle:
    """The attributes of midlevel/ovehicle.py:119-131 the builders read."""

    def __init__(self, pred_positions, pred_yaws, bbox):
        self.pred_positions, self.pred_yaws = list(pred_positions), list(pred_yaws)
        self.bbox = np.asarray(bbox, dtype=np.float64)
        self.n_states = len(self.pred_positions)


def make_ovehicles(pred, z, minpos, pos_last, K_per_obstacle, bbox):
    from oracle import restate
    pos = pred.astype(np.float64) + minpos                                    # v8ideal:486
    ovs, cells, yaws = [], [], []
    for o in range(pos.shape[0]):
        pp = [pos[o][z[o] == k] for k in range(K_per_obstacle[o])]            # v8ideal:489-494
        yy = [restate.yaws_from_positions(p, pos_last[o]) for p in pp]
        ovs.append(FakeOVehicle(pp, yy, bbox))
        cells.append(pp)
        yaws.append(yy)
    return ovs, cells, yaws


def params_of(ovs, frame):
    return types.SimpleNamespace(O=len(ovs), K=np.array([ov.n_states for ov in ovs]), frame=frame)


def dense_tko(nested, tail):
    T, Kmax, O = len(nested), len(nested[0]), len(nested[0][0])
    out = np.full((T, Kmax, O) + tail, np.nan)
    for t in range(T):
        for k in range(Kmax):
            for o in range(O):
                if nested[t][k][o] is not None:
                    out[t, k, o] = nested[t][k][o]
    return out


def dense_ok(nested):
    O, Kmax = len(nested), len(nested[0])
    out = np.full((O, Kmax), np.nan)
    for o in range(O):
        for k in range(Kmax):
            if nested[o][k] is not None:
                out[o, k] = nested[o][k]
    return out


def check_tuple(built, ref, pre="", rtol=ROW_RTOL, affine=False):
    """``built``: the 9-tuple of a builder; ``ref``: golden npz or oracle dict with keys pre+name."""
    hp, vertices, A_union, b_union, OVc, direct, ov_mean, ov_cov, last = built
    g = lambda k: ref[pre + k]
    R = g("n").shape[0]
    np.testing.assert_array_equal(hp.t.reshape(-1), g("t"))
    np.testing.assert_array_equal(hp.side.reshape(-1), g("side"))                        # bit-exact
    np.testing.assert_allclose(hp.n.reshape(R, 2), g("n"), rtol=rtol, atol=0)
    np.testing.assert_allclose(hp.rhs.reshape(-1), g("rhs"), rtol=rtol, atol=0)
    assert not hp.status.any()
    np.testing.assert_allclose(dense_tko(A_union, (4, 2)), g("A_union"), rtol=1e-9, atol=1e-12, equal_nan=True)
    np.testing.assert_allclose(dense_tko(b_union, (4,)), g("b_union"), rtol=1e-9, atol=1e-9, equal_nan=True)
    np.testing.assert_allclose(np.stack([dense_ok(a) for a in ov_mean]), g("ov_mean"), rtol=1e-10, equal_nan=True)
    np.testing.assert_allclose(np.stack([dense_ok(a) for a in ov_cov]), g("ov_cov"), rtol=1e-7, equal_nan=True)
    assert bool(OVc) == bool(g("OVconstraint"))
    assert all(d is None for d in direct) and last == 0
    if not affine:
        np.testing.assert_allclose(hp.extra["prob_lower_save"], g("prob_lower_save"), rtol=rtol)
    return hp


def check_moments_pickle(path, ref_mean, ref_cov, ref_cross, rtol=1e-8):
    """The file the repo wrote, read with THE REFERENCE'S indexing (v8ideal:2662-2681):
    mean_p0p1[o][k][t], cov_p0p1[o][k][t+1], cross_cov[o][k][t+1][t]; None in the unused slots (:2580-2582)."""
    with open(path, "rb") as f:
        d = pickle.load(f)
    O, Kmax, T = ref_mean.shape[:3]
    for key in ("mean_p0p1", "cov_p0p1", "cross_cov"):
        assert len(d[key]) == O and all(len(x) == Kmax for x in d[key]) and all(len(y) == T for x in d[key] for y in x)
    assert all(len(d["cross_cov"][o][k][t]) == T - 1 for o in range(O) for k in range(Kmax) for t in range(T))
    for o in range(O):
        for k in range(Kmax):
            if np.isnan(ref_mean[o, k, 0, 0]):
                assert d["mean_p0p1"][o][k][0] is None and d["cov_p0p1"][o][k][0] is None
                continue
            for t in range(T):
                np.testing.assert_allclose(d["mean_p0p1"][o][k][t], ref_mean[o, k, t], rtol=rtol)
                np.testing.assert_allclose(d["cov_p0p1"][o][k][t], ref_cov[o, k, t], rtol=rtol, atol=1e-12)
                assert d["mean_p0p1"][o][k][t].shape == (2,) and d["cov_p0p1"][o][k][t].shape == (2, 2)
                for tau in range(T - 1):
                    if tau < t:
                        np.testing.assert_allclose(d["cross_cov"][o][k][t][tau], ref_cross[o, k, t, tau], rtol=rtol,
                                                   atol=1e-12)
                    else:
                        assert d["cross_cov"][o][k][t][tau] is None
    return d


def replay(g, n_cells, steps_T, dev):
    """Z tensors [C, n, Tsh, 2] per shrinking step, from the seeded legacy stream the reference run consumed."""
    from oracle import restate_planner as rpl
    rs, off, out = None, 0, []
    for Tsh in steps_T:
        Z, sums, rs = rpl.replay_noise(int(g["noise_seed"]), n_cells, Tsh, int(g["n_ideal"]), rs)
        np.testing.assert_array_equal(sums, g["noise_sums"][off:off + n_cells * Tsh])
        off += n_cells * Tsh
        out.append(torch.from_numpy(Z).to(dev))
    return out


# --------------------------------------------------------------------------- c1: three shrinking-horizon steps
def _c1(golden):
    g = golden("ref_planner_c1")
    K = [int(k) for k in g["s1_K"]]
    ovs, cells, yaws = make_ovehicles(g["pred"], g["z"], g["minpos"], g["pos_last"], K, g["bbox"])
    return g, K, ovs, cells, yaws


def test_c1_step1_matches_reference_methods(dev, golden, tmp_path):
    from ccmpc_b200 import midlevel
    g, K, ovs, _, _ = _c1(golden)
    T = 8
    mc = midlevel.MidlevelConstraints(T, ego_vehicle_id=0, out_dir=str(tmp_path), device=dev)
    built = mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs, 100), ovs, None, None, None, g["s1_eps_ura"], None, T, g["ref_traj"])
    hp = check_tuple(built, g, "s1_")
    np.testing.assert_allclose(mc.prob_lower_save, g["s1_prob_lower_save"], rtol=ROW_RTOL)      # v8ideal:961-962
    # (A, b) the reference saves (:931-939) and the stacked block are the same rows
    A, b = hp.A_b()
    blk = hp.block()
    Ax = blk["A_x"].reshape(-1, T, 4)
    rows = np.arange(Ax.shape[0])
    np.testing.assert_array_equal(Ax[rows, hp.t.reshape(-1), :2], A.reshape(-1, 2))
    np.testing.assert_array_equal(blk["b_x"], b.reshape(-1))
    assert np.count_nonzero(Ax) <= 2 * Ax.shape[0]
    # vertices are materialised lazily, same nesting [t][k][o] (v8ideal:631)
    np.testing.assert_allclose(built[1]()[0][0][0].cpu().numpy(), g["s1_vertices_t0k0o0"], rtol=1e-12, atol=1e-9)
    # the _moments pickle: reference layout, reference values
    check_moments_pickle(os.path.join(str(tmp_path), "agent0_frame100_moments"), g["s1_mom_mean"], g["s1_mom_cov"],
                         g["s1_mom_cross"])


@pytest.mark.parametrize("handoff", ["device_cache", "own_pickle", "reference_pickle"])
def test_c1_shrinking_horizon_steps_match_reference_methods(dev, golden, tmp_path, handoff):
    """Step 2 (frame 110, Tsh 7) takes the predict_ideal branch from the frame-100 moments - resident on the
    device, re-read from the pickle the repo wrote, or read from the pickle THE REFERENCE wrote - and step 3
    (frame 120, Tsh 6) continues from the moments of the ideal trajectories (v8ideal:824-825, 885-888, 960,
    2626-2635).  The 1e6-sample noise is the reference run's own (replayed from its seed)."""
    from ccmpc_b200 import midlevel
    g, K, ovs, _, _ = _c1(golden)
    T, C, n = 8, sum(K), int(g["n_ideal"])
    out_dir = str(tmp_path)
    mc = midlevel.MidlevelConstraints(T, ego_vehicle_id=0, out_dir=out_dir, device=dev)
    if handoff == "reference_pickle":
        with open(os.path.join(out_dir, "agent0_frame100_moments"), "wb") as f:
            f.write(g["s1_mom_pickle"].tobytes())
    else:
        mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
            params_of(ovs, 100), ovs, None, None, None, g["s1_eps_ura"], None, T, g["ref_traj"])
        if handoff == "own_pickle":
            mc = midlevel.MidlevelConstraints(T, ego_vehicle_id=0, out_dir=out_dir, device=dev)      # empty cache
    Z2, Z3 = replay(g, C, (T - 1, T - 2), dev)
    built2 = mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs, 110), ovs, None, None, None, g["s2_eps_ura"], None, T - 1, g["ref_traj"],
        ideal_noise=dict(x0=g["s2_x0"], z=Z2, n_samples=n))
    check_tuple(built2, g, "s2_")
    assert mc.prob_lower_save is None or handoff == "device_cache"             # only a T == ph step stores it (:961)
    check_moments_pickle(os.path.join(out_dir, "agent0_frame110_moments"), g["s2_mom_mean"], g["s2_mom_cov"],
                         g["s2_mom_cross"])
    del Z2
    built3 = mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs, 120), ovs, None, None, None, g["s3_eps_ura"], None, T - 2, g["ref_traj"],
        ideal_noise=dict(x0=g["s3_x0"], z=Z3, n_samples=n))
    check_tuple(built3, g, "s3_")
    check_moments_pickle(os.path.join(out_dir, "agent0_frame120_moments"), g["s3_mom_mean"], g["s3_mom_cov"],
                         g["s3_mom_cross"])
    # a step whose predecessor's moments are missing fails like the reference's failed load (:2628-2635)
    with pytest.raises(midlevel.InSimulationException):
        mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
            params_of(ovs, 500), ovs, None, None, None, g["s2_eps_ura"], None, T - 1, g["ref_traj"])


def test_c1_affine_builder_matches_reference_methods(dev, golden, tmp_path):
    from ccmpc_b200 import midlevel
    g, K, ovs, _, _ = _c1(golden)
    mc = midlevel.MidlevelConstraints(8, out_dir=str(tmp_path), device=dev)
    for pre, Tsh in (("af_", 8), ("af5_", 5)):
        built = mc.compute_obstacle_constraints_GMM_affine(params_of(ovs, 100), ovs, None, None, None,
                                                           g[pre + "eps_ura"], None, Tsh, g["ref_traj"])
        hp = check_tuple(built, g, pre, affine=True)
        assert built[4] is False                                                # out_junction = False (v8ideal:1424)
        A, b = hp.A_b()
        blk = hp.block()
        Ax = blk["A_x"].reshape(-1, Tsh, 4)
        np.testing.assert_array_equal(Ax[np.arange(Ax.shape[0]), hp.t.reshape(-1), :2], A.reshape(-1, 2))
        np.testing.assert_array_equal(blk["b_x"], b.reshape(-1))


def test_k_grows_between_steps_matches_reference_methods(dev, golden, tmp_path):
    """v8ideal:2652-2656: latent 2 of obstacle 0 did not exist at the previous step -> last saved mode."""
    from ccmpc_b200 import midlevel
    g = golden("ref_planner_kgrow")
    Ka, Kb = [int(k) for k in g["s1_K"]], [int(k) for k in g["s2_K"]]
    ovs_a, _, _ = make_ovehicles(g["pred"], g["z_a"], g["minpos"], g["pos_last"], Ka, g["bbox"])
    ovs_b, _, _ = make_ovehicles(g["pred"], g["z_b"], g["minpos"], g["pos_last"], Kb, g["bbox"])
    T, n = 8, int(g["n_ideal"])
    mc = midlevel.MidlevelConstraints(T, out_dir=str(tmp_path), device=dev)
    check_tuple(mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs_a, 100), ovs_a, None, None, None, g["s1_eps_ura"], None, T, g["ref_traj"]), g, "s1_")
    (Z,) = replay(g, sum(Kb), (T - 1,), dev)
    check_tuple(mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs_b, 110), ovs_b, None, None, None, g["s2_eps_ura"], None, T - 1, g["ref_traj"],
        ideal_noise=dict(x0=g["s2_x0"], z=Z, n_samples=n)), g, "s2_")
    check_moments_pickle(os.path.join(str(tmp_path), "agent0_frame110_moments"), g["s2_mom_mean"], g["s2_mom_cov"],
                         g["s2_mom_cross"])


# --------------------------------------------------------------------------- c2 (O=4, K=25, ph8, N=5000)
def test_c2_builders_match_reference_methods_and_block_handover(dev, golden, tmp_path):
    from ccmpc_b200 import midlevel
    g = golden("ref_planner_c2")
    K = [int(k) for k in g["s1_K"]]
    ovs, _, _ = make_ovehicles(g["pred"], g["z"], g["minpos"], g["pos_last"], K, g["bbox"])
    T = 8
    mc = midlevel.MidlevelConstraints(T, out_dir=str(tmp_path), device=dev)
    built = mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs, 100), ovs, None, None, None, g["s1_eps_ura"], None, T, g["ref_traj"])
    hp = check_tuple(built, g, "s1_")
    assert hp.n.shape[:2] == (100, 28)                                          # 2 800 rows
    check_moments_pickle(os.path.join(str(tmp_path), "agent0_frame100_moments"), g["s1_mom_mean"], g["s1_mom_cov"],
                         g["s1_mom_cross"])
    check_tuple(mc.compute_obstacle_constraints_GMM_affine(params_of(ovs, 100), ovs, None, None, None,
                                                           g["af_eps_ura"], None, T, g["ref_traj"]),
                g, "af_", affine=True)
    # one stacked block instead of 2 800 cvxpy rows; with the LTV model also in the decision variable u
    rng = np.random.default_rng(0)
    nx, nu = 4, 2
    Gamma = rng.normal(size=(nx * T, nu * T)); x_bar = rng.normal(size=nx * T) * 50; u_bar = rng.normal(size=nu * T)
    blk = hp.block(nx=nx, nu=nu, Gamma=Gamma, x_bar=x_bar, u_bar=u_bar)
    A, b = hp.A_b()
    A_x = np.zeros((2800, nx * T))
    r = np.arange(2800); t = hp.t.reshape(-1)
    A_x[r, nx * t] = A.reshape(-1, 2)[:, 0]
    A_x[r, nx * t + 1] = A.reshape(-1, 2)[:, 1]
    np.testing.assert_array_equal(blk["A_x"], A_x)
    np.testing.assert_array_equal(blk["b_x"], b.reshape(-1))
    np.testing.assert_allclose(blk["A_u"], A_x @ Gamma, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(blk["b_u"], b.reshape(-1) - A_x @ (x_bar - Gamma @ u_bar), rtol=1e-11, atol=1e-9)
    # any u: A_u u <= b_u  <=>  A_x x(u) <= b_x with x = Gamma (u - u_bar) + x_bar (v8ideal:2887)
    u = rng.normal(size=nu * T)
    x = Gamma @ (u - u_bar) + x_bar
    np.testing.assert_allclose(blk["A_u"] @ u - blk["b_u"], A_x @ x - b.reshape(-1), rtol=1e-9, atol=1e-8)


def test_c2_shrinking_step_matches_oracle(dev, golden, tmp_path):
    """c2 step 2 (100 cells through predict_ideal) against the restated driver with supplied noise
    (20 000 ideal samples per cell here; the 1e6-sample form of this branch is pinned by the c1 tests)."""
    from ccmpc_b200 import midlevel
    from oracle import restate_planner as rpl
    g = golden("ref_planner_c2")
    K = [int(k) for k in g["s1_K"]]
    ovs, cells, yaws = make_ovehicles(g["pred"], g["z"], g["minpos"], g["pos_last"], K, g["bbox"])
    T, n, C = 8, 20_000, sum(K)
    bb = [g["bbox"]] * len(ovs)
    mc = midlevel.MidlevelConstraints(T, out_dir=str(tmp_path), device=dev, n_ideal_samples=n)
    mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs, 100), ovs, None, None, None, g["s1_eps_ura"], None, T, g["ref_traj"])
    ref1 = rpl.minkowski_step(cells, yaws, bb, g["ref_traj"], g["s1_eps_ura"], T, T)
    rng = np.random.default_rng(5)
    Z = rng.standard_normal((C, n, T - 1, 2))
    x0 = np.stack([ref1["moments"]["mean_p0p1"][o][k][0] for o in range(len(K)) for k in range(K[o])]) + 0.05
    ref2 = rpl.minkowski_step(cells, yaws, bb, g["ref_traj"], g["s1_eps_ura"], T - 1, T,
                              moments_prev=ref1["moments"], x0=x0, Z=Z)
    built2 = mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs, 110), ovs, None, None, None, g["s1_eps_ura"], None, T - 1, g["ref_traj"],
        ideal_noise=dict(x0=x0, z=torch.from_numpy(Z).to(dev)))
    check_tuple(built2, ref2)
    mean, cov, cross = (np.full((4, 25, T - 1) + s, np.nan) for s in ((2,), (2, 2), (T - 2, 2, 2)))
    for o in range(4):
        for k in range(25):
            for t in range(T - 1):
                mean[o, k, t] = ref2["moments"]["mean_p0p1"][o][k][t]
                cov[o, k, t] = ref2["moments"]["cov_p0p1"][o][k][t]
                for tau in range(t):
                    cross[o, k, t, tau] = ref2["moments"]["cross_cov"][o][k][t][tau]
    check_moments_pickle(os.path.join(str(tmp_path), "agent0_frame110_moments"), mean, cov, cross)


def test_c1_at_100k_predictions_matches_oracle(dev, tmp_path):
    """c1 as tests/Hz20/params.py:377 actually runs it: n_predictions = 100 000 (the id says np5000)."""
    from ccmpc_b200 import midlevel
    from oracle import restate_planner as rpl, synth
    O, K, T, N = 1, 3, 8, 100_000
    sc = synth.make_scene(1, O, K, T)
    z, pos = synth.sample_scene(sc, N, 7)
    pred = (pos - synth.MINPOS).astype(np.float32)
    ovs, cells, yaws = make_ovehicles(pred, z, synth.MINPOS, sc.pos_last, [K], (3.7, 1.79))
    eps = midlevel.eps_allocation([K])
    mc = midlevel.MidlevelConstraints(T, out_dir=str(tmp_path), device=dev)
    built = mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs, 100), ovs, None, None, None, eps, None, T, sc.ref_traj)
    check_tuple(built, rpl.minkowski_step(cells, yaws, [np.array([3.7, 1.79])], sc.ref_traj, eps, T, T))
    built = mc.compute_obstacle_constraints_GMM_affine(params_of(ovs, 100), ovs, None, None, None, eps, None, T,
                                                       sc.ref_traj)
    check_tuple(built, rpl.affine_step(cells, yaws, [np.array([3.7, 1.79])], sc.ref_traj, eps, T, T), affine=True)


# --------------------------------------------------------------------------- error behaviour, modes
def test_shrinking_step_with_singular_saved_moments_raises_linalgerror(dev, golden):
    """predict_ideal inverts cov_p0p1[t] (v8ideal:2683): a saved covariance that is singular raises LinAlgError there.
    The fused rollout reports it as a per-cell status that the builder reads with its other results at the end of
    the step - the exception must still be the reference's."""
    from ccmpc_b200 import midlevel
    g, K, ovs, _, _ = _c1(golden)
    T = 8
    mc = midlevel.MidlevelConstraints(T, ego_vehicle_id=0, out_dir=None, device=dev, n_ideal_samples=4096)
    mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs, 100), ovs, None, None, None, g["s1_eps_ura"], None, T, g["ref_traj"])
    mc._moments[100].cov[1, 2].zero_()                      # cell 1, timestep 2: Sigma_t = 0
    with pytest.raises(np.linalg.LinAlgError):
        mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
            params_of(ovs, 110), ovs, None, None, None, g["s1_eps_ura"], None, T - 1, g["ref_traj"])
    # the stand-alone method raises at once
    with pytest.raises(np.linalg.LinAlgError):
        mc._predict_ideal_fused(midlevel.CellBatch.from_ovehicles(ovs, dev), T - 1, params_of(ovs, 110), g["ref_traj"])


def test_degenerate_cells_raise_what_the_reference_raises(dev, golden, tmp_path):
    from ccmpc_b200 import midlevel
    g, K, ovs, _, _ = _c1(golden)
    T = 8
    mc = midlevel.MidlevelConstraints(T, out_dir=None, device=dev)
    call = lambda o, ref=g["ref_traj"]: mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(o, 100), o, None, None, None, None, None, T, ref)
    # all samples of one mode identical: Sigma_tau singular -> np.linalg.inv raises (makeconstraint:65)
    dup = [FakeOVehicle([ovs[0].pred_positions[0], np.repeat(ovs[0].pred_positions[1][:1], 40, axis=0)],
                        [ovs[0].pred_yaws[0], np.repeat(ovs[0].pred_yaws[1][:1], 40, axis=0)], g["bbox"])]
    with pytest.raises(np.linalg.LinAlgError):
        call(dup)
    # a single sample: np.cov is NaN (ddof = 1) and scipy.linalg.solve rejects it (ValueError; LinAlgError is one)
    one = [FakeOVehicle([ovs[0].pred_positions[0][:1]], [ovs[0].pred_yaws[0][:1]], g["bbox"])]
    with pytest.raises(ValueError):
        call(one)
    with pytest.raises(ValueError):
        mc.compute_obstacle_constraints_GMM_affine(params_of(one, 100), one, None, None, None, None, None, T,
                                                   g["ref_traj"])
    # reference point at the obstacle mean's y: infinite slope (v8ideal:923)
    good = [FakeOVehicle(ovs[0].pred_positions[:1], ovs[0].pred_yaws[:1], g["bbox"])]
    hp = call(good)[0]
    ref = np.array(g["ref_traj"], dtype=np.float64)
    ref[3, 1] = hp.extra["mean"][0, 3, 1]
    with pytest.raises(midlevel.InSimulationException):
        call(good, ref)
    # rows with a status never reach the solver hand-over either
    bad = midlevel.HalfPlanes(n=np.zeros((1, 2)), d=np.zeros(1), rhs=np.zeros(1), idx=np.zeros(1, np.int32),
                              side=np.ones(1, np.int32), t=np.zeros(1, np.int32), status=np.array([2], np.uint8),
                              cell_obstacle=np.zeros(1, int), cell_mode=np.zeros(1, int), dev={})
    with pytest.raises(TypeError):
        bad.block()


def test_accel_modes_through_the_planner_api(dev, golden):
    from ccmpc_b200 import midlevel
    g, K, ovs, _, _ = _c1(golden)
    for bad in (3, -1, 1.5, "2"):
        with pytest.raises((ValueError, TypeError)):
            midlevel.MidlevelConstraints(8, out_dir=None, device=dev, accel=bad)
    for accel in (1, 2):
        mc = midlevel.MidlevelConstraints(8, out_dir=None, device=dev, accel=accel)
        assert mc.accel == accel
        built = mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
            params_of(ovs, 100), ovs, None, None, None, g["s1_eps_ura"], None, 8, g["ref_traj"])
        check_tuple(built, g, "s1_", rtol=1e-6)             # direct solve: beta within ~1e-8, contract 1e-5
    plan = midlevel.StepPlan.from_samples(g["z"], g["pred"].astype(np.float64) + g["minpos"], g["pos_last"], 3,
                                          prediction_horizon=8, device=dev, accel=2)
    assert plan.accel == 2
    res = plan.run(ref_traj=g["ref_traj"])
    np.testing.assert_allclose(res["nd"].reshape(-1, 3)[:, 2], g["s1_rhs"], rtol=1e-6)
    np.testing.assert_array_equal(np.where(res["idx_side"].reshape(-1) & 2, -1, 1), g["s1_side"])


# --------------------------------------------------------------------------- MC risk estimate (stage 1)
def _risk_inputs():
    from oracle import synth
    O, K, T = 2, 2, 8
    sc = synth.make_scene(3, O, K, T + 1)
    mean, cov, cross = (a.reshape((O * K,) + a.shape[2:]) for a in sc.moments())
    ego = sc.ref_traj[:T] + np.array([0.0, 2.0])
    return O, K, T, mean, cov, cross, ego


def _moments_obj(midlevel, mean, cov, cross, O, K, dev):
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return midlevel.Moments(t(mean), t(cov), t(cross), np.repeat(np.arange(O), K), np.tile(np.arange(K), O),
                            [K] * O, mean.shape[1])


def test_mc_collision_risk_matches_oracle_with_supplied_noise(dev):
    from ccmpc_b200 import midlevel
    from oracle import restate
    O, K, T, mean, cov, cross, ego = _risk_inputs()
    n = 4099
    rng = np.random.default_rng(9)
    Z = rng.standard_normal((O * K, n, T, 2))
    x0 = mean[:, 0] + 0.1
    mc = midlevel.MidlevelConstraints(T + 1, out_dir=None, device=dev)
    res = mc.mc_collision_risk(_moments_obj(midlevel, mean, cov, cross, O, K, dev), [K] * O, T, ego, n_samples=n,
                               x0=x0, z=torch.from_numpy(Z).to(dev))
    counts = res["counts"].cpu().numpy()
    traj = res["cells"].samples.reshape(O * K, n, T, 2).cpu().numpy()
    total = 0
    for c in range(O * K):
        ref = restate.ideal_rollout(mean[c], cov[c], cross[c], T, Z[c], x0[c])
        np.testing.assert_allclose(traj[c], ref, rtol=1e-12, atol=1e-10)
        coll, any_, _, _ = restate.collision_counts(traj[c], ego)           # counts of the SAME tensor: bit-exact
        np.testing.assert_array_equal(counts[c, :T], coll)
        assert counts[c, T] == any_
        total += int(coll.sum())
        m, cv, _ = restate.cell_moments(traj[c])
        np.testing.assert_allclose(res["moments"]["mean"][c].cpu().numpy(), m, rtol=1e-10)
        np.testing.assert_allclose(res["moments"]["cov"][c].cpu().numpy(), cv, rtol=1e-8, atol=1e-12)
    assert total > 0, "the test scene must produce collisions"


def _risk_worker(rank, world, port, q):
    """Two ranks share cuda:0 (gloo reduces the host copy of the packed sums): the sharded estimate must equal
    the unsharded one bit for bit in the counts - Philox counters are global sample indices."""
    import torch.distributed as dist
    from ccmpc_b200 import midlevel
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        dev = torch.device("cuda:0")
        torch.cuda.set_device(dev)
        O, K, T, mean, cov, cross, ego = _risk_inputs()
        n = 30_001                                  # odd: shards of different sizes
        mc = midlevel.MidlevelConstraints(T + 1, out_dir=None, device=dev)
        mom = _moments_obj(midlevel, mean, cov, cross, O, K, dev)
        sharded = mc.mc_collision_risk(mom, [K] * O, T, ego, n_samples=n, seed=77, process_group=dist.group.WORLD)
        whole = mc.mc_collision_risk(mom, [K] * O, T, ego, n_samples=n, seed=77)
        ok = bool(torch.equal(sharded["counts"], whole["counts"]) and int(whole["counts"].sum()) > 0
                  and int(sharded["sums"][0, 0].item()) == n
                  and sharded["cells"].N == (O * K) * (n // 2 + (1 if rank == 0 else 0))
                  and torch.allclose(sharded["moments"]["cov"], whole["moments"]["cov"], rtol=1e-9, atol=1e-12))
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def test_mc_collision_risk_two_ranks_equals_one(dev):
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_risk_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]


def test_shrinking_step_with_philox_noise_is_consistent(dev, golden, tmp_path):
    """Without supplied noise the shrinking-horizon branch runs predict_ideal, the trajectory store and the sums in
    ONE kernel (ccmpc_mc_rollout_sums at T = 7): the rows it returns must be the rows of the ideal trajectories it
    stored - rebuilt here from those trajectories with the two-kernel path and with the oracle."""
    from ccmpc_b200 import midlevel, ops
    from oracle import restate
    g, K, ovs, _, _ = _c1(golden)
    T, n = 8, 50_000
    mc = midlevel.MidlevelConstraints(T, out_dir=str(tmp_path), device=dev, n_ideal_samples=n, seed=3)
    mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs, 100), ovs, None, None, None, g["s1_eps_ura"], None, T, g["ref_traj"])
    l0 = ops._lib.load().ccmpc_launch_count()
    built = mc.compute_obstacle_constraints_GMM_Minkowski_idealprediction(
        params_of(ovs, 110), ovs, None, None, None, g["s2_eps_ura"], None, T - 1, g["ref_traj"])
    hp, work = built[0], mc.last_ideal
    assert work.T == T - 1 and work.N == sum(K) * n and not hp.status.any()
    # the same rows from the stored trajectories through sample_sums (other summation order: 1e-9)
    out2, _, _ = mc.minkowski(work, g["ref_traj"], g["s2_eps_ura"])
    h2 = {k: v.cpu().numpy() for k, v in out2.items()}
    np.testing.assert_array_equal(hp.idx, h2["idx_side"] & 1)
    np.testing.assert_array_equal(hp.side, np.where(h2["idx_side"] & 2, -1, 1))
    np.testing.assert_allclose(hp.n, h2["nd"][..., :2], rtol=1e-8)
    np.testing.assert_allclose(hp.rhs, h2["nd"][..., 2], rtol=1e-8)
    # ... and through the oracle on one cell
    traj = work.samples.reshape(sum(K), n, T - 1, 2)[1].cpu().numpy()
    ref = restate.minkowski_cell(traj, g["ref_traj"], g["s2_eps_ura"][0, 1], T)
    np.testing.assert_array_equal(hp.side[1], ref["side"])
    np.testing.assert_allclose(hp.rhs[1], ref["d"], rtol=1e-7)
    # the moments it saved for the next step are the moments of those trajectories
    with open(os.path.join(str(tmp_path), "agent0_frame110_moments"), "rb") as f:
        d = pickle.load(f)
    m, cv, cr = restate.cell_moments(traj)
    np.testing.assert_allclose(np.stack(d["mean_p0p1"][0][1]), m, rtol=1e-10)
    np.testing.assert_allclose(np.stack(d["cov_p0p1"][0][1]), cv, rtol=1e-8, atol=1e-12)
    np.testing.assert_allclose(d["cross_cov"][0][1][5][2], cr[5, 2], rtol=1e-8, atol=1e-12)
    del l0
