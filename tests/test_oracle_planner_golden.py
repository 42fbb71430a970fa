"""The restated loop drivers, L4, save_moments, predict_ideal, mode grouping and cost_final against goldens
produced by THE REFERENCE'S OWN CODE (oracle/make_golden_extracted.py, oracle/make_golden_planner.py: the
reference functions / MidlevelAgent methods compiled unmodified from /root/reference).  CPU only."""
import os
import pickle

import numpy as np
import pytest

from oracle import restate, restate_planner as rpl

MINK_KEYS = ("n", "t", "side", "rhs")


def cells_of(pred, z, minpos, pos_last, K_per_obstacle):
    pos = pred.astype(np.float64) + minpos
    cells = [[pos[o][z[o] == k] for k in range(K_per_obstacle[o])] for o in range(pos.shape[0])]
    yaws = [[restate.yaws_from_positions(p, pos_last[o]) for p in cells[o]] for o in range(len(cells))]
    return cells, yaws


def nested_from_dense(mean, cov, cross):
    """Dense golden moments -> the nested lists save_moments pickles (None where NaN)."""
    O, Kmax, T = mean.shape[:3]
    d = {"mean_p0p1": np.empty((O, Kmax, T), dtype=object).tolist(),
         "cov_p0p1": np.empty((O, Kmax, T), dtype=object).tolist(),
         "cross_cov": np.empty((O, Kmax, T, max(T - 1, 0)), dtype=object).tolist()}
    for o in range(O):
        for k in range(Kmax):
            for t in range(T):
                if np.isnan(mean[o, k, t, 0]):
                    continue
                d["mean_p0p1"][o][k][t] = mean[o, k, t]
                d["cov_p0p1"][o][k][t] = cov[o, k, t]
                for tau in range(t):
                    d["cross_cov"][o][k][t][tau] = cross[o, k, t, tau]
    return d


def check_step(g, pre, out, rtol=1e-9, mom_rtol=1e-9):
    np.testing.assert_array_equal(out["t"], g[pre + "t"])
    np.testing.assert_array_equal(out["side"], g[pre + "side"])
    np.testing.assert_allclose(out["n"], g[pre + "n"], rtol=rtol, atol=0)
    np.testing.assert_allclose(out["rhs"], g[pre + "rhs"], rtol=rtol, atol=0)
    np.testing.assert_allclose(out["A_union"], g[pre + "A_union"], rtol=1e-12, atol=1e-12, equal_nan=True)
    np.testing.assert_allclose(out["b_union"], g[pre + "b_union"], rtol=1e-12, atol=1e-10, equal_nan=True)
    np.testing.assert_allclose(out["ov_mean"], g[pre + "ov_mean"], rtol=1e-13, equal_nan=True)
    np.testing.assert_allclose(out["ov_cov"], g[pre + "ov_cov"], rtol=1e-11, equal_nan=True)
    assert bool(out["OVconstraint"]) == bool(g[pre + "OVconstraint"])
    if "prob_lower_save" in out:
        np.testing.assert_allclose(out["prob_lower_save"], g[pre + "prob_lower_save"], rtol=rtol)
        mean, cov, cross = g[pre + "mom_mean"], g[pre + "mom_cov"], g[pre + "mom_cross"]
        ref = nested_from_dense(mean, cov, cross)
        for key in ("mean_p0p1", "cov_p0p1", "cross_cov"):
            _same_nested(out["moments"][key], ref[key], mom_rtol)


def _same_nested(a, b, rtol):
    if isinstance(b, list):
        assert isinstance(a, list) and len(a) == len(b)
        for x, y in zip(a, b):
            _same_nested(x, y, rtol)
    elif b is None:
        assert a is None
    else:
        np.testing.assert_allclose(a, b, rtol=rtol, atol=1e-12)


def test_minkowski_and_affine_step_match_reference_c1_step1(golden):
    g = golden("ref_planner_c1")
    K = [int(k) for k in g["s1_K"]]
    cells, yaws = cells_of(g["pred"], g["z"], g["minpos"], g["pos_last"], K)
    T = cells[0][0].shape[1]
    bb = [g["bbox"]] * len(cells)
    out = rpl.minkowski_step(cells, yaws, bb, g["ref_traj"], g["s1_eps_ura"], T, T)
    check_step(g, "s1_", out)
    assert g["s1_direct_all_none"] and int(g["s1_last"]) == 0
    for pre, Tsh in (("af_", T), ("af5_", 5)):
        check_step(g, pre, rpl.affine_step(cells, yaws, bb, g["ref_traj"], g[pre + "eps_ura"], Tsh, T))
        assert not bool(g[pre + "OVconstraint"])          # out_junction = False is hard-coded (v8ideal:1424)


def test_shrinking_horizon_steps_match_reference_c1(golden):
    """Steps 2 and 3 of the shrinking horizon: predict_ideal from the previous step's pickle with the replayed
    1e6-sample noise, then the Minkowski loop over the ideal trajectories, then save_moments of those."""
    g = golden("ref_planner_c1")
    K = [int(k) for k in g["s1_K"]]
    cells, yaws = cells_of(g["pred"], g["z"], g["minpos"], g["pos_last"], K)
    T, C, n = cells[0][0].shape[1], sum(K), int(g["n_ideal"])
    bb = [g["bbox"]] * len(cells)
    moments = pickle.loads(g["s1_mom_pickle"].tobytes())        # the file the REFERENCE wrote at frame 100
    rs, off = None, 0
    for pre, Tsh in (("s2_", T - 1), ("s3_", T - 2)):
        Z, sums, rs = rpl.replay_noise(int(g["noise_seed"]), C, Tsh, n, rs)
        np.testing.assert_array_equal(sums, g["noise_sums"][off:off + C * Tsh])      # same stream as the reference saw
        off += C * Tsh
        out = rpl.minkowski_step(cells, yaws, bb, g["ref_traj"], g[pre + "eps_ura"], Tsh, T, moments_prev=moments,
                                 x0=g[pre + "x0"], Z=Z)
        check_step(g, pre, out, rtol=1e-8, mom_rtol=1e-8)
        moments = out["moments"]


def test_k_grows_between_steps_matches_reference(golden):
    """v8ideal:2652-2656: a latent index beyond the saved modes falls back to the last saved one."""
    g = golden("ref_planner_kgrow")
    Ka, Kb = [int(k) for k in g["s1_K"]], [int(k) for k in g["s2_K"]]
    assert Ka == [2, 2] and Kb == [3, 2]
    cells_a, yaws_a = cells_of(g["pred"], g["z_a"], g["minpos"], g["pos_last"], Ka)
    cells_b, yaws_b = cells_of(g["pred"], g["z_b"], g["minpos"], g["pos_last"], Kb)
    T, n = cells_a[0][0].shape[1], int(g["n_ideal"])
    bb = [g["bbox"]] * 2
    out1 = rpl.minkowski_step(cells_a, yaws_a, bb, g["ref_traj"], g["s1_eps_ura"], T, T)
    check_step(g, "s1_", out1)
    Z, sums, _ = rpl.replay_noise(int(g["noise_seed"]), sum(Kb), T - 1, n)
    np.testing.assert_array_equal(sums, g["noise_sums"])
    out2 = rpl.minkowski_step(cells_b, yaws_b, bb, g["ref_traj"], g["s2_eps_ura"], T - 1, T,
                              moments_prev=out1["moments"], x0=g["s2_x0"], Z=Z)
    check_step(g, "s2_", out2, rtol=1e-8, mom_rtol=1e-8)


def test_minkowski_and_affine_step_match_reference_c2(golden):
    g = golden("ref_planner_c2")
    K = [int(k) for k in g["s1_K"]]
    assert K == [25] * 4 and g["s1_n"].shape[0] == 2800
    cells, yaws = cells_of(g["pred"], g["z"], g["minpos"], g["pos_last"], K)
    bb = [g["bbox"]] * 4
    check_step(g, "s1_", rpl.minkowski_step(cells, yaws, bb, g["ref_traj"], g["s1_eps_ura"], 8, 8))
    check_step(g, "af_", rpl.affine_step(cells, yaws, bb, g["ref_traj"], g["af_eps_ura"], 8, 8))


def test_l4_matches_reference(golden):
    g = golden("ref_l4")
    for i in range(len(g["theta"])):
        verts = restate.box_vertices(g["centers"][i], g["yaw"][i], g["lw"])
        At, b0 = restate.l4_outer(g["theta"][i], verts)
        np.testing.assert_array_equal(At, g["At"][i])
        np.testing.assert_allclose(b0, g["b0"][i], rtol=1e-13, atol=1e-12)


def test_cell_moments_match_reference_save_moments(golden):
    g = golden("ref_moments")
    offs = g["offsets"]
    for c in range(len(offs) - 1):
        mean, cov, cross = restate.cell_moments(g["samples"][offs[c]:offs[c + 1]])
        np.testing.assert_array_equal(mean, g["mean"][c])
        np.testing.assert_array_equal(cov, g["cov"][c])
        np.testing.assert_array_equal(cross, g["cross"][c])


def test_ideal_rollout_matches_reference_predict_ideal(golden):
    g = golden("ref_ideal_rollout")
    Tsh = int(g["Tsh"])
    for c in range(g["x0"].shape[0]):
        for zk, tk in (("Z_head", "traj_head"), ("Z_tail", "traj_tail")):
            ref = restate.ideal_rollout(g["mean"][c], g["cov"][c], g["cross"][c], Tsh, g[zk][c], g["x0"][c])
            np.testing.assert_allclose(ref, g[tk][c], rtol=1e-14, atol=1e-12)


def test_group_by_mode_matches_reference_from_trajectron(golden):
    g = golden("ref_grouping")
    for ci in range(int(g["n_cases"])):
        p = lambda k: g[f"c{ci}_{k}"]
        T = p("pred").shape[1]
        pos_l, yaw_l, centers, pmf, _ = restate.group_by_mode(p("pred").astype(np.float64) + p("minpos"), p("z"),
                                                              p("pmf"), p("pos_last"), T, float(p("filter")))
        assert len(pos_l) == int(p("K"))
        np.testing.assert_array_equal(np.cumsum([0] + [x.shape[0] for x in pos_l]), p("offsets"))
        np.testing.assert_array_equal(np.concatenate(pos_l), p("positions"))
        np.testing.assert_array_equal(np.concatenate(yaw_l), p("yaws"))
        np.testing.assert_array_equal(centers, p("init_center"))
        np.testing.assert_array_equal(pmf, p("latent_pmf"))


def test_episode_cost_matches_reference_cost_final(golden):
    g = golden("ref_cost")
    for e in range(len(g["lens"])):
        L = int(g["lens"][e])
        cost, costU = restate.episode_cost(list(g["X"][e, :L]), list(g["U"][e, :L]))
        assert cost == pytest.approx(g["cost"][e], rel=1e-14)
        assert costU == pytest.approx(g["costU"][e], rel=1e-14)


def test_reference_predict_ideal_reads_a_pickle_written_by_the_repo(tmp_path, monkeypatch):
    """The other direction of the `_moments` hand-off, with the reference's OWN reader: ``Moments.nested()`` (the
    product's writer, run here on CPU tensors) -> pickle -> the reference's unmodified ``predict_ideal``
    (v8ideal:2620-2711, compiled out of /root/reference by oracle/ref_loader.extract) indexes it as
    [o][k][t+1] / cross_cov[o][k][t+1][t] and rolls out; the rows it produces equal the restated rollout on the
    same (replayed) noise.  Needs /root/reference: runs in the build container, skipped on the GPU box."""
    import torch
    from oracle import ref_loader
    if not ref_loader.available():
        pytest.skip("/root/reference is not present here")
    from ccmpc_b200 import midlevel, synth
    from oracle import ref_planner as rp
    O, K, T = 2, 2, 5
    sc = synth.make_scene(7, O, K, T)
    z, pos = synth.sample_scene(sc, 300, 7)
    flat, offs, co, ck = synth.cells_from_samples(z, pos, K)
    C = len(co)
    mean = np.zeros((C, T, 2)); cov = np.zeros((C, T, 2, 2)); cross = np.zeros((C, T, T, 2, 2))
    for c in range(C):
        mean[c], cov[c], cross[c] = restate.cell_moments(flat[offs[c]:offs[c + 1]])
    mom = midlevel.Moments(torch.from_numpy(mean), torch.from_numpy(cov), torch.from_numpy(cross), np.asarray(co),
                           np.asarray(ck), [K] * O, T)
    monkeypatch.chdir(tmp_path)
    os.makedirs("out/data")
    with open("out/data/agent0_frame40_moments", "wb") as f:
        pickle.dump(mom.nested(), f)
    ovs = [rp.FakeOVehicle([np.zeros((1, T, 2))] * K, [np.zeros((1, T))] * K, (3.7, 1.79)) for _ in range(O)]
    agent = rp.make_agent(T, 0, seed=11)
    Tsh = T - 1
    traj_all = agent.predict_ideal(ovs, Tsh, 0, rp.make_params(ovs, 50))          # reads frame 50 - 10
    x0 = np.concatenate(agent.rnd.x0_draws, axis=0)
    rs = np.random.RandomState(11)
    keep = 64
    for c in range(C):
        rs.standard_normal((1, 2))
        Z = np.stack([rs.standard_normal((1_000_000, 2))[:keep] for _ in range(Tsh)], axis=1)
        ref = restate.ideal_rollout(mean[c], cov[c], cross[c], Tsh, Z, x0[c])
        np.testing.assert_allclose(traj_all[co[c]][ck[c]][:keep], ref, rtol=1e-14, atol=1e-12)
