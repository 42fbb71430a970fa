"""Parity at BASELINE.json's FULL sizes through size-independent properties (the oracle cannot finish these
sizes in seconds, so the checks are identities the reference's arithmetic must satisfy, evaluated with plain
torch fp64 elementwise ops on the device as the checker - never as the product).

c3  1M samples / obstacle, ph16: rollout -> sums -> moments
c4  4096 x 8 x 16 x 25 MVOE footprint fits + tangent half-planes
c5  8 obstacles x 20 000 samples, ph16: one MPC step through the ingest boundary
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from ccmpc_b200 import _lib
    _lib.load(build_if_missing=False)
    return torch.device("cuda:0")


def T_(a, dev, dtype=torch.float64):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dtype).to(dev)


# ------------------------------------------------------------------------------------------ c3
def test_c3_fullsize_rollout_counts_and_moments(dev):
    from ccmpc_b200 import ops
    from oracle import synth
    O, K, T, n = 8, 4, 16, 250_000
    C = O * K
    sc = synth.make_scene(3, O, K, T + 1)
    mean_h, cov_h, cross_h = sc.moments()
    mean, cov, cross = (T_(mean_h.reshape(C, T + 1, 2), dev), T_(cov_h.reshape(C, T + 1, 2, 2), dev),
                        T_(cross_h.reshape(C, T + 1, T + 1, 2, 2), dev))
    ego = T_(synth.reference_trajectory(T), dev)
    pivot = mean[:, 1:T + 1].contiguous()
    traj, _, status, _ = ops.mc_rollout(mean, cov, cross, None, n, T, seed=7)
    assert not status.any()
    flat = traj.view(C * n, T, 2)
    offsets = torch.arange(C + 1, dtype=torch.int64, device=dev) * n
    sums = ops.sample_sums(flat, offsets, pivot, ego=ego)
    L = ops.sums_layout(T)
    assert torch.equal(sums[:, 0], torch.full((C,), float(n), dtype=torch.float64, device=dev))

    # (1) collision counts, bit-exact against the A9 predicate written as separate (unfused) torch fp64 ops
    R2 = torch.tensor(ops.R_EV_OV, dtype=torch.float64, device=dev) ** 2
    coll_ref = torch.empty((C, T), dtype=torch.int64, device=dev)
    any_ref = torch.empty((C,), dtype=torch.int64, device=dev)
    for c in range(C):                                      # per cell: bounded temporary memory
        dx = ego[None, :, 0] - traj[c, :, :, 0]
        dy = ego[None, :, 1] - traj[c, :, :, 1]
        hit = (dx * dx + dy * dy) < R2
        coll_ref[c] = hit.sum(dim=0)
        any_ref[c] = hit.any(dim=1).sum()
    counts = sums[:, L["coll"]:L["coll"] + T + 1]
    assert torch.equal(counts[:, :T].to(torch.int64), coll_ref), "per-timestep collision counts differ"
    assert torch.equal(counts[:, T].to(torch.int64), any_ref), "any-time collision counts differ"
    assert torch.equal(counts, counts.round()), "counts must be exact integers in the fp64 sums"

    # (2) linearity over sample shards (what the multi-GPU all-reduce relies on): two halves add up to the whole;
    #     counts bit-exact, power sums to fp64 round-off
    half = n // 2
    sums_a = torch.zeros_like(sums)
    for lo, hi in ((0, half), (half, n)):
        sel = traj[:, lo:hi].contiguous().view(-1, T, 2)
        o2 = torch.arange(C + 1, dtype=torch.int64, device=dev) * (hi - lo)
        sums_a += ops.sample_sums(sel, o2, pivot, ego=ego)
    assert torch.equal(sums_a[:, L["coll"]:L["coll"] + T + 1], counts)
    scale = sums.abs().clamp_min(1.0)
    assert float(((sums_a - sums).abs() / scale)[:, :L["coll"]].max()) < 1e-11

    # (3) moments: mean / cov / cross against torch on one cell, and the chain property of predict_ideal
    #     (v8ideal:2684-2703): sample moments of the rollout reproduce the conditional-Gaussian chain within
    #     CLT bounds at n = 250 000
    mom = ops.moments_from_sums(sums, pivot, want_yaw=False)
    for c in (0, C - 1):
        X = traj[c].reshape(n, 2 * T)
        m = X.mean(dim=0)
        Xc = X - m
        S = (Xc.T @ Xc) / (n - 1)
        assert float((mom["mean"][c].reshape(-1) - m).abs().max()) < 1e-9
        for t in range(T):
            blk = S[2 * t:2 * t + 2, 2 * t:2 * t + 2]
            assert float((mom["cov"][c, t] - blk).abs().max()) <= 1e-9 * float(blk.abs().max())
            for tau in range(t):
                blk = S[2 * t:2 * t + 2, 2 * tau:2 * tau + 2]
                assert float((mom["cross"][c, t, tau] - blk).abs().max()) <= 1e-9 * float(S.abs().max())
    # one-step conditional structure: regressing x_{t+1} on x_t recovers A_t = C_{t+1,t} Sigma_t^-1 of the
    # INPUT moments (the broadcast x0 draw makes marginals differ from the input, the conditionals do not)
    c = 1
    for t in (3, 10):
        St = mom["cov"][c, t].cpu().numpy()
        Ct = mom["cross"][c, t + 1, t].cpu().numpy()
        A_hat = Ct @ np.linalg.inv(St)
        Cin = cross_h.reshape(C, T + 1, T + 1, 2, 2)[c, t + 2, t + 1]
        Sin = cov_h.reshape(C, T + 1, 2, 2)[c, t + 1]
        A_in = Cin @ np.linalg.inv(Sin)
        assert np.abs(A_hat - A_in).max() < 0.02 * max(1.0, np.abs(A_in).max())


# ------------------------------------------------------------------------------------------ c4
def _c4_inputs(dev, S=4096, O=8, T=16, K=25, seed=4):
    from oracle import synth
    g = torch.Generator(device=dev).manual_seed(seed)
    u = lambda lo, hi: torch.rand((S, O, T, K), generator=g, device=dev, dtype=torch.float64) * (hi - lo) + lo
    sx, sy, rho = u(0.1, 3.0), u(0.1, 3.0), u(-0.8, 0.8)
    cov = torch.stack([sx * sx, rho * sx * sy, sy * sy], dim=-1).contiguous()
    mu = torch.stack([u(150, 200), u(-100, -60)], dim=-1).contiguous()
    ref = torch.from_numpy(np.broadcast_to(synth.reference_trajectory(T), (S, T, 2)).copy()).to(dev)
    ref[..., 1] += 30.0
    return cov, mu, ref, -2.0 * np.log(0.05 / O / T)


def test_c4_fullsize_sweep_properties(dev):
    """13.1 M fits: (i) beta satisfies the reference's stopping rule |F(beta_prev) - beta_prev| < 1e-8, checked as
    a fixed-point residual; (ii) Q is the reference's formula for that beta (makeconstraint:35); (iii) the emitted
    line is tangent to the Q ellipse (makeconstraint:144-159); (iv) best_idx / side obey the closed forms of
    SURVEY A7/A8; (v) the direct solve and the reference sequence agree to 1e-7 everywhere."""
    from ccmpc_b200 import ops
    cov, mu, ref, chi = _c4_inputs(dev)
    S, O, T, K = cov.shape[:4]
    R2 = ops.R_EV_OV ** 2
    res = {}
    for accel in (0, 2):
        q, nd, idx_side, status = ops.mvoe_halfplane_sweep(cov, mu, ref, chi, accel=accel)
        assert int(status.max()) == 0
        res[accel] = (q.clone(), nd.clone(), idx_side.clone())
    for accel, (q, nd, idx_side) in res.items():
        a = cov * chi
        # recover beta from Q's off-diagonal / trace: Q = (1 + 1/b) a + (1 + b) R2 I
        tr_a = a[..., 0] + a[..., 2]
        tr_q = q[..., 0] + q[..., 2]
        # tr_q = (1 + 1/b) tr_a + 2 (1 + b) R2  ->  2 R2 b^2 + (tr_a + 2 R2 - tr_q) b + tr_a = 0 (two positive
        # roots with product tr_a / (2 R2)); the off-diagonal (1 + 1/b) a_xy picks the right one
        inv_b = q[..., 1] / a[..., 1] - 1.0
        b = 1.0 / inv_b
        assert float(((1.0 + inv_b) * tr_a + 2.0 * (1.0 + b) * R2 - tr_q).abs().max() / float(tr_q.max())) < 1e-9
        # (i) fixed-point residual with lam = eig(a^-1 R2 I): sum 1/(1+b lam) / sum lam/(1+b lam) = b^2
        det_a = a[..., 0] * a[..., 2] - a[..., 1] * a[..., 1]
        trl = R2 * tr_a / det_a
        dtl = R2 * R2 / det_a
        Fb = torch.sqrt((2.0 + b * trl) / (trl + 2.0 * b * dtl))
        # beta is the LAST iterate; the previous one differs by < 1e-8, and F contracts, so |F(b) - b| < 1e-8 too
        assert float((Fb - b).abs().max()) < 1e-8
        # (iii) tangency: (d - n.mu)^2 = n^T Q n with n = (n0, 1)
        n0, d = nd[..., 0], nd[..., 2]
        assert torch.equal(nd[..., 1], torch.ones_like(n0))
        nQn = n0 * n0 * q[..., 0] + 2.0 * n0 * q[..., 1] + q[..., 2]
        nmu = n0 * mu[..., 0] + mu[..., 1]
        assert float((((d - nmu) ** 2 - nQn).abs() / nQn).max()) < 1e-9
        # slope (v8ideal:923): n0 = (ref_x - mu_x) / (ref_y - mu_y)
        rx = ref[:, None, :, None, 0].expand(S, O, T, K)
        ry = ref[:, None, :, None, 1].expand(S, O, T, K)
        assert float((n0 - (rx - mu[..., 0]) / (ry - mu[..., 1])).abs().max()) < 1e-9
        # (iv) best_idx = 0 iff ref_y > mu_y; side bit set iff n.mu > d, i.e. iff best_idx = 1
        idx = (idx_side & 1).bool()
        assert torch.equal(idx, ~(ry > mu[..., 1]))
        assert torch.equal((idx_side & 2).bool(), nmu > d)
    # (v) the two iterate modes against each other
    (q0, nd0, is0), (q2, nd2, is2) = res[0], res[2]
    assert torch.equal(is0, is2)
    assert float(((q0 - q2).abs() / q0.abs().clamp_min(1e-3)).max()) < 1e-7
    assert torch.equal(nd0[..., :2], nd2[..., :2])                       # the normal does not depend on beta
    # d = n.mu +- sqrt(n^T Q n): only the half-width depends on beta, so that is the scale of the comparison
    half = torch.sqrt(nd0[..., 0] ** 2 * q0[..., 0] + 2.0 * nd0[..., 0] * q0[..., 1] + q0[..., 2])
    assert float(((nd0[..., 2] - nd2[..., 2]).abs() / half).max()) < 1e-7


# ------------------------------------------------------------------------------------------ c5
def test_c5_fullsize_step_against_oracle(dev):
    """One c5 MPC step (O=8, 20 000 samples / obstacle, ph16) from the predictor's float32 outputs through device
    grouping: the oracle (NumPy) still finishes this size in seconds, so counts / idx / side are checked
    bit-exactly and the affine rows to 1e-9."""
    from ccmpc_b200 import midlevel
    from oracle import restate, synth
    O, K, T, N = 8, 4, 16, 20_000
    sc = synth.make_scene(5, O, K, T)
    z, pos = synth.sample_scene(sc, N, 5)
    plan = midlevel.StepPlan.from_samples(z, pos, sc.pos_last, K, prediction_horizon=T, device=dev, builder="affine")
    res = plan.run(ref_traj=sc.ref_traj)
    eps = restate.eps_allocation(plan.n_states)
    flat = plan.host_samples.numpy()
    for c in range(plan.C):
        cell = flat[plan.offsets_host[c]:plan.offsets_host[c + 1]]
        ref = restate.affine_cell(cell, sc.ref_traj, eps[plan.cell_obstacle[c], plan.cell_mode[c]], T)
        assert np.array_equal(res["aff_idx_side"][c] & 1, ref["idx"])
        assert np.array_equal(np.where(res["aff_idx_side"][c] & 2, -1, 1), ref["side"])
        assert np.allclose(res["aff_offset_rhs"][c][:, 1], ref["rhs"], rtol=1e-9, atol=0)
        coll, any_, _, _ = restate.collision_counts(cell, sc.ref_traj)
        assert np.array_equal(res["counts"][c][:T], coll) and res["counts"][c][T] == any_


# ------------------------------------------------------------------------------------------ ragged one-wave sums
@pytest.mark.parametrize("T", [16, 8, 5])
def test_sample_sums_one_wave_ragged_cells(dev, T):
    """Large input over few ragged cells (empty, 1 row, sizes that are not multiples of 4): sample_sums runs as ONE
    wave of balanced units.  Counts are bit-exact against plain torch fp64 elementwise ops, every sum matches a
    per-cell launch (which takes the small-input slab path) to round-off, and the row counts add up."""
    from ccmpc_b200 import ops
    sizes = [0, 1, 123_457, 700_001, 3, 250_000, 99_999, 0, 31]
    C, N = len(sizes), sum(sizes)
    g = torch.Generator(device=dev).manual_seed(17 + T)
    centre = torch.rand((C, 1, T, 2), generator=g, device=dev, dtype=torch.float64) * 40 + 150
    offsets = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]), dtype=torch.int64, device=dev)
    samples = torch.empty((N, T, 2), dtype=torch.float64, device=dev)
    for c, n in enumerate(sizes):
        if n:
            samples[offsets[c]:offsets[c + 1]] = centre[c] + torch.randn((n, T, 2), generator=g, device=dev,
                                                                         dtype=torch.float64) * 2.0
    pivot = ops.cell_pivot(samples, offsets)
    ego = (centre[:, 0] + 2.5).contiguous()                          # per-cell plan about 3.5 m from the centre
    sums = ops.sample_sums(samples, offsets, pivot, ego=ego, R=3.4)
    L = ops.sums_layout(T) if hasattr(ops, "sums_layout") else None
    s = sums.cpu().numpy()
    np.testing.assert_array_equal(s[:, 0], np.array(sizes, dtype=np.float64))
    for c, n in enumerate(sizes):
        if not n:
            assert not s[c].any()
            continue
        x = samples[offsets[c]:offsets[c + 1]]
        one = ops.sample_sums(x.contiguous(), torch.tensor([0, n], dtype=torch.int64, device=dev),
                              pivot[c:c + 1].contiguous(), ego=ego[c:c + 1].contiguous(), R=3.4).cpu().numpy()[0]
        scale = np.maximum(np.abs(one), 1.0)
        assert np.max(np.abs(s[c] - one) / scale) < 1e-10, c
        # A9 counts, same operation order as the oracle (no FMA): torch elementwise fp64
        dx, dy = ego[c, :, 0] - x[:, :, 0], ego[c, :, 1] - x[:, :, 1]
        hit = (dx * dx + dy * dy) < 3.4 * 3.4
        coll = hit.sum(dim=0).cpu().numpy().astype(np.float64)
        anyc = float(hit.any(dim=1).sum())
        if L is not None:
            np.testing.assert_array_equal(s[c, L["coll"]:L["coll"] + T], coll)
            assert s[c, L["coll"] + T] == anyc
        else:
            m = ops.moments_from_sums(sums, pivot)
            np.testing.assert_array_equal(m["counts"][c, :T].cpu().numpy().astype(np.float64), coll)
            assert float(m["counts"][c, T]) == anyc
        assert coll.sum() > 0 or n < 100
