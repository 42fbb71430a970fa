"""Generate tests/golden/ref_*.npz by running THE REFERENCE'S OWN functions.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Run in the build container
(where /root/reference exists):

    python -m oracle.make_golden

The reference functions are imported by file path (oracle/ref_loader.py) from
  collect/in_simulation/midlevel/v8ideal/makeconstraint.py   (:7-303)
  python-utility/utility/utility/npu.py                      (:308-350)
and called on seeded inputs; inputs and outputs are stored side by side so the
tests can replay them against (i) the NumPy restatement and (ii) the CUDA path
without the reference being present (it is absent on the GPU box).

The "pipeline" goldens drive the restated loop of v8ideal:893-947 / :1476-1515
with the reference's functions plugged in (restate.minkowski_cell(impl=ref)).
The loop itself cannot be imported (v8ideal/__init__.py imports carla, cvxpy,
docplex), so those files pin the per-pair arithmetic, not the loop driver.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import scipy.stats

from . import ref_loader, restate, synth

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _spd_pairs(rng, M):
    s1 = synth.sym_to_full(synth.random_spd(rng, (M,), 0.05, 4.0))
    s2 = synth.sym_to_full(synth.random_spd(rng, (M,), 0.05, 4.0))
    # hand-picked cases: KAT of SURVEY §8c, footprint fits, extreme ratios, equal shapes
    extra1 = [np.eye(2), np.array([[2, .5], [.5, 1.]]), np.eye(2) * 1e-4, np.eye(2) * 50.0,
              np.array([[1.0, 0.999], [0.999, 1.0]]), np.array([[3.0, -1.2], [-1.2, 0.7]])]
    extra2 = [4 * np.eye(2), np.array([[1, -.2], [-.2, 3.]]), np.eye(2) * 3.4 ** 2, np.eye(2) * 3.4 ** 2,
              np.eye(2) * 3.4 ** 2, np.array([[3.0, -1.2], [-1.2, 0.7]])]
    s1 = np.concatenate([np.stack(extra1), s1])
    s2 = np.concatenate([np.stack(extra2), s2])
    # second half: footprint fits Sigma2 = R^2 I  (v8ideal:917-918)
    h = s1.shape[0] // 2
    s2[h:] = np.eye(2) * restate.R_EV_OV ** 2
    return s1, s2


def gen_mvoe(mc, rng):
    s1, s2 = _spd_pairs(rng, 506)
    beta = np.zeros(len(s1))
    Q = np.zeros_like(s1)
    for i in range(len(s1)):
        beta[i], Q[i] = mc.compute_mvoe(s1[i], s2[i])
    np.savez(os.path.join(OUT, "ref_mvoe.npz"), Sigma1=s1, Sigma2=s2, beta=beta, Q=Q)


def gen_predict_moments(mc, rng):
    M, N = 64, 96
    base = rng.standard_normal((M, 4, 4))
    rows = np.einsum("mij,mjn->min", base, rng.standard_normal((M, 4, N)))
    rows += rng.uniform(-200, 200, size=(M, 4, 1))     # Town03-sized offsets
    ci = np.zeros((M, 2, 2)); cm = np.zeros((M, 2, 2)); ct = np.zeros((M, 2, 2))
    for i in range(M):
        ci[i], cm[i], ct[i] = mc.predict_moments([rows[i, 0], rows[i, 1], rows[i, 2], rows[i, 3]])
    np.savez(os.path.join(OUT, "ref_predict_moments.npz"), rows=rows, cov_infer=ci, cov_mu=cm, cov_t=ct)


def gen_tangent(mc, rng):
    M = 256
    mu = rng.uniform(-50, 250, size=(M, 2))
    Sig = synth.sym_to_full(synth.random_spd(rng, (M,), 0.3, 6.0))
    c = rng.choice([1.0, 3.4, 2.0], size=M)
    a = mu + rng.uniform(-30, 30, size=(M, 2))
    a[:, 1] += np.where(np.abs(a[:, 1] - mu[:, 1]) < 0.5, 1.0, 0.0)
    m = -(a[:, 0] - mu[:, 0]) / (a[:, 1] - mu[:, 1])
    # KAT rows of SURVEY §8c
    mu[0], Sig[0], c[0], m[0], a[0] = [1, 2], [[2, .5], [.5, 1]], 1.0, 0.5, [10, 3]
    mu[1], Sig[1], c[1], m[1], a[1] = [1, 2], [[2, .5], [.5, 1]], 1.0, 0.5, [10, 30]
    n = np.zeros((M, 2)); d1 = np.zeros(M); d2 = np.zeros(M)
    ds = np.zeros(M); idx = np.zeros(M, np.int32); dist = np.zeros((M, 2)); forced = np.zeros((M, 2))
    for i in range(M):
        (n1, e1), (n2, e2) = mc.tangent_lines_of_slope_m(mu[i], Sig[i], c[i], m[i])
        n[i], d1[i], d2[i] = n1, e1, e2
        dist[i] = [mc.distance_point_to_line(n1, e1, a[i]), mc.distance_point_to_line(n2, e2, a[i])]
        ns, dd, ii = mc.choose_closest_tangent(mu[i], Sig[i], c[i], m[i], a[i])
        ds[i], idx[i] = dd, ii
        forced[i] = [mc.choose_closest_tangent(mu[i], Sig[i], c[i], m[i], a[i], const_idx=j)[1] for j in (0, 1)]
    np.savez(os.path.join(OUT, "ref_tangent.npz"), mu=mu, Sigma=Sig, c=c, m=m, a=a,
             n=n, d1=d1, d2=d2, dist=dist, d_star=ds, idx=idx, d_forced=forced)


def gen_scale_lower(mc, rng):
    M = 128
    ci = synth.sym_to_full(synth.random_spd(rng, (M,), 0.1, 2.0))
    cm = synth.sym_to_full(synth.random_spd(rng, (M,), 0.1, 2.0))
    ci[0] = [[.5, .1], [.1, .4]]; cm[0] = [[.3, .05], [.05, .2]]
    ct = ci + cm
    eps = rng.choice([0.05 / 8, 0.05 / 4 / 8, 0.05 / 8 / 16, 0.0125 / 8], size=M)
    eps[0] = 0.05 / 8
    Gam = scipy.stats.norm.ppf(1 - eps)
    scale = np.array([mc.compute_scale(ci[i], cm[i], ct[i], Gam[i]) for i in range(M)])
    lower = np.array([mc.compute_lower_bound(ci[i], cm[i], ct[i], eps[i]) for i in range(M)])
    np.savez(os.path.join(OUT, "ref_scale_lower.npz"), cov_infer=ci, cov_mu=cm, cov_t=ct,
             eps=eps, Gamma=Gam, scale=scale, lower=lower)


def gen_vertices(npu, rng):
    N = 300
    centers = rng.uniform(-200, 200, size=(N, 2))
    theta = rng.uniform(-np.pi, np.pi, size=N)
    lw = np.array([3.7, 1.79])
    v = npu.vertices_of_bboxes(centers, theta, lw)
    v_scalar = npu.vertices_of_bboxes(centers[:7], 0.3, [4.5, 2.5])
    np.savez(os.path.join(OUT, "ref_vertices.npz"), centers=centers, theta=theta, lw=lw, vertices=v,
             vertices_scalar_theta=v_scalar)


def gen_scalars():
    eps = np.array([0.05 / 8, 0.05 / 4 / 8, 0.05 / 8 / 16, 0.05 / 8 / 8, 0.05 / 3 / 8, 1e-3, 1e-6])
    np.savez(os.path.join(OUT, "ref_scalars.npz"), eps=eps,
             chi2_ppf_1m_eps=scipy.stats.chi2.ppf(1 - eps, df=2),
             chi2_ppf_target=scipy.stats.chi2.ppf(0.9999, df=2),
             norm_ppf_1m_eps=scipy.stats.norm.ppf(1 - eps),
             chi2_cdf_x=np.array([0.01, 0.5, 2.0, 9.0, 30.0]),
             chi2_cdf=scipy.stats.chi2.cdf(np.array([0.01, 0.5, 2.0, 9.0, 30.0]), df=2))


def gen_pipeline(mc):
    """Small scene through the restated loops with the reference's functions."""
    O, K, T, N = 2, 3, 6, 240
    scene = synth.make_scene(11, O, K, T, flat=True)
    z, pos = synth.sample_scene(scene, N, 12)
    flat, offs, co, ck = synth.cells_from_samples(z, pos, K)
    eps_ura = restate.eps_allocation([K] * O)
    keys = ["n", "d", "idx", "side", "A", "b", "Q", "Q1", "beta1", "beta2", "cov_infer", "cov_mu",
            "prob_pair", "prob_lower", "mean"]
    mk = {k: [] for k in keys}
    af = {k: [] for k in ["n", "d", "idx", "side", "offset", "rhs", "mean", "cov"]}
    for c in range(len(co)):
        cell = flat[offs[c]:offs[c + 1]]
        r = restate.minkowski_cell(cell, scene.ref_traj, eps_ura[co[c], ck[c]], T, impl=mc)
        for k in keys:
            mk[k].append(r[k])
        r = restate.affine_cell(cell, scene.ref_traj, eps_ura[co[c], ck[c]], T, impl=mc)
        for k in af:
            af[k].append(r[k])
    np.savez(os.path.join(OUT, "ref_pipeline.npz"), samples=flat, offsets=offs, cell_obstacle=co, cell_mode=ck,
             ref_traj=scene.ref_traj, eps_ura=eps_ura, T=T,
             **{"mk_" + k: np.stack(v) for k, v in mk.items()},
             **{"af_" + k: np.stack(v) for k, v in af.items()})


def main():
    if not ref_loader.available():
        sys.exit("reference not found under %s" % ref_loader.REFERENCE_ROOT)
    os.makedirs(OUT, exist_ok=True)
    mc, npu = ref_loader.makeconstraint(), ref_loader.npu()
    rng = np.random.default_rng(20261104)
    gen_mvoe(mc, rng)
    gen_predict_moments(mc, rng)
    gen_tangent(mc, rng)
    gen_scale_lower(mc, rng)
    gen_vertices(npu, rng)
    gen_scalars()
    gen_pipeline(mc)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
