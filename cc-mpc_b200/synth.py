"""Seeded synthetic inputs of the BASELINE.json config shapes.

BENCH / TEST INPUT GENERATORS (no reference arithmetic, no oracle code).  Generates the *inputs*
both the CUDA path and the CPU oracle consume - synthetic GMM predictions in
Town03 scene-4 coordinates, their samples, a reference trajectory, closed-form
moments for the rollout, and random SPD covariances for the batched MVOE
sweep.  No reference arithmetic lives here.

Config seeds: c1 = 1, c2 = 2, c3 = 3, c4 = 4, c5 = 5.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

DT = 0.5                      # tests/conftest.py:37 (prediction step, s)
GOAL = np.array([167.174698, -81.759842])   # dataForCost_ref.py:209
AR1 = 0.9                     # cross-time correlation of the synthetic modes
MINPOS = np.array([100.0, -150.0])          # synthetic scene.x_min / y_min (v8ideal:476)


@dataclass
class GmmScene:
    """Per-obstacle synthetic GMM prediction over T timesteps."""
    weights: np.ndarray     # (O, K)
    mean: np.ndarray        # (O, K, T, 2)
    chol: np.ndarray        # (O, K, T, 2, 2) lower Cholesky factor of Sigma_t
    ref_traj: np.ndarray    # (T, 2)
    pos_last: np.ndarray    # (O, 2) last observed position of each obstacle
    T: int

    @property
    def O(self):
        return self.weights.shape[0]

    @property
    def K(self):
        return self.weights.shape[1]

    def cov(self):
        """Sigma_t per (o,k,t): (O,K,T,2,2)."""
        return self.chol @ np.swapaxes(self.chol, -1, -2)

    def cross(self, t, tau):
        """Cov(x_t, x_tau) per (o,k): S_t (AR1^|t-tau|) S_tau^T."""
        return (AR1 ** abs(t - tau)) * (self.chol[:, :, t] @ np.swapaxes(self.chol[:, :, tau], -1, -2))

    def moments(self):
        """Closed-form (mean (O,K,T,2), cov (O,K,T,2,2), cross (O,K,T,T,2,2))."""
        O, K, T = self.O, self.K, self.T
        cross = np.zeros((O, K, T, T, 2, 2))
        for t in range(T):
            for tau in range(T):
                cross[:, :, t, tau] = self.cross(t, tau)
        return self.mean.copy(), self.cov(), cross


def reference_trajectory(T, start=(176.0, -73.5), speed=6.0):
    """Straight line from ``start`` toward the goal at ``speed`` m/s (T,2)."""
    start = np.asarray(start, dtype=np.float64)
    u = (GOAL - start) / np.linalg.norm(GOAL - start)
    steps = np.arange(1, T + 1)[:, None] * (speed * DT)
    return start[None, :] + steps * u[None, :]


def make_scene(seed, O, K, T, flat=True, margin=0.5):
    """Synthetic GMM predictions: constant-turn-rate mode means from
    p0 in U([150,200] x [-100,-60]), speed U(3,10) m/s, sigma growing linearly
    from 0.1 to 3 m, rho in U(-0.8, 0.8).  Mode means keep |ref_y - mu_y| >=
    ``margin`` at every timestep (resampled otherwise) so that the tangent
    slope and best_idx of the reference are well conditioned."""
    rng = np.random.default_rng(seed)
    ref = reference_trajectory(T)
    weights = np.full((O, K), 1.0 / K) if flat else rng.dirichlet(np.ones(K), size=O)
    mean = np.zeros((O, K, T, 2))
    chol = np.zeros((O, K, T, 2, 2))
    pos_last = np.zeros((O, 2))
    tt = np.arange(1, T + 1) * DT
    for o in range(O):
        p0 = np.array([rng.uniform(150, 200), rng.uniform(-100, -60)])
        pos_last[o] = p0
        for k in range(K):
            for _ in range(1000):
                v = rng.uniform(3, 10)
                hdg = rng.uniform(-np.pi, np.pi)
                omega = rng.uniform(-0.3, 0.3)
                th = hdg + omega * tt
                # constant-turn-rate path (closed form, small-omega safe)
                if abs(omega) > 1e-3:
                    px = p0[0] + v / omega * (np.sin(th) - np.sin(hdg))
                    py = p0[1] - v / omega * (np.cos(th) - np.cos(hdg))
                else:
                    px = p0[0] + v * tt * np.cos(hdg)
                    py = p0[1] + v * tt * np.sin(hdg)
                if np.all(np.abs(ref[:, 1] - py) >= margin + 0.25):
                    break
            else:
                raise RuntimeError("could not place a mode with the requested margin")
            mean[o, k, :, 0] = px
            mean[o, k, :, 1] = py
            sx = np.linspace(0.1, 3.0, T) * rng.uniform(0.7, 1.0)
            sy = np.linspace(0.1, 3.0, T) * rng.uniform(0.7, 1.0)
            rho = rng.uniform(-0.8, 0.8)
            chol[o, k, :, 0, 0] = sx
            chol[o, k, :, 1, 0] = rho * sy
            chol[o, k, :, 1, 1] = sy * np.sqrt(1 - rho * rho)
    return GmmScene(weights, mean, chol, ref, pos_last, T)


def sample_scene(scene: GmmScene, N, seed, through_float32=True):
    """Draw N samples per obstacle: z ~ Cat(w), positions ~ N(mu_k, Sigma_k)
    with AR(1) noise across time.  Positions are rounded through float32 in
    scene-local coordinates and shifted back by ``MINPOS`` in float64, the way
    the reference receives them (prediction.py:104 float32; v8ideal:486
    ``predictions[idx] + minpos``).
    Returns z (O,N) int64 and positions (O,N,T,2) float64."""
    rng = np.random.default_rng(seed)
    O, K, T = scene.O, scene.K, scene.T
    z = np.stack([rng.choice(K, size=N, p=scene.weights[o]) for o in range(O)])
    w = rng.standard_normal((O, N, T, 2))
    e = np.empty_like(w)
    e[:, :, 0] = w[:, :, 0]
    s = np.sqrt(1 - AR1 * AR1)
    for t in range(1, T):
        e[:, :, t] = AR1 * e[:, :, t - 1] + s * w[:, :, t]
    pos = np.empty((O, N, T, 2))
    for o in range(O):
        L = scene.chol[o][z[o]]            # (N,T,2,2)
        pos[o] = scene.mean[o][z[o]] + np.einsum("ntij,ntj->nti", L, e[o])
    if through_float32:
        pos = (pos - MINPOS).astype(np.float32).astype(np.float64) + MINPOS
    return z, pos


def cells_from_samples(z, pos, K):
    """Mode-sorted CSR layout of the samples (stable in sample order, the order
    v8ideal:491-492 appends them).  Returns
      flat (sum N, T, 2), offsets (C+1,) int64, cell_obstacle (C,), cell_mode (C,)
    keeping only non-empty (o,k) cells, obstacle-major."""
    O = z.shape[0]
    chunks, offs, co, ck = [], [0], [], []
    for o in range(O):
        for k in range(K):
            sel = np.nonzero(z[o] == k)[0]
            if sel.size == 0:
                continue
            chunks.append(pos[o, sel])
            offs.append(offs[-1] + sel.size)
            co.append(o)
            ck.append(k)
    flat = np.concatenate(chunks, axis=0)
    return flat, np.asarray(offs, np.int64), np.asarray(co, np.int32), np.asarray(ck, np.int32)


def random_spd(rng, shape, smin=0.1, smax=3.0):
    """Random symmetric 2x2 covariances as (xx, xy, yy) triples, shape + (3,)."""
    sx = rng.uniform(smin, smax, size=shape)
    sy = rng.uniform(smin, smax, size=shape)
    rho = rng.uniform(-0.8, 0.8, size=shape)
    return np.stack([sx * sx, rho * sx * sy, sy * sy], axis=-1)


def sym_to_full(s3):
    """(…,3) (xx,xy,yy) -> (…,2,2)."""
    s3 = np.asarray(s3)
    out = np.empty(s3.shape[:-1] + (2, 2))
    out[..., 0, 0] = s3[..., 0]
    out[..., 0, 1] = s3[..., 1]
    out[..., 1, 0] = s3[..., 1]
    out[..., 1, 1] = s3[..., 2]
    return out


def mvoe_sweep_inputs(seed, S, O, T, K):
    """c4: one mode covariance per (scenario, obstacle, timestep, mode),
    (S,O,T,K,3) float64 symmetric triples."""
    rng = np.random.default_rng(seed)
    return random_spd(rng, (S, O, T, K))


def ego_plan(scene: GmmScene):
    """Ego plan = the reference trajectory."""
    return scene.ref_traj.copy()


def gmm2d_params(scene: GmmScene, Z=25, slots=None, local=True):
    """The scene as the predictor would hand it over (prediction.py:76-86): per vehicle a latent pmf over Z
    discrete states and, per (latent, timestep), ONE-component GMM2D parameters in float32 (Trajectron++ emits
    float32; vehicles use GMM_components = 1).  The scene's K modes occupy latent slots ``slots`` (default: the
    first K), every other latent has pmf 0.  ``local``: means in scene-local coordinates (minus MINPOS), the
    way predictions arrive before v8ideal:486 adds minpos.
    -> latent_pmf (O,Z) f64, mus (O,Z,T,1,2) f32, log_sigmas (O,Z,T,1,2) f32, corrs (O,Z,T,1) f32."""
    O, K, T = scene.O, scene.K, scene.T
    slots = np.arange(K) if slots is None else np.asarray(slots)
    pmf = np.zeros((O, Z))
    mus = np.zeros((O, Z, T, 1, 2), np.float32)
    log_sigmas = np.zeros((O, Z, T, 1, 2), np.float32)
    corrs = np.zeros((O, Z, T, 1), np.float32)
    sx = scene.chol[..., 0, 0]
    sy = np.sqrt(scene.chol[..., 1, 0] ** 2 + scene.chol[..., 1, 1] ** 2)
    for j, sl in enumerate(slots):
        pmf[:, sl] = scene.weights[:, j]
        mus[:, sl, :, 0, :] = scene.mean[:, j] - (MINPOS if local else 0.0)
        log_sigmas[:, sl, :, 0, 0] = np.log(sx[:, j])
        log_sigmas[:, sl, :, 0, 1] = np.log(sy[:, j])
        corrs[:, sl, :, 0] = scene.chol[:, j, :, 1, 0] / sy[:, j]
    return pmf, mus, log_sigmas, corrs


def unicycle_control_params(V, Z, T, seed, G=1, dt=0.5):
    """Synthetic predictor output in CONTROL space, the way Trajectron++ emits it for vehicles
    (dynamic = Unicycle): per vehicle a latent pmf over Z states (Dirichlet(0.3), 1-5 states above the 0.1
    filter of ovehicle.py:58) and, per (latent, timestep), G-component GMM2D parameters over
    u = (heading rate [rad/s], acceleration [m/s^2]); each latent is a manoeuvre: heading rate in
    U(-0.35, 0.35) with every fifth latent driving straight (|dphi| <= 1e-2, the small-angle branch),
    acceleration in U(-1.5, 1.5), standard deviations 0.02-0.08 rad/s and 0.1-0.5 m/s^2.  Initial states
    x0 = (x, y, phi, v): local-frame positions in [0,50] x [0,40], speed U(3,10) m/s (max_v = 10, v8ideal:90).
    -> latent_pmf (V,Z) f64, log_pis (V,Z,T,G) f32 or None, mus, log_sigmas (V,Z,T,G,2) f32, corrs (V,Z,T,G) f32,
       x0 (V,4) f32."""
    rng = np.random.default_rng(seed)
    pmf = rng.dirichlet(np.full(Z, 0.3), size=V)
    turn = rng.uniform(-0.35, 0.35, (V, Z, 1, G))
    turn[:, ::5] = rng.uniform(-8e-3, 8e-3, turn[:, ::5].shape)
    acc = rng.uniform(-1.5, 1.5, (V, Z, 1, G))
    mus = np.zeros((V, Z, T, G, 2), np.float32)
    mus[..., 0] = turn + rng.normal(0, 0.01, (V, Z, T, G))
    mus[:, ::5, :, :, 0] = turn[:, ::5]
    mus[..., 1] = acc + rng.normal(0, 0.1, (V, Z, T, G))
    log_sigmas = np.zeros((V, Z, T, G, 2), np.float32)
    log_sigmas[..., 0] = np.log(rng.uniform(0.02, 0.08, (V, Z, T, G)))
    log_sigmas[:, ::5, :, :, 0] = np.log(1e-4)          # straight manoeuvres stay inside the small-angle branch
    log_sigmas[..., 1] = np.log(rng.uniform(0.1, 0.5, (V, Z, T, G)))
    corrs = rng.uniform(-0.5, 0.5, (V, Z, T, G)).astype(np.float32)
    log_pis = rng.normal(size=(V, Z, T, G)).astype(np.float32) if G > 1 else None
    x0 = np.stack([rng.uniform(0, 50, V), rng.uniform(0, 40, V), rng.uniform(-np.pi, np.pi, V),
                   rng.uniform(3, 10, V)], axis=1).astype(np.float32)
    return pmf, log_pis, mus, log_sigmas, corrs, x0
