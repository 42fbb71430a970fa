#!/usr/bin/env python
"""Tiny drivers for `ncu --set full` captures (one kernel, a few launches, no timing).

  python tools/prof_driver.py sweep [accel]     # c4 MVOE sweep, accel = 0 | 1 | 2
  python tools/prof_driver.py fit   [accel]     # c4-sized stand-alone footprint fit
  python tools/prof_driver.py sums | risk | rollout | fused | c5s   # c3 / c5 kernels
Run WITHOUT ncu first (must exit 0), then under
  ncu --set full --clock-control none --import-source on -k regex:<kernel> -s 2 -c 1 -o gpurun_out/<name> python tools/prof_driver.py ...
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccmpc_b200 import ops  # noqa: E402
from oracle import synth  # noqa: E402  (generators only)


def c4_inputs(dev, S=4096, O=8, T=16, K=25, seed=4):
    g = torch.Generator(device=dev).manual_seed(seed)
    u = lambda lo, hi: torch.rand((S, O, T, K), generator=g, device=dev, dtype=torch.float64) * (hi - lo) + lo
    sx, sy, rho = u(0.1, 3.0), u(0.1, 3.0), u(-0.8, 0.8)
    cov = torch.stack([sx * sx, rho * sx * sy, sy * sy], dim=-1).contiguous()
    mu = torch.stack([u(150, 200), u(-100, -60)], dim=-1).contiguous()
    ref = torch.from_numpy(np.broadcast_to(synth.reference_trajectory(T), (S, T, 2)).copy()).to(dev)
    ref[..., 1] += 30.0
    return cov, mu, ref, -2.0 * np.log(0.05 / O / T)


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "sweep"
    accel = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    dev = torch.device("cuda:0")
    if what == "c5s":
        from ccmpc_b200 import midlevel
        O, K, T, N, Z = 8, 4, 16, 20_000, 25
        sc = synth.make_scene(5, O, K, T)
        pmf, mus, ls, rho = synth.gmm2d_params(sc, Z=Z)
        plan = midlevel.StepPlan.from_gmm(O, N, T, sc.pos_last, prediction_horizon=T, Z=Z, minpos=synth.MINPOS,
                                          device=dev, builder="affine", seed=5, use_graph=False)
        for i in range(3):
            plan.run_gmm(pmf, mus, ls, rho, ref_traj=sc.ref_traj, step_seed=i)
        print("ok c5s")
        return
    if what == "sums":
        O, K, T, n = 8, 4, 16, 250_000
        C = O * K
        sc = synth.make_scene(3, O, K, T + 1)
        m, c, x = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in sc.moments())
        m, c, x = m.reshape(C, T + 1, 2), c.reshape(C, T + 1, 2, 2), x.reshape(C, T + 1, T + 1, 2, 2)
        ego = torch.from_numpy(synth.reference_trajectory(T)).to(dev)
        traj = torch.empty((C, n, T, 2), dtype=torch.float64, device=dev)
        ops.mc_rollout(m, c, x, None, n, T, seed=1, out=traj)
        flat = traj.view(C * n, T, 2)
        offsets = torch.arange(C + 1, dtype=torch.int64, device=dev) * n
        pv = m[:, 1:T + 1].contiguous()
        for i in range(4):
            ops.sample_sums(flat, offsets, pv, ego=ego)
        torch.cuda.synchronize()
        print("ok sums")
        return
    if what == "fused":
        O, K, T, n = 8, 4, 16, 250_000
        C = O * K
        sc = synth.make_scene(3, O, K, T + 1)
        m, c, x = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in sc.moments())
        m, c, x = m.reshape(C, T + 1, 2), c.reshape(C, T + 1, 2, 2), x.reshape(C, T + 1, T + 1, 2, 2)
        ego = torch.from_numpy(synth.reference_trajectory(T)).to(dev)
        traj = torch.empty((C, n, T, 2), dtype=torch.float64, device=dev)
        pv = m[:, 1:T + 1].contiguous()
        ws = ops.mc_rollout_sums_workspace(C, n, T, dev)
        for i in range(4):
            ops.mc_rollout_sums(m, c, x, None, n, T, pv, ego, seed=i, workspace=ws, out=traj)
        torch.cuda.synchronize()
        print("ok fused")
        return
    if what in ("risk", "rollout"):
        O, K, T, n = 8, 4, 16, 250_000
        C = O * K
        sc = synth.make_scene(3, O, K, T + 1)
        m, c, x = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in sc.moments())
        m, c, x = m.reshape(C, T + 1, 2), c.reshape(C, T + 1, 2, 2), x.reshape(C, T + 1, T + 1, 2, 2)
        ego = torch.from_numpy(synth.reference_trajectory(T)).to(dev)
        traj = torch.empty((C, n, T, 2), dtype=torch.float64, device=dev) if what == "rollout" else None
        for i in range(4):
            if what == "risk":
                ops.mc_risk(m, c, x, None, n, T, ego, seed=i)
            else:
                ops.mc_rollout(m, c, x, None, n, T, seed=i, out=traj)
        torch.cuda.synchronize()
        print("ok", what)
        return
    cov, mu, ref, chi = c4_inputs(dev)
    out = None
    for _ in range(4):
        if what == "sweep":
            out = ops.mvoe_halfplane_sweep(cov, mu, ref, chi, accel=accel, out=out)
        else:
            ops.mvoe_fit(cov.reshape(-1, 3), None, footprint_r2=3.4 ** 2, accel=accel, want_iters=False)
    torch.cuda.synchronize()
    print("ok", what, accel)


if __name__ == "__main__":
    main()
