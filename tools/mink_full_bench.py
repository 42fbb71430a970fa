"""c4 in its FULL Minkowski form: every (scenario, obstacle, mode) cell runs the whole (t, tau) loop of
v8ideal:893-947 (120 pairs at ph16, two MVOE fits per pair): 4096 x 8 x 25 cells = 1.97e8 fits per launch."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccmpc_b200 import ops  # noqa: E402
from oracle import synth  # noqa: E402  (generators only)


def main():
    dev = torch.device("cuda:0")
    S, O, K, T = (int(sys.argv[1]) if len(sys.argv) > 1 else 4096), 8, 25, 16
    C = S * O * K
    # 256 distinct cells (512 samples each) from one synthetic scene, tiled over the scenarios
    n = 512
    sc = synth.make_scene(4, 8, 4, T)
    _, pos = synth.sample_scene(sc, n * 32, 4)                # (8, n * 32, T, 2)
    pos = pos.reshape(8 * 32, n, T, 2)
    base_c = pos.shape[0]
    flat = torch.from_numpy(pos.reshape(-1, T, 2)).to(dev)
    offs = torch.arange(base_c + 1, dtype=torch.int64, device=dev) * n
    piv = ops.cell_pivot(flat, offs)
    sums_b = ops.sample_sums(flat, offs, piv)
    rep = (C + base_c - 1) // base_c
    sums = sums_b.repeat(rep, 1)[:C].contiguous()
    pivot = piv.repeat(rep, 1, 1)[:C].contiguous()
    ref = torch.from_numpy(synth.reference_trajectory(T)).to(dev).contiguous()
    eps = torch.full((C,), 0.05 / O, dtype=torch.float64, device=dev)
    P = T * (T - 1) // 2
    print(f"cells {C}, pairs {C * P}, fits {2 * C * P}, sums {sums.numel() * 8 / 1e9:.2f} GB")
    for accel in (0, 2):
        out = ops.minkowski_alloc(C, T, dev, fields=("nd", "ab", "idx_side", "prob_lower", "status"))
        for _ in range(2):
            ops.minkowski_halfplanes(sums, pivot, ref, eps, T, accel=accel, out=out)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            ops.minkowski_halfplanes(sums, pivot, ref, eps, T, accel=accel, out=out)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        bad = int((out["status"] != 0).sum())
        byt = sums.numel() * 8 + sum(v.numel() * v.element_size() for v in out.values())
        print(f"accel={accel}: {ms:.3f} ms  {2 * C * P / ms / 1e6:.1f} Gfit/s  {byt / ms / 1e6:.0f} GB/s  failed rows {bad}")
        del out


if __name__ == "__main__":
    main()
