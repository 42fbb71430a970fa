// rollout.cuh - per-cell step parameters of the conditional-Gaussian chain of predict_ideal
// (v8ideal:2662-2709), shared by rollout.cu (materialising rollout) and risk.cu (fused rollout + evaluation).
#pragma once
#include "common.cuh"
#include "philox.cuh"

namespace ccmpc {

// per (cell, t): x' = mu_{t+1} + A (x - mu_t) + L z  (v8ideal:2688-2703) with the constant part folded,
//   b = mu_{t+1} - A mu_t,   x' = b + A x + L z
// so one step is 7 dependent-depth-2 FMAs instead of 11 FP64 instructions of depth 5 (the chain over t is the
// serial part of every rollout kernel and shares the FP64 pipe with the Gram update of the fused kernel).
// 9 doubles + pad -> five 16-byte shared-memory loads.  |b| <= ~|mu| (1 + |A|): the fold costs a few ulp of the
// position (1e-13 m at Town03 coordinates), far inside the oracle tolerance (tests: 1e-13 relative).
struct __align__(16) StepParams {
    double bx, by, a00, a01, a10, a11, l00, l10, l11, pad;
};
constexpr int SP_D = sizeof(StepParams) / sizeof(double);

// one step of the chain, the same instruction sequence in rollout.cu, risk.cu and fused.cu (bit-identical samples)
__device__ __forceinline__ void rollout_step(const StepParams& p, double& x, double& y, double zx, double zy) {
    const double nx = fma(p.a00, x, fma(p.a01, y, fma(p.l00, zx, p.bx)));
    const double ny = fma(p.a10, x, fma(p.a11, y, fma(p.l10, zx, fma(p.l11, zy, p.by))));
    x = nx;
    y = ny;
}


// Built once per CTA: thread NT-1 fixes the start point x0 (supplied, or the ONE draw shared by all samples of
// the cell, v8ideal:2664-2665), threads t < Tsh build (b_t, A_t, L_t) of step t; `write_out`: this CTA reports status / x0 of the cell.  Caller syncs.
template <int NT>
__device__ __forceinline__ void rollout_setup(const double* __restrict__ mean, const double* __restrict__ cov,
                                              const double* __restrict__ cross, int Tm, const double* __restrict__ x0,
                                              int64_t c, int Tsh, uint32_t seed_lo, uint32_t seed_hi, StepParams* sp,
                                              double* s_x0, uint8_t* __restrict__ status, double* __restrict__ x0_out,
                                              int tid, bool write_out) {
    if (tid == NT - 1) {
        if (x0) {
            s_x0[0] = x0[2 * c];
            s_x0[1] = x0[2 * c + 1];
        } else {
            // the ONE draw x0 ~ N(mean_0, cov_0) shared by all samples of the cell (v8ideal:2664-2665),
            // from a reserved Philox counter so that every CTA (and every GPU shard) derives the same value
            Philox r = philox4x32_10(0xFFFFFFFFu, 0xFFFFFFFFu, (uint32_t)c, 0xFFFFFFFFu, seed_lo, seed_hi);
            float za, zb;
            box_muller(r.c[0], r.c[1], za, zb);
            const double* S0 = cov + (c * Tm) * 4;
            const double l00 = sqrt(S0[0]);
            const double l10 = S0[2] / l00;
            const double l11 = sqrt(S0[3] - l10 * l10);
            s_x0[0] = mean[(c * Tm) * 2] + l00 * (double)za;
            s_x0[1] = mean[(c * Tm) * 2 + 1] + fma(l10, (double)za, l11 * (double)zb);
        }
        if (x0_out && write_out) { x0_out[2 * c] = s_x0[0]; x0_out[2 * c + 1] = s_x0[1]; }
    }
    if (tid < Tsh) {
        const int t = tid;
        const double* mt = mean + (c * Tm + t) * 2;
        const double* St = cov + (c * Tm + t) * 4;
        const double* Sn = cov + (c * Tm + t + 1) * 4;
        const double* Cx = cross + ((c * Tm + t + 1) * Tm + t) * 4;
        StepParams p;
        int st = 0;
        // inverse of Sigma_t (np.linalg.inv of a general 2x2)
        const double det = diff_of_products(St[0], St[3], St[1], St[2]);
        if (det == 0.0 || !isfinite(det)) st = CCMPC_ST_SINGULAR;
        const double id = 1.0 / det;
        const double i00 = St[3] * id, i01 = -St[1] * id, i10 = -St[2] * id, i11 = St[0] * id;
        p.a00 = fma(Cx[0], i00, Cx[1] * i10);
        p.a01 = fma(Cx[0], i01, Cx[1] * i11);
        p.a10 = fma(Cx[2], i00, Cx[3] * i10);
        p.a11 = fma(Cx[2], i01, Cx[3] * i11);
        // cond_cov = Sigma_{t+1} - A C^T ; Cholesky reads the lower triangle
        const double k00 = Sn[0] - fma(p.a00, Cx[0], p.a01 * Cx[1]);
        const double k10 = Sn[2] - fma(p.a10, Cx[0], p.a11 * Cx[1]);
        const double k11 = Sn[3] - fma(p.a10, Cx[2], p.a11 * Cx[3]);
        if (!(k00 > 0.0) && st == 0) st = CCMPC_ST_NO_TANGENT;
        p.l00 = sqrt(k00);
        p.l10 = k10 / p.l00;
        const double r = k11 - p.l10 * p.l10;
        if (!(r > 0.0) && st == 0) st = CCMPC_ST_NO_TANGENT;
        p.l11 = sqrt(r);
        p.bx = mt[2] - fma(p.a00, mt[0], p.a01 * mt[1]);
        p.by = mt[3] - fma(p.a10, mt[0], p.a11 * mt[1]);
        p.pad = 0.0;
        sp[t] = p;
        if (status && write_out && st) status[c] = (uint8_t)st;
    }
}

}  // namespace ccmpc
