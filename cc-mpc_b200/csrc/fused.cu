// fused.cu - K1 + K2 + K6 in one kernel: Monte Carlo rollout, trajectory store and the one-pass sums.
//
// ccmpc_mc_rollout writes the (C, n, T, 2) trajectory tensor and ccmpc_sample_sums reads it straight back
// (32 B of HBM traffic per trajectory-step).  Here the CTA that rolls 128 samples out keeps them in shared
// memory (the transpose tile the rollout kernel stages its coalesced store through anyway), writes them to
// HBM once, and feeds the SAME tile to the FP64 mma.sync Gram update and the A9 predicate of sample_sums:
// 16 B per trajectory-step, and the issue-bound rollout phase of one CTA overlaps the FP64-pipe-bound sums
// phase of its neighbours on the SM.
//
//   rollout:  predict_ideal body v8ideal:2662-2709 - same Philox counters and arithmetic as rollout.cu
//             (sample i of cell c is bit-identical to ccmpc_mc_rollout's)
//   sums:     np.mean / np.cov of every (t, tau) pair (makeconstraint:55, v8ideal:2586-2606) + A9 counts,
//             same fragment layout, predicate order and packed layout as stats.cu (ccmpc.h "sums layout")
// Restrictions of this fast path: Philox noise, horizon 8 or 16, no yaw sums.
#include "rollout.cuh"

namespace ccmpc {

constexpr int FU_THREADS = 128;
constexpr int FU_WARPS = FU_THREADS / 32;
#ifndef FU_CTAS_PER_SM
#define FU_CTAS_PER_SM 4
#endif

__device__ __forceinline__ void fu_dmma8x8x4(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

template <int NF>
struct FusedSmem {                                   // per-warp epilogue area (doubles), as StatsSmem in stats.cu
    static constexpr int NT = NF * (NF + 1) / 2;
    static constexpr int TP = 4 * NF;
    static constexpr int O_S1 = NT * 64;
    static constexpr int O_COLL = O_S1 + 8 * NF;
    static constexpr int O_PEN = O_COLL + TP;
    static constexpr int O_D2S = O_PEN + TP;
    static constexpr int O_ANY = O_D2S + TP;
    static constexpr int PW = O_ANY + 2;
};

// TSH = 4 NF timesteps.  grid (gx, C): CTA (bx, c) rolls out and reduces tiles bx, bx + gx, ... of cell c.
template <int NF>
__global__ void __launch_bounds__(FU_THREADS, FU_CTAS_PER_SM)
rollout_sums_kernel(const double* __restrict__ mean, const double* __restrict__ cov, const double* __restrict__ cross,
                    int Tm, const double* __restrict__ x0, int64_t n, uint32_t seed_lo, uint32_t seed_hi,
                    uint64_t sample_offset, const double* __restrict__ pivot, const double* __restrict__ ego,
                    int64_t ego_cell_stride, double r2, double* __restrict__ traj, double* __restrict__ partials,
                    uint8_t* __restrict__ status, double* __restrict__ x0_out) {
    constexpr int TSH = 4 * NF;
    constexpr int NT = FusedSmem<NF>::NT;
    constexpr int TS = NF / 2;                       // timesteps per lane
    constexpr int STRIDE2 = TSH | 1;                 // double2 per staged row (odd: conflict-free row writes)
    using SM = FusedSmem<NF>;
    extern __shared__ __align__(16) double smem[];
    StepParams* sp = reinterpret_cast<StepParams*>(smem);                        // [TSH]
    double2* stage = reinterpret_cast<double2*>(smem + TSH * SP_D);               // [128][STRIDE2]
    __shared__ double s_x0[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int r = lane & 3, q = lane >> 2;
    const int64_t c = blockIdx.y;

    rollout_setup<FU_THREADS>(mean, cov, cross, Tm, x0, c, TSH, seed_lo, seed_hi, sp, s_x0, status, x0_out, tid);

    // ---- per-lane constants of the sums phase: its TS timesteps
    double pivx[TS], pivy[TS], ex[TS], ey[TS];
#pragma unroll
    for (int i = 0; i < TS; ++i) {
        const int t = TS * q + i;
        pivx[i] = pivot[(c * TSH + t) * 2];
        pivy[i] = pivot[(c * TSH + t) * 2 + 1];
        ex[i] = ego[c * ego_cell_stride + 2 * t];
        ey[i] = ego[c * ego_cell_stride + 2 * t + 1];
    }
    double acc[NT][2];
#pragma unroll
    for (int i = 0; i < NT; ++i) acc[i][0] = acc[i][1] = 0.0;
    double s1x[TS], s1y[TS], pen[TS], d2s[TS];
    int coll[TS];
#pragma unroll
    for (int i = 0; i < TS; ++i) { s1x[i] = s1y[i] = pen[i] = d2s[i] = 0.0; coll[i] = 0; }
    int any_cnt = 0;
    int64_t rows_done = 0;
    __syncthreads();

    const int64_t ntiles = (n + FU_THREADS - 1) / FU_THREADS;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t i0 = tile * FU_THREADS;
        const int rows = (int)min((int64_t)FU_THREADS, n - i0);
        // ---- phase 1: one sample per thread into the staged tile (absent rows hold the pivot: u = 0)
        {
            double2* my_stage = stage + tid * STRIDE2;
            if (tid < rows) {
                double x = s_x0[0], y = s_x0[1];
                const uint64_t gi = sample_offset + (uint64_t)(i0 + tid);
                const uint32_t gi_lo = (uint32_t)gi, gi_hi = (uint32_t)(gi >> 32);
                auto step = [&](int t, double zx, double zy) {
                    const StepParams& p = sp[t];
                    const double dx = x - p.mux, dy = y - p.muy;
                    const double cmx = p.mnx + fma(dx, p.a00, dy * p.a01);
                    const double cmy = p.mny + fma(dx, p.a10, dy * p.a11);
                    x = cmx + zx * p.l00;
                    y = cmy + fma(zx, p.l10, zy * p.l11);
                    my_stage[t] = make_double2(x, y);
                };
#pragma unroll
                for (int t2 = 0; t2 < TSH / 2; ++t2) {
                    Philox rr = philox4x32_10(gi_lo, gi_hi, (uint32_t)c, (uint32_t)t2, seed_lo, seed_hi);
                    float za, zb, zc, zd;
                    box_muller(rr.c[0], rr.c[1], za, zb);
                    box_muller(rr.c[2], rr.c[3], zc, zd);
                    step(2 * t2, (double)za, (double)zb);
                    step(2 * t2 + 1, (double)zc, (double)zd);
                }
            } else {
                for (int t = 0; t < TSH; ++t)
                    my_stage[t] = make_double2(pivot[(c * TSH + t) * 2], pivot[(c * TSH + t) * 2 + 1]);
            }
        }
        __syncwarp();                                 // every phase works on the warp's own 32 rows: no CTA barrier,
                                                      // so the issue-bound rollout of one warp overlaps the FP64-pipe-bound
                                                      // sums of the others
        // ---- phase 2: contiguous store of the warp's (wrows, TSH, 2) block
        if (traj) {
            const int wrows = min(32, rows - 32 * warp);
            double2* out2 = reinterpret_cast<double2*>(traj) + (c * n + i0 + 32 * warp) * TSH;
            const double2* wstage = stage + 32 * warp * STRIDE2;
            constexpr int RPP = 32 / TSH;
            const int t = lane % TSH, r0 = lane / TSH;
#pragma unroll 4
            for (int rw = r0; rw < wrows; rw += RPP) __stcs(out2 + rw * TSH + t, wstage[rw * STRIDE2 + t]);
        }
        // ---- phase 3: this warp's 32 rows, 4-sample groups: shift, S1, A9 predicate, Gram update
#pragma unroll 4
        for (int g = 0; g < 8; ++g) {
            const int row = 32 * warp + 4 * g + r;
            const bool rv = row < rows;
            const double2* src = stage + row * STRIDE2 + TS * q;
            double f[NF];
            bool hit_any = false;
#pragma unroll
            for (int i = 0; i < TS; ++i) {
                const double2 cur = src[i];
                const double ux = cur.x - pivx[i], uy = cur.y - pivy[i];
                s1x[i] += ux;
                s1y[i] += uy;
                f[2 * i] = ux;
                f[2 * i + 1] = uy;
                if (rv) {
                    // A9 predicate, fixed order, no FMA contraction (identical to stats.cu)
                    const double dx = __dsub_rn(ex[i], cur.x), dy = __dsub_rn(ey[i], cur.y);
                    const double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
                    d2s[i] += d2;
                    if (d2 < r2) {
                        ++coll[i];
                        pen[i] += r2 - d2;
                        hit_any = true;
                    }
                }
            }
            const unsigned bal = __ballot_sync(0xffffffffu, hit_any);
            if (bal) {
#pragma unroll
                for (int rr = 0; rr < 4; ++rr) any_cnt += (bal & (0x11111111u << rr)) != 0;
            }
            int ti = 0;
#pragma unroll
            for (int bi = 0; bi < NF; ++bi)
#pragma unroll
                for (int bj = 0; bj <= bi; ++bj, ++ti) fu_dmma8x8x4(acc[ti][0], acc[ti][1], f[bi], f[bj]);
        }
        rows_done += rows;
        __syncwarp();                                 // the warp's rows are overwritten by its next rollout
    }
    __syncthreads();                                  // the epilogue area aliases the staged tiles of all warps

    // ---- warp epilogue (as stats.cu): fixed-order shuffles over the 4 sample rows, then shared memory
    //      (the staged tiles are dead: barrier above)
    double* wsm = smem + (size_t)warp * SM::PW;
#pragma unroll
    for (int ti = 0; ti < NT; ++ti)
        *reinterpret_cast<double2*>(wsm + ti * 64 + q * 8 + 2 * r) = make_double2(acc[ti][0], acc[ti][1]);
#pragma unroll
    for (int i = 0; i < TS; ++i) {
        double v[5] = {s1x[i], s1y[i], (double)coll[i], pen[i], d2s[i]};
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            v[k] += __shfl_xor_sync(0xffffffffu, v[k], 1);
            v[k] += __shfl_xor_sync(0xffffffffu, v[k], 2);
        }
        if (r == 0) {
            const int t = TS * q + i;
            wsm[SM::O_S1 + 2 * t] = v[0];
            wsm[SM::O_S1 + 2 * t + 1] = v[1];
            wsm[SM::O_COLL + t] = v[2];
            wsm[SM::O_PEN + t] = v[3];
            wsm[SM::O_D2S + t] = v[4];
        }
    }
    if (lane == 0) wsm[SM::O_ANY] = (double)any_cnt;
    __syncthreads();

    // ---- CTA epilogue: add the warps in warp order into this CTA's partial vector (packed sums layout)
    const SumsLayout L(TSH);
    double* out = partials + (c * gridDim.x + blockIdx.x) * (int64_t)L.len;
    for (int e = tid; e < L.len; e += FU_THREADS) {
        int src;
        if (e == 0) {
            out[0] = (double)rows_done;
            continue;
        } else if (e < L.o_s2) {
            src = SM::O_S1 + (e - L.o_s1);
        } else if (e < L.o_coll) {
            const int idx = e - L.o_s2;
            int i = (int)((sqrtf(8.0f * (float)idx + 1.0f) - 1.0f) * 0.5f);
            while (i * (i + 1) / 2 > idx) --i;
            while ((i + 1) * (i + 2) / 2 <= idx) ++i;
            const int j = idx - i * (i + 1) / 2;
            int bi = i % NF, bj = j % NF, m = i / NF, nn = j / NF;
            if (bi < bj) { int t0 = bi; bi = bj; bj = t0; t0 = m; m = nn; nn = t0; }
            src = (bi * (bi + 1) / 2 + bj) * 64 + m * 8 + nn;
        } else if (e < L.o_any) {
            src = SM::O_COLL + (e - L.o_coll);
        } else if (e == L.o_any) {
            src = SM::O_ANY;
        } else if (e < L.o_d2s) {
            src = SM::O_PEN + (e - L.o_pen);
        } else if (e < L.o_yaw) {
            src = SM::O_D2S + (e - L.o_d2s);
        } else {
            out[e] = 0.0;                                 // no yaw sums on this path
            continue;
        }
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < FU_WARPS; ++w) s += smem[(size_t)w * SM::PW + src];
        out[e] = s;
    }
}

// sums[c] = sum over the cell's CTA partials, in CTA order
__global__ void fused_finalize_kernel(const double* __restrict__ partials, int nparts, int len,
                                      double* __restrict__ sums) {
    const int64_t c = blockIdx.x;
    const int j = blockIdx.y * blockDim.x + threadIdx.x;
    if (j >= len) return;
    double s = 0.0;
    for (int p = 0; p < nparts; ++p) s += partials[(c * nparts + p) * (int64_t)len + j];
    sums[c * len + j] = s;
}

static int fused_grid_x(int64_t n, int64_t C) {
    int64_t want = (148 * FU_CTAS_PER_SM) / C;         // one wave of resident CTAs, split over the cells
    int64_t need = (n + FU_THREADS - 1) / FU_THREADS;
    int64_t g = want < need ? want : need;
    return (int)(g < 1 ? 1 : g);
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" size_t ccmpc_mc_rollout_sums_workspace_bytes(int64_t C, int64_t n, int Tsh) {
    if (C <= 0 || n <= 0 || Tsh < 1 || Tsh > CCMPC_MAX_T) return 16;
    return (size_t)C * (size_t)fused_grid_x(n, C) * (size_t)SumsLayout(Tsh).len * sizeof(double);
}

extern "C" int ccmpc_mc_rollout_sums(const double* mean, const double* cov, const double* cross, int Tm,
                                     const double* x0, int64_t C, int64_t n, int Tsh, uint64_t seed,
                                     uint64_t sample_offset, const double* pivot, const double* ego,
                                     int64_t ego_cell_stride, double R, double* traj, double* sums, uint8_t* status,
                                     double* x0_out, void* workspace, size_t workspace_bytes, void* stream) {
    CCMPC_REQUIRE(Tsh == 16 || Tsh == 8, "ccmpc_mc_rollout_sums: horizon must be 8 or 16 (got %d)", Tsh);
    CCMPC_REQUIRE(Tm >= Tsh + 1, "ccmpc_mc_rollout_sums: moments must cover Tsh + 1 steps (Tm = %d)", Tm);
    CCMPC_REQUIRE(C >= 0 && n >= 0 && C < 65536, "ccmpc_mc_rollout_sums: bad C / n");
    if (C == 0) return CCMPC_OK;
    CCMPC_REQUIRE(mean && cov && cross && pivot && ego && sums, "ccmpc_mc_rollout_sums: null pointer");
    CCMPC_REQUIRE(ego_cell_stride == 0 || ego_cell_stride == 2 * Tsh, "ccmpc_mc_rollout_sums: ego stride must be 0 or 2 Tsh");
    CCMPC_REQUIRE(((uintptr_t)traj & 15) == 0, "ccmpc_mc_rollout_sums: traj must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const SumsLayout L(Tsh);
    if (n == 0) {
        cudaError_t e = cudaMemsetAsync(sums, 0, (size_t)C * L.len * sizeof(double), st);
        if (e != cudaSuccess) { set_error("ccmpc_mc_rollout_sums: memset: %s", cudaGetErrorString(e)); return CCMPC_E_LAUNCH; }
        return CCMPC_OK;
    }
    CCMPC_REQUIRE(workspace && workspace_bytes >= ccmpc_mc_rollout_sums_workspace_bytes(C, n, Tsh),
                  "ccmpc_mc_rollout_sums: workspace too small");
    if (status) {
        cudaError_t e = cudaMemsetAsync(status, 0, (size_t)C, st);
        if (e != cudaSuccess) { set_error("ccmpc_mc_rollout_sums: memset: %s", cudaGetErrorString(e)); return CCMPC_E_LAUNCH; }
    }
    const int gx = fused_grid_x(n, C);
    dim3 grid((unsigned)gx, (unsigned)C);
    double* partials = reinterpret_cast<double*>(workspace);
    const size_t smem = (size_t)Tsh * sizeof(StepParams) + (size_t)FU_THREADS * (Tsh | 1) * 16;
    if (Tsh == 16)
        rollout_sums_kernel<4><<<grid, FU_THREADS, smem, st>>>(mean, cov, cross, Tm, x0, n, (uint32_t)seed,
                                                               (uint32_t)(seed >> 32), sample_offset, pivot, ego,
                                                               ego_cell_stride, R * R, traj, partials, status, x0_out);
    else
        rollout_sums_kernel<2><<<grid, FU_THREADS, smem, st>>>(mean, cov, cross, Tm, x0, n, (uint32_t)seed,
                                                               (uint32_t)(seed >> 32), sample_offset, pivot, ego,
                                                               ego_cell_stride, R * R, traj, partials, status, x0_out);
    int rc = check_launch("rollout_sums_kernel");
    if (rc != CCMPC_OK) return rc;
    dim3 fg((unsigned)C, (unsigned)((L.len + 127) / 128));
    fused_finalize_kernel<<<fg, 128, 0, st>>>(partials, gx, L.len, sums);
    return check_launch("fused_finalize_kernel");
}
