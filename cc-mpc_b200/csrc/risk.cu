// risk.cu - fused Monte Carlo collision-risk estimate: rollout + evaluation with NO trajectory tensor.
//
// BASELINE config "MC collision-risk estimate, 1M samples / obstacle, ph16" only needs the per-sample
// evaluation of the rolled-out obstacle against the ego plan, not the samples themselves.  This kernel runs the
// predict_ideal chain (v8ideal:2662-2709, same arithmetic and the same Philox stream as rollout.cu, so sample i
// of cell c is bit-identical to ccmpc_mc_rollout's) and applies the A9 predicate (R = 3.4,
// v8ideal:795; same no-FMA operation order as ccmpc_sample_sums) to every (sample, timestep) in registers.
// Algorithmic HBM bytes: 0 per trajectory-step (a few KB of partials per CTA) - the kernel is bound by
// instruction issue (Philox + Box-Muller + the fp64 chain), not by memory; bench.py reports it against the
// materialising path (16 B written + 16 B read per trajectory-step).
//
// Per cell the output is the "risk vector" of 3T + 2 doubles
//     [0] n   [1, 1+T) coll[t]   [1+T] any   [2+T, 2+2T) pen[t]   [2+2T, 2+3T) d2s[t]
// with the meanings of the packed sums layout (ccmpc.h).  Counts are integers held exactly in doubles: CTA
// partials are added in a fixed order by risk_finalize_kernel, and one fp64 all-reduce over sample shards is
// exact for them; every sum is accumulated per thread and reduced in a fixed order, so the result is
// deterministic for a given grid.
#include "rollout.cuh"

namespace ccmpc {

constexpr int RK_THREADS = 128;
constexpr int RK_WARPS = RK_THREADS / 32;
constexpr int RK_CTAS_PER_SM = 5;     // 80 registers; per-thread accumulators live in shared memory (40 KB per CTA at ph16)

template <int TSH>
__global__ void __launch_bounds__(RK_THREADS, RK_CTAS_PER_SM)
risk_kernel(const double* __restrict__ mean, const double* __restrict__ cov, const double* __restrict__ cross, int Tm,
            const double* __restrict__ x0, int64_t n, int Tsh_rt, uint32_t seed_lo, uint32_t seed_hi,
            uint64_t sample_offset, const double* __restrict__ ego, int64_t ego_cell_stride, double r2,
            double* __restrict__ partial, uint8_t* __restrict__ status, double* __restrict__ x0_out) {
    extern __shared__ __align__(16) double smem[];
    constexpr int TMAX = TSH > 0 ? TSH : CCMPC_MAX_T;
    const int Tsh = TSH > 0 ? TSH : Tsh_rt;
    const int L = 3 * Tsh + 2;
    StepParams* sp = reinterpret_cast<StepParams*>(smem);                 // [Tsh]
    double2* s_ego = reinterpret_cast<double2*>(smem + Tsh * SP_D);        // [Tsh]
    double* s_red = smem + Tsh * SP_D + 2 * Tsh;                           // [RK_WARPS][L]
    __shared__ double s_x0[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t c = blockIdx.y;

    rollout_setup<RK_THREADS>(mean, cov, cross, Tm, x0, c, Tsh, seed_lo, seed_hi, sp, s_x0, status, x0_out, tid, blockIdx.x == 0);
    if (tid < Tsh) s_ego[tid] = reinterpret_cast<const double2*>(ego + c * ego_cell_stride)[tid];
    __syncthreads();

    // per-thread accumulators in SHARED memory, [t][tid] (conflict-free): sum d^2 is one LDS + DADD + STS per
    // step, hit counts and the penetration sum are only touched on a hit (a shared fp64 atomicAdd per hit was
    // measured first: it is a CAS loop, 32-way serialised inside a warp, and cells the ego drives through ran 3x longer).  Keeping them out of the register
    // file (48 registers at ph16) doubles the resident warps, which is what hides the Philox / MUFU / fp64-chain
    // latencies of this issue-bound kernel.
    double* s_d2s = s_red + RK_WARPS * L;                                  // [Tsh][RK_THREADS]
    double* s_pen = s_d2s + Tsh * RK_THREADS;                              // [Tsh][RK_THREADS]
    uint32_t* s_coll = reinterpret_cast<uint32_t*>(s_pen + Tsh * RK_THREADS);   // [Tsh][RK_THREADS]
    for (int t = 0; t < Tsh; ++t) {
        s_d2s[t * RK_THREADS + tid] = 0.0;
        s_pen[t * RK_THREADS + tid] = 0.0;
        s_coll[t * RK_THREADS + tid] = 0u;
    }
    uint32_t any_cnt = 0, n_cnt = 0;

    for (int64_t i = (int64_t)blockIdx.x * RK_THREADS + tid; i < n; i += (int64_t)gridDim.x * RK_THREADS) {
        double x = s_x0[0], y = s_x0[1];
        const uint64_t gi = sample_offset + (uint64_t)i;
        const uint32_t gi_lo = (uint32_t)gi, gi_hi = (uint32_t)(gi >> 32);
        bool hit_any = false;
        auto step = [&](int t, double zx, double zy) {
            rollout_step(sp[t], x, y, zx, zy);
            // A9 predicate, fixed order, no FMA contraction (identical to stats.cu)
            const double2 e = s_ego[t];
            const double ex = __dsub_rn(e.x, x), ey = __dsub_rn(e.y, y);
            const double d2 = __dadd_rn(__dmul_rn(ex, ex), __dmul_rn(ey, ey));
            s_d2s[t * RK_THREADS + tid] += d2;
            if (d2 < r2) {
                ++s_coll[t * RK_THREADS + tid];
                s_pen[t * RK_THREADS + tid] += r2 - d2;
                hit_any = true;
            }
        };
#pragma unroll 2
        for (int t2 = 0; t2 < (TSH > 0 ? (TSH + 1) / 2 : CCMPC_MAX_T / 2); ++t2) {
            const int t = 2 * t2;
            if (TSH == 0 && t >= Tsh) break;
            Philox r = philox4x32_10(gi_lo, gi_hi, (uint32_t)c, (uint32_t)t2, seed_lo, seed_hi);
            float za, zb, zc, zd;
            box_muller(r.c[0], r.c[1], za, zb);
            box_muller(r.c[2], r.c[3], zc, zd);
            step(t, (double)za, (double)zb);
            if (t + 1 < Tsh) step(t + 1, (double)zc, (double)zd);
        }
        any_cnt += hit_any ? 1u : 0u;
        ++n_cnt;
    }

    // warp reduction in a fixed order, then the CTA's partial vector
    auto wsum_d = [](double v) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        return v;
    };
    auto wsum_u = [](uint32_t v) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        return v;
    };
    double* wr = s_red + warp * L;
    {
        const uint32_t nn = wsum_u(n_cnt), aa = wsum_u(any_cnt);
        if (lane == 0) { wr[0] = (double)nn; wr[1 + Tsh] = (double)aa; }
    }
#pragma unroll
    for (int t = 0; t < TMAX; ++t) {
        if (TSH == 0 && t >= Tsh) break;
        const uint32_t cc = wsum_u(s_coll[t * RK_THREADS + tid]);
        const double dd = wsum_d(s_d2s[t * RK_THREADS + tid]), pp = wsum_d(s_pen[t * RK_THREADS + tid]);
        if (lane == 0) { wr[1 + t] = (double)cc; wr[2 + Tsh + t] = pp; wr[2 + 2 * Tsh + t] = dd; }
    }
    __syncthreads();
    double* out = partial + (c * gridDim.x + blockIdx.x) * L;
    for (int e = tid; e < L; e += RK_THREADS) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < RK_WARPS; ++w) v += s_red[w * L + e];
        out[e] = v;
    }
}

// risk[c][e] = sum over the cell's CTA partials, fixed order
__global__ void risk_finalize_kernel(const double* __restrict__ partial, int nparts, int L, double* __restrict__ risk) {
    const int64_t c = blockIdx.x;
    for (int e = threadIdx.x; e < L; e += blockDim.x) {
        double v = 0.0;
        for (int p = 0; p < nparts; ++p) v += partial[(c * nparts + p) * L + e];
        risk[c * L + e] = v;
    }
}

static int risk_grid_x(int64_t n, int64_t C) {
    // one wave: 148 SMs x RK_CTAS_PER_SM resident 128-thread CTAs, split evenly over the cells, never more than
    // the samples need
    int64_t want = ((int64_t)sm_count() * RK_CTAS_PER_SM) / C;
    int64_t need = (n + RK_THREADS - 1) / RK_THREADS;
    int64_t g = want < need ? want : need;
    return (int)(g < 1 ? 1 : g);
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" size_t ccmpc_mc_risk_workspace_bytes(int64_t C, int64_t n, int Tsh) {
    if (C <= 0 || n <= 0 || Tsh < 1) return 16;
    return (size_t)C * (size_t)risk_grid_x(n, C) * (size_t)(3 * Tsh + 2) * sizeof(double);
}

extern "C" int ccmpc_mc_risk(const double* mean, const double* cov, const double* cross, int Tm, const double* x0,
                             int64_t C, int64_t n, int Tsh, uint64_t seed, uint64_t sample_offset, const double* ego,
                             int64_t ego_cell_stride, double R, double* risk, uint8_t* status, double* x0_out,
                             void* workspace, size_t workspace_bytes, void* stream) {
    CCMPC_REQUIRE(Tsh >= 1 && Tsh <= CCMPC_MAX_T, "ccmpc_mc_risk: need 1 <= Tsh <= %d", CCMPC_MAX_T);
    CCMPC_REQUIRE(Tm >= Tsh + 1, "ccmpc_mc_risk: moments must cover Tsh + 1 steps (Tm = %d, Tsh = %d)", Tm, Tsh);
    CCMPC_REQUIRE(C >= 0 && n >= 0 && C < 65536, "ccmpc_mc_risk: bad C / n");
    if (C == 0) return CCMPC_OK;
    CCMPC_REQUIRE(mean && cov && cross && ego && risk, "ccmpc_mc_risk: null pointer");
    CCMPC_REQUIRE(CCMPC_ALIGNED16(ego), "ccmpc_mc_risk: ego must be 16-byte aligned");
    CCMPC_REQUIRE(ego_cell_stride == 0 || ego_cell_stride == 2 * Tsh, "ccmpc_mc_risk: ego stride must be 0 or 2 Tsh");
    cudaStream_t st = (cudaStream_t)stream;
    const int L = 3 * Tsh + 2;
    if (n == 0) {
        cudaError_t e = cudaMemsetAsync(risk, 0, (size_t)C * L * sizeof(double), st);
        if (e != cudaSuccess) { set_error("ccmpc_mc_risk: memset: %s", cudaGetErrorString(e)); return CCMPC_E_LAUNCH; }
        return CCMPC_OK;
    }
    CCMPC_REQUIRE(workspace && workspace_bytes >= ccmpc_mc_risk_workspace_bytes(C, n, Tsh),
                  "ccmpc_mc_risk: workspace too small");
    if (status) {
        cudaError_t e = cudaMemsetAsync(status, 0, (size_t)C, st);
        if (e != cudaSuccess) { set_error("ccmpc_mc_risk: memset: %s", cudaGetErrorString(e)); return CCMPC_E_LAUNCH; }
    }
    const int gx = risk_grid_x(n, C);
    dim3 grid((unsigned)gx, (unsigned)C);
    const size_t smem = ((size_t)Tsh * SP_D + 2 * (size_t)Tsh + (size_t)RK_WARPS * L) * sizeof(double) +
                        (size_t)Tsh * RK_THREADS * (2 * sizeof(double) + sizeof(uint32_t));
    double* partial = reinterpret_cast<double*>(workspace);
#define CCMPC_RK_LAUNCH(TSH)                                                                                   \
    risk_kernel<TSH><<<grid, RK_THREADS, smem, st>>>(mean, cov, cross, Tm, x0, n, Tsh, (uint32_t)seed,           \
                                                     (uint32_t)(seed >> 32), sample_offset, ego, ego_cell_stride, \
                                                     R * R, partial, status, x0_out)
    if (Tsh == 16) CCMPC_RK_LAUNCH(16);
    else if (Tsh == 8) CCMPC_RK_LAUNCH(8);
    else CCMPC_RK_LAUNCH(0);
#undef CCMPC_RK_LAUNCH
    int rc = check_launch("risk_kernel");
    if (rc != CCMPC_OK) return rc;
    risk_finalize_kernel<<<(unsigned)C, 64, 0, st>>>(partial, gx, L, risk);
    return check_launch("risk_finalize_kernel");
}
