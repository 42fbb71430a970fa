// stats.cu - K2 (sample moments) + K6 (collision counts / cost) of libccmpc.
//
// ONE read of the (N, T, 2) fp64 sample tensor produces every sum over samples
// the reference path needs:
//   np.mean / np.cov of all (t, tau) pairs   makeconstraint.py:55, v8ideal/__init__.py:894-904,
//                                            :2586-2606 (save_moments), :1485-1493 (affine)
//   np.mean(yaws)                            v8ideal/__init__.py:723
//   collision counts / soft cost (A9)        restated from R = 3.4, v8ideal/__init__.py:795
// The reference re-reads the samples (T-1) times (np.cov per (t, tau)); here the
// full 2T x 2T second-moment matrix is a register-tiled rank-k update out of a
// shared-memory tile, so algorithmic HBM traffic is 16 B per (sample, timestep).
//
// Work decomposition: unit = up to `slab` consecutive samples of one cell;
// one CTA per unit writes a partial sums vector, a second kernel adds the
// partials of each cell in unit order (deterministic, no fp64 atomics).
#include "sums_common.cuh"

namespace ccmpc {

constexpr int ST_THREADS = 256;
constexpr int SS_WARPS = ST_THREADS / 32;
#ifndef SS_DEPTH_DEF
#define SS_DEPTH_DEF 3
#endif
constexpr int SS_DEPTH = SS_DEPTH_DEF;            // 4-sample groups in flight per warp (register prefetch ring)
constexpr int SS_DEPTH_YAW = 2;        // ... when the yaw tensor is streamed as well (register budget)
#ifndef SS_SLAB_DEF
#define SS_SLAB_DEF 2048
#endif
constexpr int SS_SLAB = SS_SLAB_DEF;          // samples per unit (= per CTA) for large inputs; see pick_slab()

// ---------------------------------------------------------------------------
// The 2T x 2T second-moment matrix is a Gram matrix  G = U^T U  over samples
// (U = pivot-shifted coordinates): a dense fp64 contraction of 2T(2T+1)/2 = 528
// unique FMAs per 256-byte sample row at T = 16.  tools/fp64_pipe_bench.cu shows
// DFMA and mma.sync.m8n8k4.f64 (DMMA) share one pipe on B200 (33.7 / 37.1 TFLOP/s),
// but the DMMA form needs ONE register operand per 8 FMAs instead of shared-memory
// operand traffic, so the fragments are loaded straight from global memory:
//
//   lane = 4*q + r   (r = sample within the 4-sample group = MMA k index,
//                     q = 0..7 = MMA m / n index)
//   the lane loads columns NF*q .. NF*q+NF-1 of sample row r (NF = 2: T <= 8,
//   NF = 4: T <= 16) - a warp reads 4 whole rows = one contiguous block;
//   fragment j = column NF*q + j serves as BOTH the A fragment (row.col layout:
//   A[m][k] -> lane 4m+k) and the B fragment (B[k][n] -> lane 4n+k) of
//   tile (bi, bj):  C[m][n] += sum_k U[k][NF*m+bi] * U[k][NF*n+bj].
//   NF(NF+1)/2 tiles (bi >= bj) cover every (i >= j) entry.
// ---------------------------------------------------------------------------
struct StatsArgs {
    const double* samples;
    const int64_t* offsets;
    const int32_t* cub;
    const double* pivot;
    const double* ego;
    int64_t ego_cell_stride;
    const double* yaws;
    double* partials;
    int64_t C;
    double r2;
    int T;
    int slab;
};

template <int NF>
struct StatsSmem : SumsWarpArea<NF> {
    static constexpr size_t BYTES = (size_t)SS_WARPS * SumsWarpArea<NF>::PW * sizeof(double);
};

template <int NF, bool YAW, bool COUNT, bool FULLT>
__global__ void __launch_bounds__(ST_THREADS, 2) sample_sums_kernel(const StatsArgs a) {
    constexpr int DEPTH = YAW ? SS_DEPTH_YAW : SS_DEPTH;
    constexpr int GPR = SS_WARPS * DEPTH;          // 4-sample groups per round of the CTA
    using SM = StatsSmem<NF>;
    constexpr int NT = SM::NT;
    constexpr int TS = NF / 2;                     // timesteps per lane
    extern __shared__ __align__(16) double smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int r = lane & 3, q = lane >> 2;
    const int T = FULLT ? 4 * NF : a.T;
    // ---- which unit am I (everything here is CTA-uniform) ------------------
    const int n_units = a.cub[a.C];
    const int u = blockIdx.x;
    if (u >= n_units) return;
    int64_t lo = 0, hi = a.C;            // largest c with cub[c] <= u
    while (hi - lo > 1) {
        int64_t mid = (lo + hi) >> 1;
        if (a.cub[mid] <= u) lo = mid; else hi = mid;
    }
    const int64_t c = lo;                // cub[c] <= u < cub[c+1]: empty cells own no unit and are never hit
    const int slab_idx = u - a.cub[c];
    const int64_t cell_end = a.offsets[c + 1];
    // balanced split: the cell's ceil(n_c / slab) units get equal shares (multiples of 4 rows) instead of
    // full slabs + a remainder, so one wave of CTAs finishes together
    const int64_t n_cell = cell_end - a.offsets[c];
    const int units_cell = a.cub[c + 1] - a.cub[c];
    const int64_t share = (((n_cell + units_cell - 1) / units_cell) + 3) & ~(int64_t)3;
    const int64_t begin = a.offsets[c] + (int64_t)slab_idx * share;
    const int rows_unit = (int)max((int64_t)0, min(share, cell_end - begin));
    const int ngroups = (rows_unit + 3) >> 2;
    const int rounds = (ngroups + GPR - 1) / GPR;          // uniform loop bounds: the DMMAs below sit in
    const int full_rounds = (rows_unit >> 2) / GPR;        // provably convergent control flow

    // ---- per-lane constants: its TS timesteps ------------------------------
    const double r2 = a.r2;
    // screening box 2^k > R: |dx| >= 2^k or |dy| >= 2^k implies d2 >= r2 (see fused.cu)
    const double box = __hiloint2double((__double2hiint(sqrt(r2)) & 0x7ff00000) + 0x00100000, 0);
    double pivx[TS], pivy[TS], ex[TS], ey[TS];
    uint32_t kx[TS], wx[TS], ky[TS], wy[TS];       // integer screen: key(x) - kx <= wx  <=>  |ex - x| < box (+ ties)
    bool tv[TS];
#pragma unroll
    for (int i = 0; i < TS; ++i) {
        const int t = TS * q + i;
        tv[i] = FULLT || t < T;
        pivx[i] = tv[i] ? a.pivot[(c * T + t) * 2] : 0.0;
        pivy[i] = tv[i] ? a.pivot[(c * T + t) * 2 + 1] : 0.0;
        ex[i] = (tv[i] && COUNT) ? a.ego[c * a.ego_cell_stride + 2 * t] : 0.0;
        ey[i] = (tv[i] && COUNT) ? a.ego[c * a.ego_cell_stride + 2 * t + 1] : 0.0;
        kx[i] = hi_key(ex[i] - box); wx[i] = hi_key(ex[i] + box) - kx[i];
        ky[i] = hi_key(ey[i] - box); wy[i] = hi_key(ey[i] + box) - ky[i];
    }

    double acc[NT][2];
#pragma unroll
    for (int i = 0; i < NT; ++i) acc[i][0] = acc[i][1] = 0.0;
    double s1x[TS], s1y[TS], pen[TS], yaw_s[TS], yaw_q[TS];
    int coll[TS];
#pragma unroll
    for (int i = 0; i < TS; ++i) { s1x[i] = s1y[i] = pen[i] = yaw_s[i] = yaw_q[i] = 0.0; coll[i] = 0; }
    int any_cnt = 0;

    // lane's element (double2 index) inside the unit for group 0 of warp 0; groups advance by 4 T
    const double2* g2 = reinterpret_cast<const double2*>(a.samples) + begin * T + (r * T + TS * q);
    const double* gy = YAW ? a.yaws + begin * T + (r * T + TS * q) : nullptr;

    // one 4-sample group: shift, S1, A9 predicate, Gram update.  `masked`: rows / timesteps may be absent
    // (absent entries are fetched as the pivot itself, so u = 0 without a select).
#define CCMPC_SS_PROCESS(cur, ycur, rv, MASKED)                                                             \
    do {                                                                                                    \
        double f[NF];                                                                                       \
        bool near_any = false;                                                                              \
        _Pragma("unroll") for (int i = 0; i < TS; ++i) {                                                    \
            const bool ok = (MASKED) ? ((rv) && tv[i]) : true;                                              \
            const double x = (cur)[i].x, y = (cur)[i].y;                                                    \
            const double ux = x - pivx[i], uy = y - pivy[i];                                                \
            s1x[i] += ux;                                                                                   \
            s1y[i] += uy;                                                                                   \
            f[2 * i] = ux;                                                                                  \
            f[2 * i + 1] = uy;                                                                              \
            /* A9 screen on the integer pipe: a sample at least `box` away from the ego in x or y cannot */ \
            /* be a hit; the exact fp64 predicate runs only for warps that hold a near sample            */ \
            if (COUNT) near_any |= ok & (hi_key(x) - kx[i] <= wx[i]) & (hi_key(y) - ky[i] <= wy[i]);        \
            if (YAW) { yaw_s[i] += (ycur)[i]; yaw_q[i] = fma((ycur)[i], (ycur)[i], yaw_q[i]); }             \
        }                                                                                                   \
        if (COUNT && __any_sync(0xffffffffu, near_any)) {                                                   \
            bool hit_any = false;                                                                           \
            _Pragma("unroll") for (int i = 0; i < TS; ++i) {                                                \
                const bool ok = (MASKED) ? ((rv) && tv[i]) : true;                                          \
                /* A9 predicate, fixed order, no FMA contraction */                                         \
                const double dx = __dsub_rn(ex[i], (cur)[i].x), dy = __dsub_rn(ey[i], (cur)[i].y);          \
                const double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));                          \
                if (ok && d2 < r2) {                                                                        \
                    ++coll[i];                                                                              \
                    pen[i] += r2 - d2;                                                                      \
                    hit_any = true;                                                                         \
                }                                                                                           \
            }                                                                                               \
            const unsigned bal = __ballot_sync(0xffffffffu, hit_any);                                       \
            _Pragma("unroll") for (int rr = 0; rr < 4; ++rr) any_cnt += (bal & (0x11111111u << rr)) != 0;   \
        }                                                                                                   \
        gram_update<NF>(acc, f);                                                                            \
    } while (0)

    double2 buf[DEPTH][TS];
    double ybuf[DEPTH][YAW ? TS : 1];
    // ---- full rounds: no masks, next round's loads in flight during the Gram update
    auto fetch_full = [&](int rd) {
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) {
            const int off = (((rd * DEPTH + d) * SS_WARPS + warp) * 4) * T;
#pragma unroll
            for (int i = 0; i < TS; ++i) {
                if (FULLT || tv[i]) {
                    buf[d][i] = __ldcs(g2 + off + i);
                    if (YAW) ybuf[d][i] = __ldcs(gy + off + i);
                } else {
                    buf[d][i] = make_double2(0.0, 0.0);
                    if (YAW) ybuf[d][i] = 0.0;
                }
            }
        }
    };
    if (full_rounds > 0) fetch_full(0);
    for (int rd = 0; rd < full_rounds; ++rd) {
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) {
            double2 cur[TS];
            double ycur[YAW ? TS : 1];
#pragma unroll
            for (int i = 0; i < TS; ++i) { cur[i] = buf[d][i]; if (YAW) ycur[i] = ybuf[d][i]; }
            if (rd + 1 < full_rounds) {
                const int off = ((((rd + 1) * DEPTH + d) * SS_WARPS + warp) * 4) * T;
#pragma unroll
                for (int i = 0; i < TS; ++i) {
                    if (FULLT || tv[i]) {
                        buf[d][i] = __ldcs(g2 + off + i);
                        if (YAW) ybuf[d][i] = __ldcs(gy + off + i);
                    }
                }
            }
            CCMPC_SS_PROCESS(cur, ycur, true, !FULLT);
        }
    }
    // ---- tail rounds: masked rows (absent entries read as the pivot / zero yaw)
    for (int rd = full_rounds; rd < rounds; ++rd) {
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) {
            const int gg = (rd * DEPTH + d) * SS_WARPS + warp;
            const bool rv = 4 * gg + r < rows_unit;
            const int off = gg * 4 * T;
#pragma unroll
            for (int i = 0; i < TS; ++i) {
                const bool ok = rv && tv[i];
                buf[d][i] = ok ? __ldcs(g2 + off + i) : make_double2(pivx[i], pivy[i]);
                if (YAW) ybuf[d][i] = ok ? __ldcs(gy + off + i) : 0.0;
            }
        }
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) {
            const int gg = (rd * DEPTH + d) * SS_WARPS + warp;
            const bool rv = 4 * gg + r < rows_unit;
            CCMPC_SS_PROCESS(buf[d], ybuf[d], rv, true);
        }
    }

#undef CCMPC_SS_PROCESS
    // ---- warp epilogue: fixed-order shuffles over the 4 sample rows, then shared memory
    double* wsm = smem + (size_t)warp * SM::PW;
    store_gram_tiles<NF>(wsm, acc, q, r);
#pragma unroll
    for (int i = 0; i < TS; ++i) {
        double v[7] = {s1x[i], s1y[i], (double)coll[i], pen[i], 0.0, yaw_s[i], yaw_q[i]};
#pragma unroll
        for (int k = 0; k < 7; ++k) {
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
            wsm[SM::O_YAW + t] = v[5];
            wsm[SM::O_YAW2 + t] = v[6];
        }
    }
    if (lane == 0) wsm[SM::O_ANY] = (double)any_cnt;
    __syncthreads();

    // ---- unit epilogue: add the warps in warp order into the partial vector
    const SumsLayout L(T);
    double* out = a.partials + (int64_t)u * L.len;
    for (int e = tid; e < L.len; e += ST_THREADS) {
        const int src = packed_to_area<NF>(e, L, true, true);
        if (src < 0) {
            out[e] = (double)rows_unit;
            continue;
        }
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < SS_WARPS; ++w) s += smem[(size_t)w * SM::PW + src];
        out[e] = s;
    }
}

// cell_unit_begin[c] = number of units of cells < c  (exclusive scan of ceil(n_c / slab))
__global__ void unit_scan_kernel(const int64_t* __restrict__ offsets, int64_t C, int slab,
                                 int32_t* __restrict__ cub) {
    __shared__ int warp_tot[32];
    __shared__ int carry;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int64_t base = 0; base < C; base += blockDim.x) {
        const int64_t c = base + tid;
        int v = 0;
        if (c < C) {
            int64_t n = offsets[c + 1] - offsets[c];
            v = n > 0 ? (int)((n + slab - 1) / slab) : 0;
        }
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[w] = x;
        __syncthreads();
        if (w == 0) {
            int tsum = lane < (int)(blockDim.x >> 5) ? warp_tot[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, tsum, o);
                if (lane >= o) tsum += y;
            }
            warp_tot[lane] = tsum;
        }
        __syncthreads();
        const int before = carry + (w > 0 ? warp_tot[w - 1] : 0) + (x - v);
        if (c < C) cub[c] = before;
        __syncthreads();
        if (tid == (int)blockDim.x - 1) carry = before + v;
        __syncthreads();
    }
    if (tid == 0) cub[C] = carry;
}

// sums[c] = sum over the cell's units, in unit order
// The d2s slots (sum of squared ego distances, a cost term) follow from the power sums (d2s_from_sums) instead of two
// FP64-pipe instructions per trajectory-step in the main loop.
__global__ void sums_finalize_kernel(const double* __restrict__ partials, const int32_t* __restrict__ cub, int T,
                                     const double* __restrict__ pivot, const double* __restrict__ ego,
                                     int64_t ego_cell_stride, double* __restrict__ sums) {
    extern __shared__ double s_sum[];
    const SumsLayout L(T);
    const int64_t c = blockIdx.x;
    const int u0 = cub[c], u1 = cub[c + 1];
    for (int j = threadIdx.x; j < L.len; j += blockDim.x) {
        double s = 0.0;
        for (int u = u0; u < u1; ++u) s += partials[(int64_t)u * L.len + j];
        s_sum[j] = s;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < L.len; j += blockDim.x) {
        double v = s_sum[j];
        if (ego && j >= L.o_d2s && j < L.o_yaw) {
            const int t = j - L.o_d2s;
            v = d2s_from_sums(s_sum, L, t, ego[c * ego_cell_stride + 2 * t] - pivot[(c * T + t) * 2],
                              ego[c * ego_cell_stride + 2 * t + 1] - pivot[(c * T + t) * 2 + 1]);
        }
        sums[c * (int64_t)L.len + j] = v;
    }
}

// ---------------------------------------------------------------------------
// save_moments layout from the sums (v8ideal/__init__.py:2586-2606)
// ---------------------------------------------------------------------------
__global__ void moments_kernel(const double* __restrict__ sums, const double* __restrict__ pivot, int T,
                               double* __restrict__ mean, double* __restrict__ cov, double* __restrict__ cross,
                               int64_t* __restrict__ counts, double* __restrict__ yaw_mean,
                               double* __restrict__ yaw_var) {
    const SumsLayout L(T);
    const int64_t c = blockIdx.x;
    const double* s = sums + c * (int64_t)L.len;
    const double n = s[0];
    const double inv_nm1 = 1.0 / (n - 1.0);
    for (int e = threadIdx.x; e < T * T; e += blockDim.x) {
        const int t = e / T, u = e - t * T;
        const int it = 2 * t, iu = 2 * u;
        if (cross) {
            double* o = cross + ((c * T + t) * T + u) * 4;
            if (u < t) {
                o[0] = cov_from_sums(s, L, it, iu, n, inv_nm1);
                o[1] = cov_from_sums(s, L, it, iu + 1, n, inv_nm1);
                o[2] = cov_from_sums(s, L, it + 1, iu, n, inv_nm1);
                o[3] = cov_from_sums(s, L, it + 1, iu + 1, n, inv_nm1);
            } else {
                o[0] = o[1] = o[2] = o[3] = 0.0;
            }
        }
        if (u == 0) {
            if (mean) {
                mean[(c * T + t) * 2] = pivot[c * 2 * T + it] + s[L.o_s1 + it] / n;
                mean[(c * T + t) * 2 + 1] = pivot[c * 2 * T + it + 1] + s[L.o_s1 + it + 1] / n;
            }
            if (cov) {
                double* o = cov + (c * T + t) * 4;
                const double xy = cov_from_sums(s, L, it, it + 1, n, inv_nm1);
                o[0] = cov_from_sums(s, L, it, it, n, inv_nm1);
                o[1] = xy;
                o[2] = xy;
                o[3] = cov_from_sums(s, L, it + 1, it + 1, n, inv_nm1);
            }
            if (counts) counts[c * (T + 1) + t] = (int64_t)llrint(s[L.o_coll + t]);
            if (yaw_mean) yaw_mean[c * T + t] = s[L.o_yaw + t] / n;
            // np.cov(yawData[:, t]) (v8ideal:876), ddof = 1
            if (yaw_var) yaw_var[c * T + t] = (s[L.o_yaw2 + t] - s[L.o_yaw + t] * s[L.o_yaw + t] / n) * inv_nm1;
        }
    }
    if (threadIdx.x == 0 && counts) counts[c * (T + 1) + T] = (int64_t)llrint(s[L.o_any]);
}

// ---------------------------------------------------------------------------
// A9 (ii): violation counts of the emitted half-planes
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(ST_THREADS)
violation_kernel(const double* __restrict__ samples, const int64_t* __restrict__ offsets, int T,
                 const double* __restrict__ nd, const int8_t* __restrict__ idx_side,
                 unsigned long long* __restrict__ viol) {
    __shared__ double s_row[CCMPC_MAX_T * (CCMPC_MAX_T - 1) / 2 * 4];
    __shared__ int s_cnt[CCMPC_MAX_T * (CCMPC_MAX_T - 1) / 2];
    const int64_t c = blockIdx.x;
    const int tid = threadIdx.x;
    const int P = T * (T - 1) / 2;
    for (int p = tid; p < P; p += ST_THREADS) {
        const int64_t o = c * P + p;
        s_row[4 * p + 0] = nd[3 * o];
        s_row[4 * p + 1] = nd[3 * o + 1];
        s_row[4 * p + 2] = nd[3 * o + 2];
        s_row[4 * p + 3] = (idx_side[o] & 2) ? -1.0 : 1.0;
        s_cnt[p] = 0;
    }
    __syncthreads();
    const int rpp = ST_THREADS / T;
    const int rl = tid / T, t = tid - rl * T;
    const int64_t b = offsets[c], e = offsets[c + 1];
    int cnt[CCMPC_MAX_T - 1];
#pragma unroll
    for (int i = 0; i < CCMPC_MAX_T - 1; ++i) cnt[i] = 0;
    const double2* g2 = reinterpret_cast<const double2*>(samples);
    const int p0 = t * (t - 1) / 2;
    if (rl < rpp) {
        for (int64_t row = b + (int64_t)blockIdx.y * rpp + rl; row < e; row += (int64_t)gridDim.y * rpp) {
            const double2 v = __ldcs(g2 + row * T + t);
#pragma unroll
            for (int u = 0; u < CCMPC_MAX_T - 1; ++u) {
                if (u < t) {
                    const double* r = s_row + 4 * (p0 + u);
                    const double val = __dadd_rn(__dmul_rn(r[0], v.x), __dmul_rn(r[1], v.y));
                    if (__dmul_rn(r[3], __dsub_rn(val, r[2])) > 0.0) ++cnt[u];
                }
            }
        }
#pragma unroll
        for (int u = 0; u < CCMPC_MAX_T - 1; ++u)
            if (u < t && cnt[u]) atomicAdd(&s_cnt[p0 + u], cnt[u]);
    }
    __syncthreads();
    for (int p = tid; p < P; p += ST_THREADS)
        if (s_cnt[p]) atomicAdd(viol + c * P + p, (unsigned long long)s_cnt[p]);
}

// pivot[c] = first sample of cell c (zeros when the cell is empty)
__global__ void cell_pivot_kernel(const double* __restrict__ samples, const int64_t* __restrict__ offsets,
                                  int64_t C, int T, double* __restrict__ pivot) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= C * 2 * T) return;
    const int64_t c = i / (2 * T);
    const int j = (int)(i - c * 2 * T);
    const int64_t b = offsets[c], e = offsets[c + 1];
    pivot[i] = e > b ? samples[b * 2 * T + j] : 0.0;
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" int64_t ccmpc_sums_len(int T) {
    if (T < 1 || T > CCMPC_MAX_T) return -1;
    return SumsLayout(T).len;
}

static int64_t max_units(int64_t n_total, int64_t C, int slab) { return n_total / slab + C; }

// Samples per CTA.  2048 amortises the CTA epilogue at Monte Carlo sizes; one MPC step (c2: 20 000 rows,
// c5: 160 000 rows) would then fill only 10-80 of the 296 resident CTA slots, so the slab shrinks (power of
// two, >= 256) until the grid covers the machine about once.
// Large inputs over few cells (c3: 8M rows, 32 cells): ONE wave - the slab grows until the at most
// n_total / slab + C units fit the 296 resident CTA slots, so the per-unit prologue / epilogue (~3.6 us: unit
// lookup, pipeline fill, CTA reduction, partial vector) is paid once per CTA instead of 13 times.
static int pick_slab(int64_t n_total, int64_t C) {
    const int64_t slots = (int64_t)sm_count() * 2;
    if (C <= slots / 2) {
        const int64_t one_wave = ((n_total + (slots - C) - 1) / (slots - C) + 3) & ~(int64_t)3;
        if (one_wave >= SS_SLAB && one_wave < ((int64_t)1 << 24)) return (int)one_wave;
    }
    int slab = SS_SLAB;
    while (slab > 256 && n_total / slab < slots) slab >>= 1;
    return slab;
}

extern "C" size_t ccmpc_sample_sums_workspace_bytes(int64_t n_total, int64_t C, int T) {
    if (T < 1 || T > CCMPC_MAX_T || n_total < 0 || C < 0) return 0;
    size_t cub = ((size_t)(C + 1) * sizeof(int32_t) + 255) / 256 * 256;
    return cub + (size_t)max_units(n_total, C, pick_slab(n_total, C)) * SumsLayout(T).len * sizeof(double) + 256;
}

extern "C" int ccmpc_sample_sums(const double* samples, const int64_t* offsets, int64_t n_total, int64_t C, int T,
                                 const double* pivot, const double* ego, int64_t ego_cell_stride, double R,
                                 const double* yaws, double* sums, void* workspace, size_t workspace_bytes,
                                 void* stream) {
    CCMPC_REQUIRE(T >= 1 && T <= CCMPC_MAX_T, "ccmpc_sample_sums: need 1 <= T <= %d", CCMPC_MAX_T);
    CCMPC_REQUIRE(C >= 0 && n_total >= 0, "ccmpc_sample_sums: negative size");
    CCMPC_REQUIRE(offsets && pivot && sums && (samples || n_total == 0), "ccmpc_sample_sums: null pointer");
    CCMPC_REQUIRE(((uintptr_t)samples & 15) == 0, "ccmpc_sample_sums: samples must be 16-byte aligned");
    CCMPC_REQUIRE(ego_cell_stride == 0 || ego_cell_stride == 2 * T, "ccmpc_sample_sums: ego_cell_stride must be 0 or 2T");
    if (C == 0) return CCMPC_OK;
    const size_t need = ccmpc_sample_sums_workspace_bytes(n_total, C, T);
    if (!workspace || workspace_bytes < need) {
        set_error("ccmpc_sample_sums: workspace %zu < %zu bytes", workspace_bytes, need);
        return CCMPC_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    SumsLayout L(T);
    int32_t* cub = reinterpret_cast<int32_t*>(workspace);
    size_t cub_bytes = ((size_t)(C + 1) * sizeof(int32_t) + 255) / 256 * 256;
    double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(workspace) + cub_bytes);
    const int slab = pick_slab(n_total, C);
    unit_scan_kernel<<<1, 1024, 0, st>>>(offsets, C, slab, cub);
    int rc = check_launch("unit_scan_kernel");
    if (rc) return rc;
    StatsArgs a;
    a.samples = samples; a.offsets = offsets; a.cub = cub; a.pivot = pivot; a.ego = ego;
    a.ego_cell_stride = ego_cell_stride; a.yaws = yaws; a.partials = partials; a.C = C; a.r2 = R * R;
    a.T = T;
    a.slab = slab;
    const int64_t mu = max_units(n_total, C, slab);
    CCMPC_REQUIRE(mu < (int64_t)1 << 31, "ccmpc_sample_sums: too many units");
    {
        const unsigned grid = (unsigned)mu;
        const bool y = yaws != nullptr, cn = ego != nullptr;
#define CCMPC_SS_LAUNCH(NF, Y, CN, FT)                                                                      \
    sample_sums_kernel<NF, Y, CN, FT><<<grid, ST_THREADS, StatsSmem<NF>::BYTES, st>>>(a)
#define CCMPC_SS_PICK(NF, FT)                                                                               \
    do {                                                                                                    \
        if (y && cn) CCMPC_SS_LAUNCH(NF, true, true, FT);                                                   \
        else if (y) CCMPC_SS_LAUNCH(NF, true, false, FT);                                                   \
        else if (cn) CCMPC_SS_LAUNCH(NF, false, true, FT);                                                  \
        else CCMPC_SS_LAUNCH(NF, false, false, FT);                                                         \
    } while (0)
        if (T == 16) CCMPC_SS_PICK(4, true);
        else if (T > 8) CCMPC_SS_PICK(4, false);
        else if (T == 8) CCMPC_SS_PICK(2, true);
        else CCMPC_SS_PICK(2, false);
#undef CCMPC_SS_PICK
#undef CCMPC_SS_LAUNCH
    }
    rc = check_launch("sample_sums_kernel");
    if (rc) return rc;
    sums_finalize_kernel<<<(unsigned)C, 256, (size_t)L.len * sizeof(double), st>>>(partials, cub, T, pivot, ego,
                                                                                   ego_cell_stride, sums);
    return check_launch("sums_finalize_kernel");
}

extern "C" int ccmpc_moments_from_sums(const double* sums, const double* pivot, int64_t C, int T, double* mean,
                                       double* cov, double* cross, int64_t* counts, double* yaw_mean,
                                       double* yaw_var, void* stream) {
    CCMPC_REQUIRE(T >= 1 && T <= CCMPC_MAX_T && C >= 0 && sums && pivot, "ccmpc_moments_from_sums: bad arguments");
    if (C == 0) return CCMPC_OK;
    moments_kernel<<<(unsigned)C, 256, 0, (cudaStream_t)stream>>>(sums, pivot, T, mean, cov, cross, counts, yaw_mean,
                                                                  yaw_var);
    return check_launch("moments_kernel");
}

extern "C" int ccmpc_violation_counts(const double* samples, const int64_t* offsets, int64_t n_total, int64_t C,
                                      int T, const double* nd, const int8_t* idx_side, int64_t* viol,
                                      void* stream) {
    CCMPC_REQUIRE(T >= 1 && T <= CCMPC_MAX_T && C >= 0 && n_total >= 0, "ccmpc_violation_counts: bad dims");
    CCMPC_REQUIRE(offsets && nd && idx_side && viol, "ccmpc_violation_counts: null pointer");
    CCMPC_REQUIRE(CCMPC_ALIGNED16(samples), "ccmpc_violation_counts: samples must be 16-byte aligned");
    const int P = T * (T - 1) / 2;
    if (C == 0 || P == 0) return CCMPC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(viol, 0, (size_t)C * P * sizeof(int64_t), st);
    if (e != cudaSuccess) { set_error("ccmpc_violation_counts: memset: %s", cudaGetErrorString(e)); return CCMPC_E_LAUNCH; }
    const int rpp = ST_THREADS / T;
    int64_t per_cell = n_total / C + 1;
    int64_t y = (per_cell + (int64_t)rpp * 32 - 1) / ((int64_t)rpp * 32);   // ~32 rows per thread
    if (y < 1) y = 1;
    if (y > 1024) y = 1024;
    dim3 grid((unsigned)C, (unsigned)y);
    violation_kernel<<<grid, ST_THREADS, 0, st>>>(samples, offsets, T, nd, idx_side,
                                                  reinterpret_cast<unsigned long long*>(viol));
    return check_launch("violation_kernel");
}

extern "C" int ccmpc_cell_pivot(const double* samples, const int64_t* offsets, int64_t C, int T, double* pivot,
                                void* stream) {
    CCMPC_REQUIRE(T >= 1 && T <= CCMPC_MAX_T && C >= 0 && offsets && pivot, "ccmpc_cell_pivot: bad arguments");
    if (C == 0) return CCMPC_OK;
    const int64_t tot = C * 2 * T;
    cell_pivot_kernel<<<(unsigned)((tot + 127) / 128), 128, 0, (cudaStream_t)stream>>>(samples, offsets, C, T, pivot);
    return check_launch("cell_pivot_kernel");
}

// ---------------------------------------------------------------------------------------------------------------
// ccmpc_sample_sums_static: ccmpc_sample_sums for callers that may not allocate device memory (CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The workspace (unit table + per-unit partial sums) is a file-scope device array
// of CCMPC_STATIC_SCRATCH_BYTES: every GPU that loads the library holds it until the process exits.  All calls
// share it - two calls that overlap on the GPU (different streams, or threads) race on it.  Nothing in it has to
// start at zero: unit_scan_kernel rewrites the unit table and every partial vector the finalize kernel reads is
// written by sample_sums_kernel of the same call, so the contents left by the previous call are never read.
static __device__ __align__(256) unsigned char g_static_scratch[CCMPC_STATIC_SCRATCH_BYTES];

// In host code `g_static_scratch` is only a handle: its device address comes from cudaGetSymbolAddress, once per
// device, with that device current.
static void* static_scratch_address(int* rc_out) {
    constexpr int MAX_DEV = 64;
    static void* addr[MAX_DEV] = {};
    *rc_out = CCMPC_OK;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
        set_error("ccmpc_sample_sums_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        *rc_out = CCMPC_E_LAUNCH;
        return nullptr;
    }
    void* p = __atomic_load_n(&addr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) {
            set_error("ccmpc_sample_sums_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            *rc_out = CCMPC_E_LAUNCH;
            return nullptr;
        }
        __atomic_store_n(&addr[dev], p, __ATOMIC_RELEASE);      // every thread resolves the same address
    }
    return p;
}

extern "C" size_t ccmpc_sample_sums_static_max_bytes(void) { return (size_t)CCMPC_STATIC_SCRATCH_BYTES; }

extern "C" int ccmpc_sample_sums_static(const double* samples, const int64_t* offsets, int64_t n_total, int64_t C,
                                        int T, const double* pivot, const double* ego, int64_t ego_cell_stride,
                                        double R, const double* yaws, double* sums, void* stream) {
    CCMPC_REQUIRE(T >= 1 && T <= CCMPC_MAX_T, "ccmpc_sample_sums_static: need 1 <= T <= %d", CCMPC_MAX_T);
    CCMPC_REQUIRE(C >= 0 && n_total >= 0, "ccmpc_sample_sums_static: negative size");
    const size_t need = ccmpc_sample_sums_workspace_bytes(n_total, C, T);
    if (need > (size_t)CCMPC_STATIC_SCRATCH_BYTES) {
        set_error("ccmpc_sample_sums_static: n_total = %lld, C = %lld, T = %d needs %zu bytes of scratch, the static "
                  "array holds %zu: use ccmpc_sample_sums with a caller-owned workspace",
                  (long long)n_total, (long long)C, T, need, (size_t)CCMPC_STATIC_SCRATCH_BYTES);
        return CCMPC_E_WORKSPACE;
    }
    int rc = CCMPC_OK;
    void* scratch = static_scratch_address(&rc);
    if (rc) return rc;
    return ccmpc_sample_sums(samples, offsets, n_total, C, T, pivot, ego, ego_cell_stride, R, yaws, sums, scratch,
                             (size_t)CCMPC_STATIC_SCRATCH_BYTES, stream);
}
