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
#include "common.cuh"

namespace ccmpc {

constexpr int ST_THREADS = 256;
constexpr int ST_MAXP = 8;          // register-prefetched passes per chunk
constexpr int ST_TILE_BYTES = 32768;
constexpr int ST_CHUNKS_PER_UNIT = 16;

struct StatsGeom {
    int T, D, DP, nb, ntile, ns, rpp, ch, npass, slab;
    __host__ explicit StatsGeom(int T_) {
        T = T_;
        D = 2 * T_;
        DP = (D + 3) / 4 * 4;
        nb = DP / 4;
        ntile = nb * (nb + 1) / 2;
        ns = ST_THREADS / ntile;
        rpp = ST_THREADS / T;
        int c = ST_MAXP * rpp;
        int cap = ST_TILE_BYTES / (DP * 8);
        if (c > cap) c = cap;
        if (c > ns * 16) c = ns * 16;
        if (c >= ns) c = c / ns * ns;
        ch = c;
        npass = (ch + rpp - 1) / rpp;
        slab = ch * ST_CHUNKS_PER_UNIT;
    }
    __host__ size_t smem_bytes() const {
        size_t tile = (size_t)ch * DP;
        size_t red = (size_t)ns * ntile * 16;
        size_t red2 = (size_t)ST_THREADS * 7;
        size_t m = tile > red ? tile : red;
        if (red2 > m) m = red2;
        return m * 8 + (size_t)ch * 4 + (size_t)4 * T * 8 + 16;
    }
};

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
    int T, D, DP, nb, ntile, ns, rpp, ch, npass, slab;
};

__global__ void __launch_bounds__(ST_THREADS, 2) sample_sums_kernel(const StatsArgs a) {
    extern __shared__ __align__(16) double smem[];
    const int tid = threadIdx.x;
    const int T = a.T, DP = a.DP;
    // ---- which unit am I -------------------------------------------------
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
    const int64_t cell_begin = a.offsets[c], cell_end = a.offsets[c + 1];
    const int64_t begin = cell_begin + (int64_t)slab_idx * a.slab;
    const int rows_unit = (int)min((int64_t)a.slab, cell_end - begin);

    // ---- shared memory carve-up -----------------------------------------
    size_t tile_d = (size_t)a.ch * DP;
    size_t red_d = (size_t)a.ns * a.ntile * 16;
    size_t big = tile_d > red_d ? tile_d : red_d;
    if ((size_t)ST_THREADS * 7 > big) big = (size_t)ST_THREADS * 7;
    double* tile = smem;
    double* s_piv = smem + big;                 // [2T]
    double* s_ego = s_piv + 2 * T;              // [2T]
    int* s_any = reinterpret_cast<int*>(s_ego + 2 * T);   // [ch]
    __shared__ int s_anycnt;

    for (int i = tid; i < (int)tile_d; i += ST_THREADS) tile[i] = 0.0;
    for (int i = tid; i < a.ch; i += ST_THREADS) s_any[i] = 0;
    if (tid < 2 * T) {
        s_piv[tid] = a.pivot[c * 2 * T + tid];
        s_ego[tid] = a.ego ? a.ego[c * a.ego_cell_stride + tid] : 0.0;
    }
    if (tid == 0) s_anycnt = 0;
    __syncthreads();

    // ---- roles -----------------------------------------------------------
    const int rl = tid / T, t = tid - rl * T;
    const bool stager = rl < a.rpp;
    const int tl = tid % a.ntile, subset = tid / a.ntile;
    const bool worker = subset < a.ns;
    int rb = 0;
    while ((rb + 1) * (rb + 2) / 2 <= tl) ++rb;
    const int cb = tl - rb * (rb + 1) / 2;

    const double pivx = s_piv[2 * t], pivy = s_piv[2 * t + 1];
    const double ex = s_ego[2 * t], ey = s_ego[2 * t + 1];
    const bool do_count = a.ego != nullptr;
    const bool do_yaw = a.yaws != nullptr;
    const double r2 = a.r2;

    double acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = 0.0;
    double s1x = 0.0, s1y = 0.0, pen = 0.0, d2s = 0.0, yaw_sum = 0.0, yaw_sq = 0.0;
    int coll = 0, any_cnt = 0;

    double2 pre[ST_MAXP];
    double ypre[ST_MAXP];
    const double2* g2 = reinterpret_cast<const double2*>(a.samples);

    auto prefetch = [&](int chunk0) {
        const int rows = min(a.ch, rows_unit - chunk0);
#pragma unroll
        for (int p = 0; p < ST_MAXP; ++p) {
            const int row = p * a.rpp + rl;
            if (stager && p < a.npass && row < rows) {
                const int64_t e = (begin + chunk0 + row) * T + t;
                pre[p] = __ldcs(g2 + e);
                if (do_yaw) ypre[p] = __ldcs(a.yaws + e);
            }
        }
    };

    prefetch(0);
    for (int chunk0 = 0; chunk0 < rows_unit; chunk0 += a.ch) {
        const int rows = min(a.ch, rows_unit - chunk0);
        // ---- registers -> shared tile (pivot shift), predicate on the raw values
#pragma unroll
        for (int p = 0; p < ST_MAXP; ++p) {
            const int row = p * a.rpp + rl;
            if (stager && p < a.npass && row < rows) {
                const double x = pre[p].x, y = pre[p].y;
                const double ux = x - pivx, uy = y - pivy;
                s1x += ux;
                s1y += uy;
                *reinterpret_cast<double2*>(tile + (size_t)row * DP + 2 * t) = make_double2(ux, uy);
                if (do_count) {
                    // A9 predicate, fixed order, no FMA contraction
                    const double dx = __dsub_rn(ex, x), dy = __dsub_rn(ey, y);
                    const double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
                    d2s += d2;
                    if (d2 < r2) {
                        ++coll;
                        pen += r2 - d2;
                        s_any[row] = 1;
                    }
                }
                if (do_yaw) { yaw_sum += ypre[p]; yaw_sq = fma(ypre[p], ypre[p], yaw_sq); }
            }
        }
        __syncthreads();
        if (chunk0 + a.ch < rows_unit) prefetch(chunk0 + a.ch);   // loads in flight during the rank-k update
        // ---- any-time hits of this chunk
        if (do_count) {
            for (int i = tid; i < rows; i += ST_THREADS) {
                any_cnt += s_any[i];
                s_any[i] = 0;
            }
        }
        // ---- rank-k update: 4x4 register tile per thread, samples strided over subsets
        if (worker) {
            const double* ap = tile + 4 * rb;
            const double* bp = tile + 4 * cb;
#pragma unroll 2
            for (int s = subset; s < rows; s += a.ns) {
                const double2 a01 = *reinterpret_cast<const double2*>(ap + (size_t)s * DP);
                const double2 a23 = *reinterpret_cast<const double2*>(ap + (size_t)s * DP + 2);
                const double2 b01 = *reinterpret_cast<const double2*>(bp + (size_t)s * DP);
                const double2 b23 = *reinterpret_cast<const double2*>(bp + (size_t)s * DP + 2);
                const double av[4] = {a01.x, a01.y, a23.x, a23.y};
                const double bv[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i * 4 + j] = fma(av[i], bv[j], acc[i * 4 + j]);
            }
        }
        __syncthreads();
    }

    // ---- unit epilogue: fixed-order reductions into the partial vector ----
    const SumsLayout L(T);
    double* out = a.partials + (int64_t)u * L.len;
    double* red = tile;
    if (worker) {
#pragma unroll
        for (int i = 0; i < 16; ++i) red[((size_t)subset * a.ntile + tl) * 16 + i] = acc[i];
    }
    __syncthreads();
    for (int e = tid; e < a.ntile * 16; e += ST_THREADS) {
        const int tt = e >> 4, ij = e & 15;
        int trb = 0;
        while ((trb + 1) * (trb + 2) / 2 <= tt) ++trb;
        const int tcb = tt - trb * (trb + 1) / 2;
        const int i = 4 * trb + (ij >> 2), j = 4 * tcb + (ij & 3);
        if (i < a.D && j < a.D && i >= j) {
            double s = 0.0;
            for (int k = 0; k < a.ns; ++k) s += red[((size_t)k * a.ntile + tt) * 16 + ij];
            out[L.o_s2 + i * (i + 1) / 2 + j] = s;
        }
    }
    __syncthreads();
    red[tid * 7 + 0] = s1x;
    red[tid * 7 + 1] = s1y;
    red[tid * 7 + 2] = (double)coll;
    red[tid * 7 + 3] = pen;
    red[tid * 7 + 4] = d2s;
    red[tid * 7 + 5] = yaw_sum;
    red[tid * 7 + 6] = yaw_sq;
    if (any_cnt) atomicAdd(&s_anycnt, any_cnt);
    __syncthreads();
    if (tid < T) {
        double r[7] = {0, 0, 0, 0, 0, 0, 0};
        for (int k = 0; k < a.rpp; ++k)
#pragma unroll
            for (int q = 0; q < 7; ++q) r[q] += red[(k * T + tid) * 7 + q];
        out[L.o_s1 + 2 * tid] = r[0];
        out[L.o_s1 + 2 * tid + 1] = r[1];
        out[L.o_coll + tid] = r[2];
        out[L.o_pen + tid] = r[3];
        out[L.o_d2s + tid] = r[4];
        out[L.o_yaw + tid] = r[5];
        out[L.o_yaw2 + tid] = r[6];
    }
    if (tid == 0) {
        out[0] = (double)rows_unit;
        out[L.o_any] = (double)s_anycnt;
    }
}

// sums[c] = sum over the cell's units, in unit order
__global__ void sums_finalize_kernel(const double* __restrict__ partials, const int32_t* __restrict__ cub,
                                     int len, double* __restrict__ sums) {
    const int64_t c = blockIdx.x;
    const int j = blockIdx.y * blockDim.x + threadIdx.x;
    if (j >= len) return;
    const int u0 = cub[c], u1 = cub[c + 1];
    double s = 0.0;
    for (int u = u0; u < u1; ++u) s += partials[(int64_t)u * len + j];
    sums[c * len + j] = s;
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

extern "C" size_t ccmpc_sample_sums_workspace_bytes(int64_t n_total, int64_t C, int T) {
    if (T < 1 || T > CCMPC_MAX_T || n_total < 0 || C < 0) return 0;
    StatsGeom g(T);
    size_t cub = ((size_t)(C + 1) * sizeof(int32_t) + 255) / 256 * 256;
    return cub + (size_t)max_units(n_total, C, g.slab) * SumsLayout(T).len * sizeof(double) + 256;
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
    StatsGeom g(T);
    SumsLayout L(T);
    int32_t* cub = reinterpret_cast<int32_t*>(workspace);
    size_t cub_bytes = ((size_t)(C + 1) * sizeof(int32_t) + 255) / 256 * 256;
    double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(workspace) + cub_bytes);
    unit_scan_kernel<<<1, 1024, 0, st>>>(offsets, C, g.slab, cub);
    int rc = check_launch("unit_scan_kernel");
    if (rc) return rc;
    StatsArgs a;
    a.samples = samples; a.offsets = offsets; a.cub = cub; a.pivot = pivot; a.ego = ego;
    a.ego_cell_stride = ego_cell_stride; a.yaws = yaws; a.partials = partials; a.C = C; a.r2 = R * R;
    a.T = T; a.D = g.D; a.DP = g.DP; a.nb = g.nb; a.ntile = g.ntile; a.ns = g.ns; a.rpp = g.rpp; a.ch = g.ch;
    a.npass = g.npass; a.slab = g.slab;
    const int64_t mu = max_units(n_total, C, g.slab);
    CCMPC_REQUIRE(mu < (int64_t)1 << 31, "ccmpc_sample_sums: too many units");
    sample_sums_kernel<<<(unsigned)mu, ST_THREADS, g.smem_bytes(), st>>>(a);
    rc = check_launch("sample_sums_kernel");
    if (rc) return rc;
    dim3 fg((unsigned)C, (unsigned)((L.len + 127) / 128));
    sums_finalize_kernel<<<fg, 128, 0, st>>>(partials, cub, L.len, sums);
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
