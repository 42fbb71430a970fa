// group.cu - stage-1 input preparation (A2) of libccmpc.
//
// Reference: collect/in_simulation/midlevel/ovehicle.py:72-76 (per-sample yaw from consecutive
// predicted positions): yaw[n, 0] = atan2 of (p[n, 0] - last observed position),
// yaw[n, t] = atan2 of (p[n, t] - p[n, t-1]).
#include <cooperative_groups.h>

#include "common.cuh"

namespace ccmpc {

// largest c with offsets[c] <= row  (cells are contiguous row ranges; empty cells are skipped)
__device__ __forceinline__ int64_t cell_of_row(const int64_t* __restrict__ offsets, int64_t C, int64_t row) {
    int64_t lo = 0, hi = C;
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (offsets[mid] <= row) lo = mid; else hi = mid;
    }
    return lo;
}

__global__ void yaws_kernel(const double* __restrict__ samples, const int64_t* __restrict__ offsets, int64_t C,
                            int64_t N, int T, const double* __restrict__ pos_last, double* __restrict__ yaws) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * T) return;
    // 32-bit division when the element count allows (a 64-bit divide is ~100 instructions)
    const int64_t row = (N * T < ((int64_t)1 << 31)) ? (int64_t)((uint32_t)i / (uint32_t)T) : i / T;
    const int t = (int)(i - row * T);
    const double2* p2 = reinterpret_cast<const double2*>(samples);
    const double2 cur = p2[i];
    double2 prev;
    if (t == 0) {
        const int64_t c = cell_of_row(offsets, C, row);
        prev = make_double2(pos_last[2 * c], pos_last[2 * c + 1]);
    } else {
        prev = p2[i - 1];
    }
    yaws[i] = atan2(cur.y - prev.y, cur.x - prev.x);
}

// ---------------------------------------------------------------------------
// A2: bucket one vehicle's samples by latent and fold rare latents into the nearest retained cluster.
//
// Reference: v8ideal/__init__.py:486-494 (predictions + minpos -> float64, bucket by z in sample order)
// and midlevel/ovehicle.py:55-113:
//   keep latents with pmf > filter_pmf, ascending (:58); centre_j = mean final position of latent j's own
//   samples (:80); every other latent, ascending, sends each of its samples to the nearest centre by final
//   position (scipy distance_matrix + argmin, first minimum wins, :94-95), appended after what the cluster
//   already holds (:96-104); pmf_j = N_j / N (:109-113).
// The resulting row order is a STABLE sort of the samples by (cluster j, origin g) with g = 0 for the
// cluster's own latent and g = 1 + latent otherwise: two LSD counting-sort passes over sample indices.
// One 8-CTA cluster per vehicle; cell index = v * S + j (cells j >= K_v stay empty).
// ---------------------------------------------------------------------------
constexpr int GR_THREADS = 1024;
constexpr int GR_WARPS = GR_THREADS / 32;
constexpr int GR_MAXS = 32;          // latent states handled (Trajectron++ uses 25)
constexpr int GR_CLUSTER = 8;        // CTAs per vehicle (portable cluster size)

struct GroupArgs {
    const float* pred32;             // [V,N,T,2] float32 (+ minpos) or
    const double* pred64;            // [V,N,T,2] float64 (minpos already added); exactly one is non-null
    const int32_t* z;                // [V,N]
    const double* latent_pmf;        // [V,S]
    int64_t N;
    int T, S;
    double minx, miny, filter_pmf;
    uint16_t* keys;                  // [V,N] workspace
    int32_t* idx1;                   // [V,N] workspace
    int64_t* perm;                   // [V,N] out: source sample of every grouped row (vehicle-local index)
    int64_t* offsets;                // [V*S+1] out
    int32_t* n_states;               // [V] out
    int32_t* cell_latent;            // [V*S] out (-1 for an unused cell)
    double* cell_pmf;                // [V*S] out
    double* init_center;             // [V*S,2] out
    uint8_t* status;                 // [V] out: 0 ok, 2 no latent above filter_pmf (the reference fails in argmin)
};

__device__ __forceinline__ double2 final_position(const GroupArgs& a, int64_t v, int64_t n) {
    const int64_t e = ((v * a.N + n) * a.T + (a.T - 1));
    if (a.pred32) {
        const float2 f = reinterpret_cast<const float2*>(a.pred32)[e];
        return make_double2((double)f.x + a.minx, (double)f.y + a.miny);          // v8ideal:486
    }
    return reinterpret_cast<const double2*>(a.pred64)[e];
}

// One vehicle = one thread-block CLUSTER of GR_CLUSTER CTAs; CTA r owns the r-th contiguous chunk of the
// vehicle's samples (and, in the second sort pass, of the intermediate order), so "stable" means: bins in
// order, inside a bin CTAs in rank order, inside a CTA the original order.  The CTAs exchange their per-bin
// counts and per-latent partial sums through distributed shared memory.
namespace cg = cooperative_groups;

// one stable counting-sort pass over positions [lo, hi) of the vehicle's current order.  key(n) in [0, 32);
// in_idx == nullptr means the identity order.  Writes out_idx at GLOBAL sorted positions; every CTA ends with
// the vehicle-wide bin starts in s_start[33].
template <typename OutT, typename KeyFn>
__device__ void stable_pass(cg::cluster_group& cl, int64_t lo, int64_t hi, const int32_t* in_idx, OutT* out_idx,
                            KeyFn key_of, int* s_hist /*[32], DSMEM-visible*/, int* s_pos /*[32]*/,
                            int* s_warp /*[GR_WARPS][32]*/, int* s_start /*[33]*/) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned rank = cl.block_rank(), nranks = cl.num_blocks();
    if (tid < 32) s_hist[tid] = 0;
    __syncthreads();
    for (int64_t i = lo + tid; i < hi; i += GR_THREADS) atomicAdd(&s_hist[key_of(in_idx ? in_idx[i] : (int32_t)i)], 1);
    __syncthreads();
    cl.sync();                                         // every CTA's histogram is complete and visible
    if (tid < 32) {
        int before = 0, total = 0;
        for (unsigned r = 0; r < nranks; ++r) {
            const int h = cl.map_shared_rank(s_hist, r)[tid];
            if (r < rank) before += h;
            total += h;
        }
        s_pos[tid] = before;                           // + bin start, below
        s_warp[tid] = total;
    }
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int b = 0; b < 32; ++b) { s_start[b] = run; run += s_warp[b]; }
        s_start[32] = run;
    }
    __syncthreads();
    if (tid < 32) s_pos[tid] += s_start[tid];          // running write position of this CTA in every bin
    __syncthreads();
    for (int64_t base = lo; base < hi; base += GR_THREADS) {
        const int64_t i = base + tid;
        const bool live = i < hi;
        const int32_t src = live ? (in_idx ? in_idx[i] : (int32_t)i) : 0;
        const int key = live ? key_of(src) : 32 + lane;                 // dead lanes never match a live key
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        const int rk = __popc(peers & ((1u << lane) - 1u));
        s_warp[warp * 32 + lane] = 0;
        __syncwarp();
        if (live && rk == 0) s_warp[warp * 32 + key] = __popc(peers);
        __syncthreads();
        if (tid < 32) {                                                  // bin tid: exclusive scan over warps
            int run = s_pos[tid];
            for (int w = 0; w < GR_WARPS; ++w) {
                const int cnt = s_warp[w * 32 + tid];
                s_warp[w * 32 + tid] = run;
                run += cnt;
            }
            s_pos[tid] = run;
        }
        __syncthreads();
        if (live) out_idx[s_warp[warp * 32 + key] + rk] = (OutT)src;
        __syncthreads();
    }
    __threadfence();
    cl.sync();                                         // out_idx complete cluster-wide; s_hist may be reused
}

__global__ void __cluster_dims__(GR_CLUSTER, 1, 1) __launch_bounds__(GR_THREADS) group_kernel(const GroupArgs a) {
    __shared__ int s_rank[GR_MAXS];              // latent -> retained rank j, or -1
    __shared__ int s_keep[GR_MAXS];              // j -> latent
    __shared__ int s_K;
    __shared__ double s_acc[GR_WARPS][GR_MAXS][2];
    __shared__ int s_cnt[GR_WARPS][GR_MAXS];
    __shared__ double s_part[GR_MAXS][2];        // this CTA's per-latent sums (DSMEM-visible)
    __shared__ int s_pcnt[GR_MAXS];
    __shared__ double s_cx[GR_MAXS], s_cy[GR_MAXS];
    __shared__ int s_hist[32], s_pos[32], s_start[33], s_warp[GR_WARPS * 32];
    cg::cluster_group cl = cg::this_cluster();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned rank = cl.block_rank(), nranks = cl.num_blocks();
    const int64_t v = blockIdx.y;
    const int S = a.S;
    const int64_t N = a.N;
    const int64_t chunk = (N + nranks - 1) / nranks;
    const int64_t lo = min(N, (int64_t)rank * chunk), hi = min(N, lo + chunk);
    // ---- retained latents (ovehicle.py:58)
    if (tid == 0) {
        int K = 0;
        for (int s = 0; s < S; ++s) {
            const bool keep = a.latent_pmf[v * S + s] > a.filter_pmf;
            s_rank[s] = keep ? K : -1;
            if (keep) s_keep[K++] = s;
        }
        s_K = K;
    }
    for (int e = tid; e < GR_WARPS * GR_MAXS; e += GR_THREADS) {
        (&s_acc[0][0][0])[2 * e] = 0.0;
        (&s_acc[0][0][0])[2 * e + 1] = 0.0;
        (&s_cnt[0][0])[e] = 0;
    }
    __syncthreads();
    const int K = s_K;
    const int32_t* zv = a.z + v * N;
    // ---- centres: mean final position of every retained latent's own samples (:80); fixed summation order
    //      (lanes by shuffle tree, warps in order, CTAs in rank order), identical in every CTA of the cluster
    for (int64_t base = lo; base < hi; base += GR_THREADS) {
        const int64_t n = base + tid;
        int j = -1;
        double2 p = make_double2(0.0, 0.0);
        if (n < hi) {
            const int zz = zv[n];
            j = (zz >= 0 && zz < S) ? s_rank[zz] : -1;
            if (j >= 0) p = final_position(a, v, n);
        }
        unsigned todo = __ballot_sync(0xffffffffu, j >= 0);
        while (todo) {
            const int leader = __ffs(todo) - 1;
            const int lj = __shfl_sync(0xffffffffu, j, leader);
            const unsigned members = __ballot_sync(0xffffffffu, j == lj);
            double sx = j == lj ? p.x : 0.0, sy = j == lj ? p.y : 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                sx += __shfl_xor_sync(0xffffffffu, sx, o);
                sy += __shfl_xor_sync(0xffffffffu, sy, o);
            }
            if (lane == leader) {
                s_acc[warp][lj][0] += sx;
                s_acc[warp][lj][1] += sy;
                s_cnt[warp][lj] += __popc(members);
            }
            todo &= ~members;
        }
    }
    __syncthreads();
    if (tid < GR_MAXS) {
        double sx = 0.0, sy = 0.0;
        int cnt = 0;
        if (tid < K)
            for (int w = 0; w < GR_WARPS; ++w) { sx += s_acc[w][tid][0]; sy += s_acc[w][tid][1]; cnt += s_cnt[w][tid]; }
        s_part[tid][0] = sx; s_part[tid][1] = sy; s_pcnt[tid] = cnt;
    }
    __syncthreads();
    cl.sync();
    if (tid < K) {
        double sx = 0.0, sy = 0.0;
        int cnt = 0;
        for (unsigned r = 0; r < nranks; ++r) {
            const double* pr = cl.map_shared_rank(&s_part[0][0], r);
            sx += pr[2 * tid]; sy += pr[2 * tid + 1];
            cnt += cl.map_shared_rank(s_pcnt, r)[tid];
        }
        s_cx[tid] = sx / (double)cnt;          // 0/0 = NaN for a retained latent without samples, as np.mean
        s_cy[tid] = sy / (double)cnt;
    }
    __syncthreads();
    // ---- key of every sample of the chunk: cluster j (major) and origin g (minor)
    uint16_t* keys = a.keys + v * N;
    for (int64_t n = lo + tid; n < hi; n += GR_THREADS) {
        const int zz = zv[n];
        int j = (zz >= 0 && zz < S) ? s_rank[zz] : -1;
        int g = 0;
        if (j < 0) {
            g = 1 + (zz < 0 ? 0 : (zz >= S ? S - 1 : zz));
            const double2 p = final_position(a, v, n);
            // scipy.spatial.distance_matrix (p = 2) + np.argmin: first strict minimum (ovehicle.py:94-95)
            // np.argmin propagates NaN: a retained latent WITHOUT own samples has a NaN centre (np.mean of an
            // empty slice) and the first such centre wins
            double best = __longlong_as_double(0x7ff0000000000000LL);
            bool nan_seen = false;
            j = 0;
            for (int k = 0; k < K; ++k) {
                const double dx = fabs(__dsub_rn(p.x, s_cx[k])), dy = fabs(__dsub_rn(p.y, s_cy[k]));
                const double dist = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
                if (nan_seen) continue;
                if (dist != dist) { j = k; nan_seen = true; }
                else if (dist < best) { best = dist; j = k; }
            }
        }
        keys[n] = (uint16_t)(j * 32 + g);
    }
    __threadfence();
    cl.sync();                                   // keys of every chunk written (pass 2 reads other CTAs' keys)
    // ---- stable sort by (j, g): minor key first
    int32_t* idx1 = a.idx1 + v * N;
    int64_t* perm = a.perm + v * N;
    stable_pass<int32_t>(cl, lo, hi, nullptr, idx1, [&](int32_t n) { return (int)(keys[n] & 31); }, s_hist, s_pos,
                         s_warp, s_start);
    stable_pass<int64_t>(cl, lo, hi, idx1, perm, [&](int32_t n) { return (int)(keys[n] >> 5); }, s_hist, s_pos,
                         s_warp, s_start);
    // ---- per-cell outputs (s_start now holds the vehicle-wide cluster starts), written by CTA 0
    if (rank == 0 && tid < S) {
        const int64_t c = v * S + tid;
        const int start = tid < 32 ? s_start[tid] : (int)N, stop = tid + 1 <= 32 ? s_start[tid + 1] : (int)N;
        a.offsets[c] = v * N + (tid < K ? start : N);
        if (a.cell_latent) a.cell_latent[c] = tid < K ? s_keep[tid] : -1;
        if (a.cell_pmf) a.cell_pmf[c] = tid < K ? (double)(stop - start) / (double)N : 0.0;     // :109-113
        if (a.init_center) {
            a.init_center[2 * c] = tid < K ? s_cx[tid] : 0.0;
            a.init_center[2 * c + 1] = tid < K ? s_cy[tid] : 0.0;
        }
    }
    if (rank == 0 && tid == 0) {
        if (v == gridDim.y - 1) a.offsets[(v + 1) * S] = (v + 1) * N;
        a.n_states[v] = K;
        if (a.status) a.status[v] = K > 0 ? 0 : CCMPC_ST_NO_TANGENT;
    }
}

// samples[v*N + i] = (double) predictions[v, perm[v, i]] + minpos     (v8ideal:486 on the grouped order)
__global__ void gather_rows_kernel(const float* __restrict__ pred32, const double* __restrict__ pred64,
                                   const int64_t* __restrict__ perm, int64_t rows, int64_t N, int T, double minx,
                                   double miny, double* __restrict__ samples,
                                   const double* __restrict__ pos_last, double* __restrict__ yaws) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= rows * T) return;
    const bool small = rows * T < ((int64_t)1 << 31);
    const int64_t row = small ? (int64_t)((uint32_t)e / (uint32_t)T) : e / T;
    const int t = (int)(e - row * T);
    const int64_t v = small ? (int64_t)((uint32_t)row / (uint32_t)N) : row / N;
    const int64_t src = (v * N + perm[row]) * T + t;
    auto fetch = [&](int64_t i) {
        if (pred32) {
            const float2 f = reinterpret_cast<const float2*>(pred32)[i];
            return make_double2((double)f.x + minx, (double)f.y + miny);
        }
        return reinterpret_cast<const double2*>(pred64)[i];
    };
    const double2 o = fetch(src);
    reinterpret_cast<double2*>(samples)[e] = o;
    if (yaws) {
        // ovehicle.py:72-76 on the row just gathered (same arithmetic as yaws_kernel): heading of the step
        // from the previous position, the vehicle's last observed position for t = 0
        const double2 prev = t == 0 ? make_double2(pos_last[2 * v], pos_last[2 * v + 1]) : fetch(src - 1);
        yaws[e] = atan2(o.y - prev.y, o.x - prev.x);
    }
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" int ccmpc_yaws_from_positions(const double* samples, const int64_t* offsets, int64_t n_total, int64_t C,
                                         int T, const double* pos_last, double* yaws, void* stream) {
    CCMPC_REQUIRE(T >= 1 && T <= CCMPC_MAX_T && C >= 1 && n_total >= 0, "ccmpc_yaws_from_positions: bad dims");
    CCMPC_REQUIRE(offsets && pos_last && yaws && (samples || n_total == 0), "ccmpc_yaws_from_positions: null pointer");
    if (n_total == 0) return CCMPC_OK;
    const int64_t tot = n_total * T;
    yaws_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>(samples, offsets, C, n_total, T,
                                                                                pos_last, yaws);
    return check_launch("yaws_kernel");
}


extern "C" size_t ccmpc_group_workspace_bytes(int64_t V, int64_t N) {
    if (V < 0 || N < 0) return 0;
    return (size_t)V * N * (sizeof(uint16_t) + sizeof(int32_t)) + 512;
}

extern "C" int ccmpc_group_by_mode(const float* pred32, const double* pred64, const int32_t* z,
                                   const double* latent_pmf, int64_t V, int64_t N, int T, int S, double minpos_x,
                                   double minpos_y, double filter_pmf, double* samples, int64_t* offsets,
                                   int64_t* perm, int32_t* n_states, int32_t* cell_latent, double* cell_pmf,
                                   double* init_center, uint8_t* status, const double* pos_last, double* yaws,
                                   void* workspace, size_t workspace_bytes, void* stream) {
    CCMPC_REQUIRE(T >= 1 && T <= CCMPC_MAX_T, "ccmpc_group_by_mode: need 1 <= T <= %d", CCMPC_MAX_T);
    CCMPC_REQUIRE(S >= 1 && S <= GR_MAXS, "ccmpc_group_by_mode: need 1 <= S <= %d latent states", GR_MAXS);
    CCMPC_REQUIRE(V >= 0 && N >= 0 && N < ((int64_t)1 << 31), "ccmpc_group_by_mode: bad V / N");
    CCMPC_REQUIRE((pred32 != nullptr) != (pred64 != nullptr) || V * N == 0,
                  "ccmpc_group_by_mode: pass exactly one of pred32 / pred64");
    if (V == 0) return CCMPC_OK;
    CCMPC_REQUIRE(z && latent_pmf && samples && offsets && perm && n_states, "ccmpc_group_by_mode: null pointer");
    CCMPC_REQUIRE(!yaws || pos_last, "ccmpc_group_by_mode: yaws needs pos_last [V,2]");
    const size_t need = ccmpc_group_workspace_bytes(V, N);
    if (!workspace || workspace_bytes < need) {
        set_error("ccmpc_group_by_mode: workspace %zu < %zu bytes", workspace_bytes, need);
        return CCMPC_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    GroupArgs a;
    a.pred32 = pred32; a.pred64 = pred64; a.z = z; a.latent_pmf = latent_pmf; a.N = N; a.T = T; a.S = S;
    a.minx = minpos_x; a.miny = minpos_y; a.filter_pmf = filter_pmf;
    a.idx1 = reinterpret_cast<int32_t*>(workspace);
    a.keys = reinterpret_cast<uint16_t*>(reinterpret_cast<char*>(workspace) + (((size_t)V * N * 4 + 255) / 256) * 256);
    a.perm = perm; a.offsets = offsets; a.n_states = n_states; a.cell_latent = cell_latent; a.cell_pmf = cell_pmf;
    a.init_center = init_center; a.status = status;
    CCMPC_REQUIRE(V <= 65535, "ccmpc_group_by_mode: V > 65535");
    group_kernel<<<dim3(GR_CLUSTER, (unsigned)V), GR_THREADS, 0, st>>>(a);
    int rc = check_launch("group_kernel");
    if (rc || N == 0) return rc;
    const int64_t tot = V * N * T;
    gather_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(pred32, pred64, perm, V * N, N, T, minpos_x,
                                                                     minpos_y, samples, pos_last, yaws);
    return check_launch("gather_rows_kernel");
}
