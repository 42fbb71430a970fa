// sums_common.cuh - what the one-pass sums kernels share (stats.cu: samples read from HBM; fused.cu: samples
// produced in shared memory by the rollout): the FP64 mma.sync Gram update, the per-warp epilogue area and the
// mapping from the packed sums layout (ccmpc.h "sums layout") to the fragment layout.  One definition, so the
// two paths cannot drift apart ("mc_rollout_sums == mc_rollout -> sample_sums" rests on it).
//
// Fragment layout (see stats.cu): lane = 4 q + r, r = sample within the 4-sample group (MMA k index), q = MMA
// m / n index; the lane holds columns NF q .. NF q + NF - 1 of sample r (column = 2 t + {x, y}).  Fragment j
// serves as A and as B of tile (bi, bj): C[m][n] += sum_k U[k][NF m + bi] U[k][NF n + bj]; the NF (NF + 1) / 2
// tiles with bi >= bj cover every (i >= j) entry of the 2T x 2T second-moment matrix.
#pragma once
#include "common.cuh"

namespace ccmpc {

__device__ __forceinline__ void dmma8x8x4(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// all tiles of one 4-sample group: acc[ti] += f[bi] (x) f[bj]
template <int NF>
__device__ __forceinline__ void gram_update(double (&acc)[NF * (NF + 1) / 2][2], const double (&f)[NF]) {
    int ti = 0;
#pragma unroll
    for (int bi = 0; bi < NF; ++bi)
#pragma unroll
        for (int bj = 0; bj <= bi; ++bj, ++ti) dmma8x8x4(acc[ti][0], acc[ti][1], f[bi], f[bj]);
}

// per-warp epilogue area (doubles): Gram tiles, then S1 and the per-timestep scalars
template <int NF>
struct SumsWarpArea {
    static constexpr int NT = NF * (NF + 1) / 2;   // 8x8 tiles (bi >= bj)
    static constexpr int TP = 4 * NF;              // timesteps covered by a warp (padded)
    static constexpr int O_S1 = NT * 64;           // [8 NF] column sums
    static constexpr int O_COLL = O_S1 + 8 * NF;   // [TP] each: coll, pen, d2s, yaw, yaw2
    static constexpr int O_PEN = O_COLL + TP;
    static constexpr int O_D2S = O_PEN + TP;
    static constexpr int O_YAW = O_D2S + TP;
    static constexpr int O_YAW2 = O_YAW + TP;
    static constexpr int O_ANY = O_YAW2 + TP;
    static constexpr int PW = O_ANY + 2;           // doubles per warp
};

// the lane's two accumulators of tile ti go to wsm[ti * 64 + q * 8 + 2 r + {0, 1}]  (C[m = q][n = 2 r + {0, 1}])
template <int NF>
__device__ __forceinline__ void store_gram_tiles(double* wsm, const double (&acc)[NF * (NF + 1) / 2][2], int q, int r) {
#pragma unroll
    for (int ti = 0; ti < NF * (NF + 1) / 2; ++ti)
        *reinterpret_cast<double2*>(wsm + ti * 64 + q * 8 + 2 * r) = make_double2(acc[ti][0], acc[ti][1]);
}

// Lane group q <-> column group (which NF columns of a sample lane group q holds).  stats.cu: identity (a lane
// loads NF contiguous columns from global memory).  fused.cu reads its fragments from the 128-byte-swizzled
// shared-memory tile, where the identity makes the two lane groups of a quarter-warp hit the same banks; swapping
// two bits of q (an involution) spreads them: PERM 1 = bits 0 and 1 (NF = 4), PERM 2 = bits 0 and 2 (NF = 2).
template <int PERM>
__device__ __forceinline__ int col_group(int q) {
    if (PERM == 1) return (q & 4) | ((q & 1) << 1) | ((q >> 1) & 1);
    if (PERM == 2) return (q & 2) | ((q & 1) << 2) | ((q >> 2) & 1);
    return q;
}

// Where element e of the packed sums vector lives inside a SumsWarpArea; -1: e == 0 (the sample count),
// -2: a slot this kernel does not produce (HAS_YAW / HAS_D2S false).
template <int NF, int PERM = 0>
__device__ __forceinline__ int packed_to_area(int e, const SumsLayout& L, bool has_yaw, bool has_d2s) {
    using SM = SumsWarpArea<NF>;
    if (e == 0) return -1;
    if (e < L.o_s2) return SM::O_S1 + (e - L.o_s1);         // S1 is stored by column, whatever the lane mapping
    if (e < L.o_coll) {
        const int idx = e - L.o_s2;
        int i = (int)((sqrtf(8.0f * (float)idx + 1.0f) - 1.0f) * 0.5f);
        while (i * (i + 1) / 2 > idx) --i;
        while ((i + 1) * (i + 2) / 2 <= idx) ++i;
        const int j = idx - i * (i + 1) / 2;
        int bi = i % NF, bj = j % NF, m = col_group<PERM>(i / NF), n = col_group<PERM>(j / NF);
        if (bi < bj) { int t0 = bi; bi = bj; bj = t0; t0 = m; m = n; n = t0; }
        return (bi * (bi + 1) / 2 + bj) * 64 + m * 8 + n;
    }
    if (e < L.o_any) return SM::O_COLL + (e - L.o_coll);
    if (e == L.o_any) return SM::O_ANY;
    if (e < L.o_d2s) return SM::O_PEN + (e - L.o_pen);
    if (e < L.o_yaw) return has_d2s ? SM::O_D2S + (e - L.o_d2s) : -2;
    if (!has_yaw) return -2;
    if (e < L.o_yaw2) return SM::O_YAW + (e - L.o_yaw);
    return SM::O_YAW2 + (e - L.o_yaw2);
}

// hi word of a double -> unsigned key ordered like the value (ties in the low word compare equal)
__device__ __forceinline__ uint32_t hi_key(double v) {
    const int h = __double2hiint(v);
    return (uint32_t)h ^ ((uint32_t)(h >> 31) | 0x80000000u);
}

// d2s[t] = sum_n |ego_t - p_{n,t}|^2 from the shifted power sums (a = ego - pivot, u = p - pivot):
//   sum |a - u|^2 = n |a|^2 - 2 a . S1 + S2xx + S2yy.
// Used where the per-sample accumulation of d2s is not worth two FP64-pipe instructions per trajectory-step
// (the quantity is a cost term with a 1e-5 tolerance, not a count).
__device__ __forceinline__ double d2s_from_sums(const double* __restrict__ s, const SumsLayout& L, int t, double ax,
                                                double ay) {
    const int ix = 2 * t, iy = 2 * t + 1;
    const double n = s[0];
    return fma(n, fma(ax, ax, ay * ay), -2.0 * fma(ax, s[L.o_s1 + ix], ay * s[L.o_s1 + iy])) + s[L.s2(ix, ix)] +
           s[L.s2(iy, iy)];
}

}  // namespace ccmpc
