// rollout.cu - K1: Monte Carlo conditional-Gaussian rollout of libccmpc.
//
// Replaces the loop body of MidlevelAgent.predict_ideal
// (collect/in_simulation/midlevel/v8ideal/__init__.py:2662-2709): for every
// cell (obstacle, mode) and sample
//     A_t  = C_{t+1,t} Sigma_t^-1                      (:2684-2689)
//     x   <- mu_{t+1} + A_t (x - mu_t) + L_t z          (:2690, :2700-2703)
//     L_t  = chol(Sigma_{t+1} - A_t C_{t+1,t}^T)        (:2693-2696)
// starting from the single x0 draw shared by all samples (:2664-2665).
// The reference draws z with NumPy's MT19937 (`RandomState.standard_normal`,
// :2699-2700); here z comes from Philox-4x32-10 (counter-based, so a sample's
// stream depends only on (seed, cell, global sample index, step) and not on how
// samples are split over CTAs or GPUs) through a float32 Box-Muller, or from a
// caller-supplied noise tensor (deterministic parity mode).
//
// One thread per sample; per-step parameters (mu, A, L) of the cell are built
// once per CTA in shared memory; the (128, Tsh, 2) fp64 output tile is
// transposed through shared memory so global stores are contiguous 16-byte
// lanes.  Algorithmic HBM bytes: 16 per (sample, timestep) written.
#include <stdlib.h>
#include <string.h>

#include "rollout.cuh"
#include "tma.cuh"

namespace ccmpc {

constexpr int RO_THREADS = 128;

#ifndef RO_FLUSH
#define RO_FLUSH 8
#endif
#ifndef RO_MIN_CTAS
#define RO_MIN_CTAS 8
#endif
// TSH > 0: compile-time horizon (fully unrolled step loop, shift/mask index math); TSH = 0: runtime Tsh.
// TMA (TSH = 8 / 16 only): the warp stages its 32 rows x 8 timesteps as ONE [32][128 B] sub-tile in the TMA
// 128-byte swizzle (conflict-free lane-per-row writes) and one lane stores it with a single cp.async.bulk.tensor -
// no LDS + STG flush loop through the LSU; the buffer is reused after the bulk copy has read it.
template <int TSH, bool ZIN, bool ZOUT, bool TMA = false>
__global__ void __launch_bounds__(RO_THREADS, RO_MIN_CTAS)
rollout_kernel(const double* __restrict__ mean, const double* __restrict__ cov, const double* __restrict__ cross,
               int Tm, const double* __restrict__ x0, int64_t n, int Tsh_rt, uint32_t seed_lo, uint32_t seed_hi,
               uint64_t sample_offset, const double* __restrict__ z_in, double* __restrict__ traj,
               double* __restrict__ z_out, uint8_t* __restrict__ status, double* __restrict__ x0_out,
               const __grid_constant__ CUtensorMap traj_map) {
    extern __shared__ __align__(1024) double smem[];
    const int Tsh = TSH > 0 ? TSH : Tsh_rt;
    // timesteps staged between two flushes: half the horizon at TSH = 16 (128 B per row and flush - still whole
    // lines - for half the shared memory, so registers, not the tile, bound the resident warps)
    constexpr int FL = (TSH == 16 || TSH == 8) ? RO_FLUSH : TSH;
    const int fl = FL > 0 ? FL : Tsh;
    StepParams* sp = reinterpret_cast<StepParams*>(smem);                       // [Tsh]
    double2* stage = reinterpret_cast<double2*>(smem + Tsh * SP_D);              // [128][stride2]
    const int stride2 = fl | 1;
    const int tid = threadIdx.x;
    const int64_t c = blockIdx.y;
    __shared__ double s_x0[2];

    rollout_setup<RO_THREADS>(mean, cov, cross, Tm, x0, c, Tsh, seed_lo, seed_hi, sp, s_x0, status, x0_out, tid, blockIdx.x == 0);
    __syncthreads();

    // persistent over the cell's 128-sample tiles: the per-step parameters above are built once per CTA, not
    // once per tile (ncu: with one tile per CTA the setup barrier was the second largest stall)
    const int lane = tid & 31, w0 = tid & ~31;
    double2* my_stage = stage + tid * stride2;
    const double2* wstage = stage + w0 * stride2;
    // TMA: per-warp 4 KB sub-tile, 1024-byte aligned (the swizzle pattern is address based)
    unsigned char* tma_tile = reinterpret_cast<unsigned char*>(smem) + ((Tsh * SP_D * 8 + 1023) / 1024) * 1024 + (tid >> 5) * 4096;
    unsigned char* tma_row = tma_tile + lane * 128;
    const int64_t ntiles = (n + RO_THREADS - 1) / RO_THREADS;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t i0 = tile * RO_THREADS;
    const int64_t i = i0 + tid;
    const bool valid = i < n;                    // lanes past the end run along (warp-level flushes) but touch no memory
    const int rows = (int)min((int64_t)RO_THREADS, n - i0);
    const int wrows = min(32, rows - w0);
    if (wrows <= 0) continue;                    // whole warp past the end
    double2* out2 = reinterpret_cast<double2*>(traj) + (c * n + i0 + w0) * Tsh;

    // the warp stores the timesteps [tb, tb + cnt) of the 32 rows it staged: contiguous cnt * 16 B per row
    auto flush = [&](int tb, int cnt) {
        if (TMA) {
            fence_proxy_async();                 // my generic-proxy writes -> visible to the TMA engine
            __syncwarp();
            if (lane == 0) {
                tma_store_box(&traj_map, tma_tile, 2 * tb, (int)(i0 + w0), (int)c);   // rows past n are clipped
                bulk_commit();
                bulk_wait_read<0>();             // the sub-tile is overwritten by the next 8 steps
            }
            __syncwarp();
            return;
        }
        __syncwarp();
        if (FL > 0 && (32 % (FL > 0 ? FL : 1)) == 0) {
            constexpr int TT = FL > 0 ? FL : 1;
            constexpr int RPP = 32 / TT;                   // rows per pass
            const int t = lane % TT, r0 = lane / TT;
#pragma unroll 4
            for (int r = r0; r < wrows; r += RPP) __stcs(out2 + r * Tsh + tb + t, wstage[r * stride2 + t]);
        } else {
            const int total = wrows * cnt;
            for (int e = lane; e < total; e += 32) {
                const int r = e / cnt, t = e - r * cnt;
                __stcs(out2 + r * Tsh + tb + t, wstage[r * stride2 + t]);
            }
        }
        __syncwarp();
    };

    double x = s_x0[0], y = s_x0[1];
    const uint64_t gi = sample_offset + (uint64_t)i;
    const uint32_t gi_lo = (uint32_t)gi, gi_hi = (uint32_t)(gi >> 32);
    const double2* zin2 = ZIN ? reinterpret_cast<const double2*>(z_in) + (c * n + i) * Tsh : nullptr;
    double2* zout2 = ZOUT ? reinterpret_cast<double2*>(z_out) + (c * n + i) * Tsh : nullptr;

    auto step = [&](int t, double zx, double zy) {
        if (ZOUT && valid) zout2[t] = make_double2(zx, zy);
        rollout_step(sp[t], x, y, zx, zy);
        if (TMA) *reinterpret_cast<double2*>(tma_row + (((t & 7) ^ (lane & 7)) << 4)) = make_double2(x, y);
        else my_stage[FL > 0 ? (t % FL) : t] = make_double2(x, y);
    };
#pragma unroll
    for (int t2 = 0; t2 < (TSH > 0 ? (TSH + 1) / 2 : (Tsh + 1) / 2); ++t2) {
        const int t = 2 * t2;
        if (ZIN) {
            const double2 za = valid ? zin2[t] : make_double2(0.0, 0.0);
            step(t, za.x, za.y);
            if (t + 1 < Tsh) {
                const double2 zb = valid ? zin2[t + 1] : make_double2(0.0, 0.0);
                step(t + 1, zb.x, zb.y);
            }
        } else {
            // one Philox block = 4 words = 2 Box-Muller pairs = the noise of steps t and t + 1
            Philox r = philox4x32_10(gi_lo, gi_hi, (uint32_t)c, (uint32_t)t2, seed_lo, seed_hi);
            float za, zb, zc, zd;
            box_muller(r.c[0], r.c[1], za, zb);
            box_muller(r.c[2], r.c[3], zc, zd);
            step(t, (double)za, (double)zb);
            if (t + 1 < Tsh) step(t + 1, (double)zc, (double)zd);
        }
        if (FL > 0 && FL < TSH && (t + 2) % FL == 0 && t + 2 < TSH) flush(t + 2 - FL, FL);
    }
    if (FL > 0) flush(TSH - FL, FL);
    else flush(0, Tsh);
    }   // tiles of this CTA
    if (TMA && lane == 0) bulk_wait_all();       // every sub-tile has landed before the CTA retires
}

__global__ void philox_kernel(const uint32_t* __restrict__ ctr, const uint32_t* __restrict__ key, int64_t M,
                              uint32_t* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    Philox r = philox4x32_10(ctr[4 * i], ctr[4 * i + 1], ctr[4 * i + 2], ctr[4 * i + 3], key[2 * i], key[2 * i + 1]);
    out[4 * i] = r.c[0]; out[4 * i + 1] = r.c[1]; out[4 * i + 2] = r.c[2]; out[4 * i + 3] = r.c[3];
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" int ccmpc_mc_rollout(const double* mean, const double* cov, const double* cross, int Tm,
                                const double* x0, int64_t C, int64_t n, int Tsh, uint64_t seed,
                                uint64_t sample_offset, const double* z, double* traj, double* z_out,
                                uint8_t* status, double* x0_out, void* stream) {
    CCMPC_REQUIRE(Tsh >= 1 && Tsh <= CCMPC_MAX_T, "ccmpc_mc_rollout: need 1 <= Tsh <= %d", CCMPC_MAX_T);
    CCMPC_REQUIRE(Tm >= Tsh + 1, "ccmpc_mc_rollout: moments must cover Tsh + 1 steps (Tm = %d, Tsh = %d)", Tm, Tsh);
    CCMPC_REQUIRE(C >= 0 && n >= 0 && C < 65536, "ccmpc_mc_rollout: bad C / n");
    CCMPC_REQUIRE(mean && cov && cross && traj, "ccmpc_mc_rollout: null pointer");
    CCMPC_REQUIRE(CCMPC_ALIGNED16(traj, z, z_out), "ccmpc_mc_rollout: traj / z / z_out must be 16-byte aligned");
    if (C == 0 || n == 0) return CCMPC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (status) {
        cudaError_t e = cudaMemsetAsync(status, 0, (size_t)C, st);
        if (e != cudaSuccess) { set_error("ccmpc_mc_rollout: memset: %s", cudaGetErrorString(e)); return CCMPC_E_LAUNCH; }
    }
    const bool tma = (Tsh == 16 || Tsh == 8) && RO_FLUSH == 8 && n < ((int64_t)1 << 31) && !getenv("CCMPC_ROLLOUT_NO_TMA");
    CUtensorMap map;
    memset(&map, 0, sizeof(map));
    if (tma) {
        int rc = make_traj_map(&map, traj, C, n, Tsh);
        if (rc != CCMPC_OK) return rc;
    }
    const size_t smem = tma ? ((size_t)Tsh * sizeof(StepParams) + 1023) / 1024 * 1024 + (RO_THREADS / 32) * 4096
                            : (size_t)Tsh * sizeof(StepParams) + (size_t)RO_THREADS * (((Tsh == 16 || Tsh == 8) ? RO_FLUSH : Tsh) | 1) * 16;
    // one wave: at most 148 * RO_MIN_CTAS resident CTAs over all cells, each looping over its cell's tiles
    const int64_t ntiles = (n + RO_THREADS - 1) / RO_THREADS;
    const int64_t cap = std::max<int64_t>(1, ((int64_t)sm_count() * RO_MIN_CTAS) / C);
    dim3 grid((unsigned)std::min(ntiles, cap), (unsigned)C);
#define CCMPC_RO_LAUNCH(TSH, ZI, ZO)                                                                          \
    do {                                                                                                      \
        if (tma && TSH > 0)                                                                                   \
            rollout_kernel<TSH, ZI, ZO, (TSH > 0)><<<grid, RO_THREADS, smem, st>>>(                           \
                mean, cov, cross, Tm, x0, n, Tsh, (uint32_t)seed, (uint32_t)(seed >> 32), sample_offset, z, traj, \
                z_out, status, x0_out, map);                                                                  \
        else                                                                                                  \
            rollout_kernel<TSH, ZI, ZO, false><<<grid, RO_THREADS, smem, st>>>(                               \
                mean, cov, cross, Tm, x0, n, Tsh, (uint32_t)seed, (uint32_t)(seed >> 32), sample_offset, z, traj, \
                z_out, status, x0_out, map);                                                                  \
    } while (0)
#define CCMPC_RO_PICK(TSH)                                                                                    \
    do {                                                                                                      \
        if (z && z_out) CCMPC_RO_LAUNCH(TSH, true, true);                                                     \
        else if (z) CCMPC_RO_LAUNCH(TSH, true, false);                                                        \
        else if (z_out) CCMPC_RO_LAUNCH(TSH, false, true);                                                    \
        else CCMPC_RO_LAUNCH(TSH, false, false);                                                              \
    } while (0)
    if (Tsh == 16) CCMPC_RO_PICK(16);
    else if (Tsh == 8) CCMPC_RO_PICK(8);
    else CCMPC_RO_PICK(0);
#undef CCMPC_RO_PICK
#undef CCMPC_RO_LAUNCH
    return check_launch("rollout_kernel");
}

extern "C" int ccmpc_philox4x32(const uint32_t* ctr, const uint32_t* key, int64_t M, uint32_t* out, void* stream) {
    CCMPC_REQUIRE(M >= 0 && ctr && key && out, "ccmpc_philox4x32: bad arguments");
    if (M == 0) return CCMPC_OK;
    philox_kernel<<<(unsigned)((M + 127) / 128), 128, 0, (cudaStream_t)stream>>>(ctr, key, M, out);
    return check_launch("philox_kernel");
}
