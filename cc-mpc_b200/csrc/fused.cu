// fused.cu - K1 + K2 + K6 in one kernel: Monte Carlo rollout, trajectory store and the one-pass sums.
//
// ccmpc_mc_rollout writes the (C, n, T, 2) trajectory tensor and ccmpc_sample_sums reads it straight back
// (32 B of HBM traffic per trajectory-step).  Here the samples go through shared memory once: they are
// written to HBM from there (16 B per trajectory-step) and the SAME tile feeds the FP64 mma.sync Gram
// update and the A9 predicate.
//
//   rollout:  predict_ideal body v8ideal:2662-2709 - same Philox counters and arithmetic as rollout.cu
//             (sample i of cell c is bit-identical to ccmpc_mc_rollout's)
//   sums:     np.mean / np.cov of every (t, tau) pair (makeconstraint:55, v8ideal:2586-2606) + A9 counts,
//             same fragment layout, predicate order and packed layout as stats.cu (sums_common.cuh)
//
// WARP-SPECIALISED, ONE CTA PER SM (round 2).  What binds this kernel is the FP64 pipe, and
// tools/fp64_sched_bench.cu (profiles/r02_fp64_sched_bench.txt) shows how it is shared on B200:
//   * one warp per sub-partition issuing mma.sync.m8n8k4.f64 (DMMA) saturates it: 16.3 clk per DMMA;
//   * a vector FP64 instruction of ANOTHER warp only gets in between two DMMAs: ~25 clk each, whatever the
//     instruction-level parallelism of that warp, while the DMMA stream loses only ~7 %;
//   * two DMMA warps on one sub-partition halve each other, and the vector FP64 of one (shift, S1, predicate)
//     waits for the other's DMMAs.
// So: exactly ONE consumer warp per sub-partition owns the Gram update (warps 0-3, `setmaxnreg.inc` to 112
// registers); the only vector FP64 instructions in its in-order stream are the pivot shift and S1.  Everything
// else runs in 20 producer warps (`setmaxnreg.dec` to 72): the rollout - a long dependent chain of Philox
// (integer), Box-Muller (MUFU) and 7 FMAs per step that does not mind waiting for DMMA gaps - and the A9
// predicate, behind a screen on the INTEGER pipe: a sample whose x or y is at least 4 m (2^k > R) away from the
// ego cannot be a hit, which an unsigned window on the hi words of the doubles decides; the exact fp64 predicate (the
// oracle's operation order, no FMA) runs only for the warp-steps that hold a near sample, and the counts live in
// producer registers (lane t owns timestep t).  A tile = 32 samples x 8 or 16 timesteps, one sample per
// producer lane, kept as sub-tiles of [32 rows][128 B] in the TMA 128-byte swizzle (16-byte chunk j of row i
// sits at chunk j ^ (i & 7)): the lane-per-row writes of one timestep spread over all banks AND the tile is the
// image a TMA tensor store un-swizzles into the row-major (C, n, T, 2) trajectory tensor.  Tiles travel through
// a ring of 24 shared-memory buffers guarded by mbarriers (full: producer -> consumer; empty: consumer + the
// producer's own "TMA finished reading" -> next producer).  ONE lane per producer warp stores the finished tile
// with one cp.async.bulk.tensor (UTMASTG) per 4 KB sub-tile; rows past the end of the cell and columns past the
// horizon are clipped by the tensor map.
//
// Work split: the tiles of all cells form one list, CTA j takes the j-th contiguous share of it; where a share
// crosses a cell boundary the CTA finishes the cell (partial sums vector out, accumulators cleared, step
// parameters of the next cell built) and carries on: every SM gets the same number of tiles whatever C is.
//
// Measured on the way (c3, one B200; profiles/r02_fused_notes.txt): round-1 kernel 0.817 ms; warp-specialised
// with one cp.async.bulk per LANE 0.977 ms (the instruction takes uniform operands, the compiler serialises
// it over the lanes: 18 % of all warp instructions); one TMA tensor store per tile, two CTAs per SM 0.735 ms;
// this form 0.74 ms with even tile shares for any number of cells.  Storing the trajectories straight from
// registers (no raw tile) caps the kernel at 0.63 ms of scattered 32-byte stores: they have to leave through TMA.
//
// Restrictions of this fast path: Philox noise, no yaw sums.  Any horizon 1..16 (tiles are padded to 8 or 16 steps).
#include <stdlib.h>
#include <string.h>
#include "rollout.cuh"
#include "sums_common.cuh"
#include "tma.cuh"

namespace ccmpc {

#ifndef WS_CONS_DEF
#define WS_CONS_DEF 4
#endif
constexpr int WS_CONS = WS_CONS_DEF;                  // consumer warps (whole warpgroups: 1 or 2 per sub-partition)
#ifndef WS_PROD_DEF
#define WS_PROD_DEF 20
#endif
constexpr int WS_PROD = WS_PROD_DEF;                  // producer warps (whole warpgroups)
constexpr int WS_THREADS = 32 * (WS_CONS + WS_PROD);
#ifndef WS_NB_DEF
#define WS_NB_DEF 24
#endif
constexpr int WS_NB = WS_NB_DEF;                      // tile buffers in the ring
// launch allocation 80 registers x 768 threads; 128 x 112 + 640 x 72 fits it
#ifndef WS_REGS_PROD_DEF
#define WS_REGS_PROD_DEF ((WS_CONS == 4 && WS_PROD <= 20) ? 72 : 64)
#endif
constexpr int WS_REGS_CONS = 112, WS_REGS_PROD = WS_REGS_PROD_DEF;
static_assert(WS_PROD % 4 == 0, "setmaxnreg works on whole warpgroups");

__device__ __forceinline__ void cta_bar() { asm volatile("bar.sync 2, %0;" ::"n"(WS_THREADS) : "memory"); }

// EXP: compile-time experiment switches (tools/fused_time.py <exp>): which part of the kernel binds?
//   bit0 no Gram update, bit1 no A9 predicate, bit2 no Philox / Box-Muller, bit3 no chain, bit4 no pivot shift,
//   bit6 no S1 adds (16 + 64: the consumer is LDS + DMMA only).  0 in production.
static int g_ws_exp = 0;          // read only when built with -DCCMPC_WS_EXPERIMENTS
[[maybe_unused]] static int* const g_ws_exp_ref = &g_ws_exp;

template <int NF>
struct WsLayout {                                     // dynamic shared memory of rollout_sums_ws_kernel
    static constexpr int TSH = 4 * NF;
    static constexpr int H = TSH / 8;                 // 128-byte sub-tiles per tile
    static constexpr size_t SUB = 32 * 128;           // [32 rows][8 double2], 128-byte swizzle
    static constexpr size_t TILE = H * SUB;
    static constexpr size_t O_RING = 0;               // 1024-byte aligned: the swizzle pattern is address based
    // consumer epilogue areas: NOT aliased onto the ring - a consumed tile may still be read by its TMA store
    static constexpr size_t O_AREA = O_RING + WS_NB * TILE;
    static constexpr size_t O_SP = O_AREA + (size_t)WS_CONS * SumsWarpArea<NF>::PW * 8;   // StepParams[TSH]
    static constexpr size_t O_BAR = O_SP + (size_t)TSH * sizeof(StepParams);   // full[NB], empty[NB]
    static constexpr size_t O_X0 = O_BAR + 2 * WS_NB * 8;
    static constexpr size_t O_PIV = O_X0 + 16;                                  // double2 pivot[TSH]
    static constexpr size_t O_EGO = O_PIV + (size_t)TSH * 16;                   // double2 ego[TSH]
    static constexpr size_t O_SCR = O_EGO + (size_t)TSH * 16;                   // uint4 screen[TSH]: lo_x, w_x, lo_y, w_y
    static constexpr size_t O_PCNT = O_SCR + (size_t)TSH * 16;                  // producer counts [WS_PROD][PC_LEN]
    static constexpr int PC_LEN = 2 * TSH + 2;                                  // coll[T], pen[T], any, pad
    static constexpr size_t BYTES = O_PCNT + (size_t)WS_PROD * PC_LEN * 8;
    static_assert(BYTES <= 227 * 1024, "shared memory of one CTA");
};

// grid G (<= number of SMs): CTA j handles global tiles [j TT / G, (j + 1) TT / G), TT = C ceil(n / 32).
// partial sums of the segment (CTA j, cell c) go to slot j + c of `partials`, `slot_cell[j + c] = c`.
template <int NF, bool FULLT, int EXP = 0>
__global__ void __launch_bounds__(WS_THREADS, 1)
rollout_sums_ws_kernel(const double* __restrict__ mean, const double* __restrict__ cov,
                       const double* __restrict__ cross, int Tm, const double* __restrict__ x0, int64_t n,
                       uint32_t seed_lo, uint32_t seed_hi, uint64_t sample_offset,
                       const double* __restrict__ pivot, const double* __restrict__ ego, int64_t ego_cell_stride,
                       double r2, const __grid_constant__ CUtensorMap traj_map, int store_traj,
                       double* __restrict__ partials, int32_t* __restrict__ slot_cell, uint8_t* __restrict__ status,
                       double* __restrict__ x0_out, int64_t C, int T_rt) {
    // T <= TSH = 4 NF timesteps are live (the shrinking horizon runs predict_ideal at T = ph - 1, ph - 2, ...):
    // the columns past 2 T of every tile stay zero (cleared once below, never written), so they add nothing to
    // the Gram update, and the tensor map clips them from the store.
    using WL = WsLayout<NF>;
    using SM = SumsWarpArea<NF>;
    constexpr int TSH = WL::TSH, TS = NF / 2, NT = SM::NT;
    constexpr int exp_flags = EXP;
    const int T = FULLT ? TSH : T_rt;               // FULLT: the horizon is a compile-time constant (ph8 / ph16)
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* ring = smem_raw + WL::O_RING;
    double* area = reinterpret_cast<double*>(smem_raw + WL::O_AREA);
    StepParams* sp = reinterpret_cast<StepParams*>(smem_raw + WL::O_SP);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + WL::O_BAR);
    uint64_t* empty = full + WS_NB;
    double* s_x0 = reinterpret_cast<double*>(smem_raw + WL::O_X0);
    double2* s_piv = reinterpret_cast<double2*>(smem_raw + WL::O_PIV);
    double2* s_ego = reinterpret_cast<double2*>(smem_raw + WL::O_EGO);
    uint4* s_scr = reinterpret_cast<uint4*>(smem_raw + WL::O_SCR);
    double* s_pcnt = reinterpret_cast<double*>(smem_raw + WL::O_PCNT);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) {
        for (int b = 0; b < WS_NB; ++b) { mbar_init(full + b, 1); mbar_init(empty + b, 2); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < (int)(WS_NB * WL::TILE / 16); i += WS_THREADS)
        reinterpret_cast<double2*>(ring)[i] = make_double2(0.0, 0.0);
    const int64_t ntc = (n + 31) >> 5;                                  // tiles per cell
    const int64_t TT = C * ntc;
    const int64_t g_lo = (TT * blockIdx.x) / gridDim.x, g_hi = (TT * (blockIdx.x + 1)) / gridDim.x;
    __syncthreads();

    if (warp >= WS_CONS) {
        // ================================================================== producers
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(WS_REGS_PROD));
        const int p = warp - WS_CONS;
        const int sw = lane & 7;
        int prev_b = -1;
        auto retire_prev = [&]() {       // the previous tile's TMA store has read its rows: the buffer may be refilled
            if (prev_b >= 0) {
                if (lane == 0) {
                    bulk_wait_read<0>();
                    mbar_arrive(empty + prev_b);
                }
                prev_b = -1;
            }
        };
        int64_t seq0 = 0;                                               // tiles of earlier segments of this CTA
        for (int64_t gt = g_lo; gt < g_hi;) {
            const int64_t c = gt / ntc, t0 = gt - c * ntc;
            const int cnt = (int)min(g_hi - gt, ntc - t0);
            cta_bar();                                                  // (S1) the previous segment is drained
            cta_bar();                                                  // (S2) parameters of cell c are built
            // A9 counts of this warp in this segment: lane t owns timestep t (lanes >= T idle)
            int coll_mine = 0, any_cnt = 0;
            double pen_mine = 0.0;
            const int first = (int)((p - seq0 % WS_PROD + WS_PROD) % WS_PROD);   // my first tile of this segment
            for (int s = first; s < cnt; s += WS_PROD) {
                const int64_t seq = seq0 + s;
                const int b = (int)(seq % WS_NB);
                const uint32_t use = (uint32_t)(seq / WS_NB);
                retire_prev();                                          // hand my previous buffer back first
                mbar_wait(empty + b, (use & 1) ^ 1);                    // use 0 passes at once
                const int64_t i0 = (t0 + s) << 5;
                const bool valid = i0 + lane < n;
                // sample (row = lane, timestep t): sub-tile t >> 3, chunk (t & 7) ^ (lane & 7)
                unsigned char* row = ring + (size_t)b * WL::TILE + lane * 128;
                double x = s_x0[0], y = s_x0[1];
                const uint64_t gi = sample_offset + (uint64_t)(i0 + lane);
                const uint32_t gi_lo = (uint32_t)gi, gi_hi = (uint32_t)(gi >> 32);
                bool hit_mine = false;
                // one step: chain, shifted copy for the Gram update, A9 screen (integer pipe) and, for warps that
                // hold a sample within `box` of the ego in x and y, the exact predicate in the oracle's order
                // the chain: one straight-line block per tile (8 Philox blocks + Box-Muller on the integer / MUFU
                // pipes, 16 x 7 FMAs that each wait for a gap in the consumer's DMMA stream) - no votes or branches
                // in it, so the scheduler is free to fill the FP64 waits with the next steps' noise
                unsigned near_mask = 0;                 // bit t: this sample is within `box` of the ego at step t
                auto advance = [&](int t, double zx, double zy) {
                    if (!(exp_flags & 8)) rollout_step(sp[t], x, y, zx, zy);
                    *reinterpret_cast<double2*>(row + (t >> 3) * WL::SUB + (((t & 7) ^ sw) << 4)) = make_double2(x, y);
                    if (!(exp_flags & 2)) {             // A9 screen: integer pipe, no branch (see below)
                        const uint4 k = s_scr[t];
                        near_mask |= (unsigned)(((uint32_t)__double2hiint(x) - k.x <= k.y) &
                                                ((uint32_t)__double2hiint(y) - k.z <= k.w)) << t;
                    }
                };
#pragma unroll
                for (int t2 = 0; t2 < TSH / 2; ++t2) {
                    if (2 * t2 >= T) break;
                    float za = 0.3f, zb = -0.2f, zc = 0.1f, zd = 0.7f;
                    if (!(exp_flags & 4)) {
                        Philox rr = philox4x32_10(gi_lo, gi_hi, (uint32_t)c, (uint32_t)t2, seed_lo, seed_hi);
                        box_muller(rr.c[0], rr.c[1], za, zb);
                        box_muller(rr.c[2], rr.c[3], zc, zd);
                    }
                    advance(2 * t2, (double)za, (double)zb);
                    if (2 * t2 + 1 < T) advance(2 * t2 + 1, (double)zc, (double)zd);
                }
                // absent rows (last tile of a cell) hold the pivot: u = 0 in the Gram update; the tensor map clips them
                // from the store
                if (!valid) {
                    for (int t = 0; t < T; ++t)
                        *reinterpret_cast<double2*>(row + (t >> 3) * WL::SUB + (((t & 7) ^ sw) << 4)) = s_piv[t];
                }
                // A9: the chain screened every step on the integer pipe (a sample at least `box` away from the ego in
                // x or y cannot be a hit); one warp-wide OR of the per-step flags, then the exact fp64 predicate (the
                // oracle's operation order) only for the flagged steps, on the values this lane wrote to its row
                if (!(exp_flags & 2)) {
                    unsigned m = __reduce_or_sync(0xffffffffu, valid ? near_mask : 0u);
                    while (m) {
                        const int t = __ffs(m) - 1;
                        m &= m - 1;
                        const double2 v = *reinterpret_cast<const double2*>(row + (t >> 3) * WL::SUB + (((t & 7) ^ sw) << 4));
                        const double2 e = s_ego[t];
                        const double dx = __dsub_rn(e.x, v.x), dy = __dsub_rn(e.y, v.y);
                        const double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
                        const bool hit = valid && d2 < r2;
                        const unsigned bal = __ballot_sync(0xffffffffu, hit);
                        if (bal) {
                            hit_mine |= hit;
                            double pn = hit ? r2 - d2 : 0.0;            // sum over the hit lanes, fixed order
#pragma unroll
                            for (int o = 16; o; o >>= 1) pn += __shfl_xor_sync(0xffffffffu, pn, o);
                            if (lane == t) { coll_mine += __popc(bal); pen_mine += pn; }
                        }
                    }
                }
                any_cnt += __popc(__ballot_sync(0xffffffffu, hit_mine));
                if (store_traj) fence_proxy_async();    // my generic-proxy writes -> visible to the TMA engine
                __syncwarp();
                if (lane == 0) {
                    if (store_traj) {
#pragma unroll
                        for (int h = 0; h < WL::H; ++h)
                            if (8 * h < T)
                                tma_store_box(&traj_map, ring + (size_t)b * WL::TILE + h * WL::SUB, 16 * h, (int)i0, (int)c);
                        bulk_commit();
                    }
                    mbar_arrive(full + b);
                }
                prev_b = b;
            }
            retire_prev();
            // this warp's counts of the segment -> shared memory (the consumers fold them into the partial vector)
            double* pc = s_pcnt + (size_t)p * WL::PC_LEN;
            if (lane < TSH) { pc[lane] = (double)coll_mine; pc[TSH + lane] = pen_mine; }      // zero for lanes >= T
            if (lane == 0) pc[2 * TSH] = (double)any_cnt;
            cta_bar();                                                  // (S3) counts visible to the consumers
            seq0 += cnt;
            gt += cnt;
        }
        if (lane == 0) bulk_wait_all();                 // every tile has landed before the CTA retires
        return;
    }

    // ====================================================================== consumers
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(WS_REGS_CONS));
    constexpr int PERM = NF == 4 ? 1 : 2;               // lane group -> column group (sums_common.cuh: bank spreading)
    const int r = lane & 3, q = lane >> 2, cg = col_group<PERM>(q);
    int ld_off[TS];                                     // byte offset of my chunk inside its 128-byte row (even g)
#pragma unroll
    for (int i = 0; i < TS; ++i) ld_off[i] = (((TS * cg + i) & 7) ^ r) << 4;
    // screening box 2^k > R: |dx| >= 2^k or |dy| >= 2^k implies d2 >= r2
    const double box = __hiloint2double((__double2hiint(sqrt(r2)) & 0x7ff00000) + 0x00100000, 0);
    const SumsLayout L(T);
    int64_t seq0 = 0;
    for (int64_t gt = g_lo; gt < g_hi;) {
        const int64_t c = gt / ntc, t0 = gt - c * ntc;
        const int cnt = (int)min(g_hi - gt, ntc - t0);
        cta_bar();                                      // (S1) producers are done with the previous cell's parameters
        rollout_setup<32 * WS_CONS>(mean, cov, cross, Tm, x0, c, T, seed_lo, seed_hi, sp, s_x0, status, x0_out, tid,
                                    t0 == 0);
        if (tid < T) {                                  // what the producers need per timestep besides the chain
            const int t = tid;
            const double px = pivot[(c * T + t) * 2], py = pivot[(c * T + t) * 2 + 1];
            const double ex = ego[c * ego_cell_stride + 2 * t], ey = ego[c * ego_cell_stride + 2 * t + 1];
            s_piv[t] = make_double2(px, py);
            s_ego[t] = make_double2(ex, ey);
            // integer screen on the raw hi words: hi(x) - lo <= w (unsigned) whenever |ex - x| < box.  The hi word
            // of a double ascends with the value for positive numbers and with the magnitude for negative ones, so
            // the window (ex - box, ex + box) maps to one interval of hi words when it lies on one side of zero
            // (the other sign is >= 2^20 away after the unsigned wrap); a window that holds zero flags every
            // sample (the exact predicate decides).  No false negatives; ties in the low word count as near.
            auto window = [&](double e, uint32_t& lo, uint32_t& w) {
                const double a = e - box, b = e + box;
                if (a > 0.0) { lo = (uint32_t)__double2hiint(a); w = (uint32_t)__double2hiint(b) - lo; }
                else if (b < 0.0) { lo = (uint32_t)__double2hiint(b); w = (uint32_t)__double2hiint(a) - lo; }
                else { lo = 0u; w = 0xffffffffu; }
            };
            uint4 k;
            window(ex, k.x, k.y);
            window(ey, k.z, k.w);
            s_scr[t] = k;
        }
        double piv[NF];                                 // pivot of my columns NF q .. NF q + NF - 1
#pragma unroll
        for (int i = 0; i < NF; ++i) piv[i] = NF * cg + i < 2 * T ? pivot[c * 2 * T + NF * cg + i] : 0.0;
        double acc[NT][2];
#pragma unroll
        for (int i = 0; i < NT; ++i) acc[i][0] = acc[i][1] = 0.0;
        double s1[NF];
#pragma unroll
        for (int i = 0; i < NF; ++i) s1[i] = 0.0;
        cta_bar();                                      // (S2) parameters visible to the producers
        const int first = (int)((warp - seq0 % WS_CONS + WS_CONS) % WS_CONS);
        for (int s = first; s < cnt; s += WS_CONS) {
            const int64_t seq = seq0 + s;
            const int b = (int)(seq % WS_NB);
            const uint32_t use = (uint32_t)(seq / WS_NB);
            mbar_wait(full + b, use & 1);
            // lane (r, q) reads timesteps TS cg + i of row 4 g + r: sub-tile t >> 3, chunk (t & 7) ^ (row & 7);
            // row & 7 = 4 (g & 1) + r, so the two parities of g differ by bit 6 of the byte offset.
            // Fragments = loads minus the pivot (the only vector FP64 work left in this warp besides S1).
            const unsigned char* tile = ring + (size_t)b * WL::TILE + ((TS * cg) >> 3) * WL::SUB + r * 128;
            double f[2][2][NF];                         // [buffer][group of the pair][fragment]
            auto load_pair = [&](int g2, int buf) {
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int i = 0; i < TS; ++i) {
                        const double2 v = *reinterpret_cast<const double2*>(tile + g2 * 1024 + h * 512 + (ld_off[i] ^ (h << 6)));
                        f[buf][h][2 * i] = (exp_flags & 16) ? v.x : v.x - piv[2 * i];
                        f[buf][h][2 * i + 1] = (exp_flags & 16) ? v.y : v.y - piv[2 * i + 1];
                    }
            };
            load_pair(0, 0);
#pragma unroll
            for (int g2 = 0; g2 < 4; ++g2) {
                const int buf = g2 & 1;
                if (g2 < 3) load_pair(g2 + 1, buf ^ 1);  // in flight during this pair's 2 x NT DMMAs
#pragma unroll
                for (int h = 0; h < 2; ++h) {
#pragma unroll
                    for (int i = 0; i < NF; ++i) if (!(exp_flags & 64)) s1[i] += f[buf][h][i];
                    if (!(exp_flags & 1)) gram_update<NF>(acc, f[buf][h]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + b);
        }
        // ---- warp epilogue of the segment (as stats.cu)
        double* wsm = area + (size_t)warp * SM::PW;
        store_gram_tiles<NF>(wsm, acc, q, r);
#pragma unroll
        for (int i = 0; i < NF; ++i) {
            double v = s1[i];
            v += __shfl_xor_sync(0xffffffffu, v, 1);
            v += __shfl_xor_sync(0xffffffffu, v, 2);
            if (r == 0) wsm[SM::O_S1 + NF * cg + i] = v;           // column NF cg + i = 2 t + {x, y}
        }
        cta_bar();                                      // (S3) consumer areas and producer counts are written
        // ---- partial vector of segment (CTA, cell): consumer warps in warp order, then the producers' counts
        const int64_t slot = (int64_t)blockIdx.x + c;
        double* out = partials + slot * (int64_t)L.len;
        // rows of this segment: whole tiles except the last one of the cell
        const int64_t rows_seg = min((t0 + cnt) << 5, n) - (t0 << 5);
        for (int e = tid; e < L.len; e += 32 * WS_CONS) {
            double sum = 0.0;
            if (e == 0) {
                sum = (double)rows_seg;
            } else if (e < L.o_coll) {
                const int src = packed_to_area<NF, PERM>(e, L, false, false);
#pragma unroll
                for (int w = 0; w < WS_CONS; ++w) sum += area[(size_t)w * SM::PW + src];
            } else if (e < L.o_d2s) {                   // coll[T], any, pen[T]: producer counts, warp order
                const int k = e < L.o_any ? e - L.o_coll : (e == L.o_any ? 2 * TSH : TSH + (e - L.o_pen));
                for (int w = 0; w < WS_PROD; ++w) sum += s_pcnt[(size_t)w * WL::PC_LEN + k];
            }                                           // d2s: Gram identity in the finalize kernel; no yaw sums
            out[e] = sum;
        }
        if (tid == 0) slot_cell[slot] = (int32_t)c;
        seq0 += cnt;
        gt += cnt;
    }
}

// sums[c] = sum over the cell's segment partials (slots c .. c + G - 1 that carry cell c), in CTA order; the d2s
// slots follow from the Gram identity
__global__ void fused_finalize_kernel(const double* __restrict__ partials, const int32_t* __restrict__ slot_cell,
                                      int G, int T, const double* __restrict__ pivot, const double* __restrict__ ego,
                                      int64_t ego_cell_stride, double* __restrict__ sums) {
    extern __shared__ double s_sum[];
    __shared__ int s_list[256], s_cnt;                  // the cell's slots, ascending (G <= 256 SMs)
    const SumsLayout L(T);
    const int64_t c = blockIdx.x;
    __shared__ unsigned s_bal[8];
    {                                                   // ordered compaction of the matching slots (warps 0-7 cover G)
        const int g = threadIdx.x, w = g >> 5, ln = g & 31;
        const bool flag = g < G && slot_cell[c + g] == (int32_t)c;
        const unsigned bal = __ballot_sync(0xffffffffu, flag);
        if (w < 8 && ln == 0) s_bal[w] = bal;
        __syncthreads();
        if (w < 8) {
            int pos = __popc(bal & ((1u << ln) - 1u));
            for (int k = 0; k < w; ++k) pos += __popc(s_bal[k]);
            if (flag) s_list[pos] = g;
        }
        if (g == 0) {
            int tot = 0;
            for (int k = 0; k < 8; ++k) tot += __popc(s_bal[k]);
            s_cnt = tot;
        }
    }
    __syncthreads();
    const int cnt = s_cnt;
    for (int j = threadIdx.x; j < L.len; j += blockDim.x) {
        double s = 0.0;
#pragma unroll 4
        for (int k = 0; k < cnt; ++k) s += __ldcs(partials + (c + s_list[k]) * (int64_t)L.len + j);
        s_sum[j] = s;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < L.len; j += blockDim.x) {
        double v = s_sum[j];
        if (j >= L.o_d2s && j < L.o_yaw) {
            const int t = j - L.o_d2s;
            v = d2s_from_sums(s_sum, L, t, ego[c * ego_cell_stride + 2 * t] - pivot[(c * T + t) * 2],
                              ego[c * ego_cell_stride + 2 * t + 1] - pivot[(c * T + t) * 2 + 1]);
        }
        sums[c * L.len + j] = v;
    }
}

static int fused_grid(int64_t n, int64_t C) {
    const int64_t tt = C * ((n + 31) / 32);
    const int64_t sms = sm_count() < 256 ? sm_count() : 256;        // fused_finalize_kernel lists at most 256 slots
    const int64_t g = tt < sms ? tt : sms;
    return (int)(g < 1 ? 1 : g);
}

static size_t slot_bytes(int64_t slots) { return ((size_t)slots * sizeof(int32_t) + 255) / 256 * 256; }

}  // namespace ccmpc

using namespace ccmpc;

extern "C" int ccmpc_debug_ws_exp(int flags) {      // not part of the ABI (tools only)
    g_ws_exp = flags;
    return 0;
}

extern "C" size_t ccmpc_mc_rollout_sums_workspace_bytes(int64_t C, int64_t n, int Tsh) {
    if (C <= 0 || n <= 0 || Tsh < 1 || Tsh > CCMPC_MAX_T) return 16;
    const int64_t slots = C + sm_count();           // segment (CTA j, cell c) -> slot j + c
    return slot_bytes(slots) + (size_t)slots * (size_t)SumsLayout(Tsh).len * sizeof(double);
}

// (2 T doubles, n samples, C cells) view of the row-major trajectory tensor, boxes of [32 samples][16 doubles] in
// the 128-byte swizzle.  cuTensorMapEncodeTiled is a host-side encoder of the driver: fetched through the runtime
// (no link against libcuda), called once per launch (the map holds the tensor's address and extents).
int ccmpc::make_traj_map(CUtensorMap* map, double* traj, int64_t C, int64_t n, int Tsh) {
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
        if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
            set_error("cuTensorMapEncodeTiled is not available (%s)", cudaGetErrorString(e));
            return CCMPC_E_LAUNCH;
        }
        encode = reinterpret_cast<encode_fn>(fn);
    }
    const cuuint64_t dims[3] = {(cuuint64_t)(2 * Tsh), (cuuint64_t)n, (cuuint64_t)C};
    const cuuint64_t strides[2] = {(cuuint64_t)(2 * Tsh) * 8, (cuuint64_t)n * (2 * Tsh) * 8};
    const cuuint32_t box[3] = {16, 32, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, traj, dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (CUresult %d; C = %lld, n = %lld)", (int)r,
                  (long long)C, (long long)n);
        return CCMPC_E_LAUNCH;
    }
    return CCMPC_OK;
}

template <int NF>
static int launch_ws(int G, cudaStream_t st, const double* mean, const double* cov, const double* cross, int Tm,
                     const double* x0, int64_t C, int64_t n, uint64_t seed, uint64_t sample_offset,
                     const double* pivot, const double* ego, int64_t ego_cell_stride, double R, double* traj,
                     double* partials, int32_t* slot_cell, uint8_t* status, double* x0_out, int Tsh) {
    auto kernel = Tsh == 4 * NF ? rollout_sums_ws_kernel<NF, true, 0> : rollout_sums_ws_kernel<NF, false, 0>;
#ifdef CCMPC_WS_EXPERIMENTS
    if (g_ws_exp == 1) kernel = rollout_sums_ws_kernel<NF, true, 1>;
    if (g_ws_exp == 2) kernel = rollout_sums_ws_kernel<NF, true, 2>;
    if (g_ws_exp == 3) kernel = rollout_sums_ws_kernel<NF, true, 3>;
    if (g_ws_exp == 12) kernel = rollout_sums_ws_kernel<NF, true, 12>;
    if (g_ws_exp == 16) kernel = rollout_sums_ws_kernel<NF, true, 16>;
    if (g_ws_exp == 28) kernel = rollout_sums_ws_kernel<NF, true, 28>;
    if (g_ws_exp == 64) kernel = rollout_sums_ws_kernel<NF, true, 64>;
    if (g_ws_exp == 5) kernel = rollout_sums_ws_kernel<NF, true, 5>;
    if (g_ws_exp == 7) kernel = rollout_sums_ws_kernel<NF, true, 7>;
    if (g_ws_exp == 85) kernel = rollout_sums_ws_kernel<NF, true, 85>;
    if (g_ws_exp == 87) kernel = rollout_sums_ws_kernel<NF, true, 87>;
    if (g_ws_exp == 4) kernel = rollout_sums_ws_kernel<NF, true, 4>;
    if (g_ws_exp == 8) kernel = rollout_sums_ws_kernel<NF, true, 8>;
    if (g_ws_exp == 6) kernel = rollout_sums_ws_kernel<NF, true, 6>;
    if (g_ws_exp == 10) kernel = rollout_sums_ws_kernel<NF, true, 10>;
    if (g_ws_exp == 14) kernel = rollout_sums_ws_kernel<NF, true, 14>;
    if (g_ws_exp == 84) kernel = rollout_sums_ws_kernel<NF, true, 84>;
    if (g_ws_exp == 88) kernel = rollout_sums_ws_kernel<NF, true, 88>;
    if (g_ws_exp == 80) kernel = rollout_sums_ws_kernel<NF, true, 80>;
    if (g_ws_exp == 92) kernel = rollout_sums_ws_kernel<NF, true, 92>;
#endif
    {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)WsLayout<NF>::BYTES);
        if (e != cudaSuccess) { set_error("ccmpc_mc_rollout_sums: smem attribute: %s", cudaGetErrorString(e)); return CCMPC_E_LAUNCH; }
    }
    CUtensorMap map;
    memset(&map, 0, sizeof(map));
    if (traj) {
        int rc = make_traj_map(&map, traj, C, n, Tsh);
        if (rc != CCMPC_OK) return rc;
    }
    kernel<<<(unsigned)G, WS_THREADS, WsLayout<NF>::BYTES, st>>>(
        mean, cov, cross, Tm, x0, n, (uint32_t)seed, (uint32_t)(seed >> 32), sample_offset, pivot, ego,
        ego_cell_stride, R * R, map, traj ? 1 : 0, partials, slot_cell, status, x0_out, C, Tsh);
    return check_launch("rollout_sums_ws_kernel");
}

extern "C" int ccmpc_mc_rollout_sums(const double* mean, const double* cov, const double* cross, int Tm,
                                     const double* x0, int64_t C, int64_t n, int Tsh, uint64_t seed,
                                     uint64_t sample_offset, const double* pivot, const double* ego,
                                     int64_t ego_cell_stride, double R, double* traj, double* sums, uint8_t* status,
                                     double* x0_out, void* workspace, size_t workspace_bytes, void* stream) {
    CCMPC_REQUIRE(Tsh >= 1 && Tsh <= CCMPC_MAX_T, "ccmpc_mc_rollout_sums: need 1 <= Tsh <= %d", CCMPC_MAX_T);
    CCMPC_REQUIRE(Tm >= Tsh + 1, "ccmpc_mc_rollout_sums: moments must cover Tsh + 1 steps (Tm = %d)", Tm);
    CCMPC_REQUIRE(C >= 0 && n >= 0 && C < 65536 && n < ((int64_t)1 << 31), "ccmpc_mc_rollout_sums: bad C / n");
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
    const int G = fused_grid(n, C);
    const int64_t slots = C + sm_count();
    int32_t* slot_cell = reinterpret_cast<int32_t*>(workspace);
    double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(workspace) + slot_bytes(slots));
    {
        cudaError_t e = cudaMemsetAsync(slot_cell, 0xFF, (size_t)slots * sizeof(int32_t), st);      // -1: empty slot
        if (e != cudaSuccess) { set_error("ccmpc_mc_rollout_sums: memset: %s", cudaGetErrorString(e)); return CCMPC_E_LAUNCH; }
    }
    int rc = Tsh > 8 ? launch_ws<4>(G, st, mean, cov, cross, Tm, x0, C, n, seed, sample_offset, pivot, ego,
                                    ego_cell_stride, R, traj, partials, slot_cell, status, x0_out, Tsh)
                     : launch_ws<2>(G, st, mean, cov, cross, Tm, x0, C, n, seed, sample_offset, pivot, ego,
                                    ego_cell_stride, R, traj, partials, slot_cell, status, x0_out, Tsh);
    if (rc != CCMPC_OK) return rc;
    fused_finalize_kernel<<<(unsigned)C, 704, (size_t)L.len * sizeof(double), st>>>(partials, slot_cell, G, Tsh, pivot,
                                                                                   ego, ego_cell_stride, sums);
    return check_launch("fused_finalize_kernel");
}
