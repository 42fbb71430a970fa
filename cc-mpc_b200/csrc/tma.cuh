// tma.cuh - mbarrier and bulk-tensor-copy (TMA) PTX used by fused.cu and rollout.cu, and the host-side tensor map of
// the trajectory tensor.
#pragma once
#include <cuda.h>          // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint)

#include "common.cuh"

namespace ccmpc {

// (2 T doubles, n samples, C cells) view of the row-major (C, n, T, 2) trajectory tensor with boxes of
// [32 samples][16 doubles] in the 128-byte swizzle.  Host side, called once per launch (fused.cu).
int make_traj_map(CUtensorMap* map, double* traj, int64_t C, int64_t n, int Tsh);

// ---- mbarrier / bulk-copy PTX ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(smem_u32(bar)) : "memory");
}
// try_wait suspends the warp until the phase completes or the time hint (ns) runs out, so a long wait costs a handful
// of polls instead of a spin: without the hint the producers' waits for a free tile were 24 % of all warp instructions
// the fused kernel issued (ncu source page: 3 x 27.6 M of 342 M on the poll loop).  The kernel time is the same with
// and without (0.676 ms: the polls went into issue slots nobody else wanted); the hint keeps the slots and the power.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\n"
        "WAIT_%=:\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
        " @p bra DONE_%=;\n"
        " bra WAIT_%=;\n"
        "DONE_%=:\n}" ::"r"(smem_u32(bar)), "r"(parity), "r"(0x989680u)
        : "memory");
}
// one [32 rows][16 doubles] box of the (2 T doubles, n samples, C cells) view of the trajectory tensor
__device__ __forceinline__ void tma_store_box(const CUtensorMap* map, const void* ssrc, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(map),
                 "r"(smem_u32(ssrc)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
}  // namespace ccmpc
