// philox.cuh - counter-based random numbers shared by the sampling kernels of libccmpc
// (rollout.cu: predict_ideal, v8ideal:2699-2700; gmm.cu: GMM2D sampling of the predictor's output).
// Philox-4x32-10 (Salmon et al., SC'11; known answers in tests/test_gpu_parity.py) + a float32 Box-Muller.
#pragma once
#include <stdint.h>

namespace ccmpc {

struct Philox {
    uint32_t c[4];
};

__device__ __forceinline__ Philox philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0;
        k1 += W1;
    }
    return Philox{{c0, c1, c2, c3}};
}

__device__ __forceinline__ float lg2_approx(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float sqrt_approx(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Box-Muller on two 32-bit words: u1 in (0, 1] (>= 2^-33, never denormal), angle uniform in [-pi, pi).
// MUFU-only transcendentals (lg2 / sqrt / sin / cos approximations, ~1e-6 relative): sampling noise, not parity.
__device__ __forceinline__ void box_muller(uint32_t r0, uint32_t r1, float& z0, float& z1) {
    const float u1 = fmaf(__uint2float_rn(r0), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
    const float rad = sqrt_approx(-1.3862943611198906f * lg2_approx(u1));    // sqrt(-2 ln u1)
    const float ang = __int2float_rn((int)r1) * 1.4629180792671596e-9f;       // pi / 2^31
    z0 = rad * __cosf(ang);
    z1 = rad * __sinf(ang);
}

}  // namespace ccmpc
