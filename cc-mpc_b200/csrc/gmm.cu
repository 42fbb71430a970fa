// gmm.cu - stage 1 of the path: Monte Carlo sampling of the predictor's GMM2D output on the device.
//
// Replaces, at the boundary of generate_vehicle_latents (collect/in_simulation/midlevel/prediction.py:76-104),
// the two sampling calls whose code lives in the Trajectron++ submodule that is MISSING from the reference
// tree (fireofearth/Trajectron-plus-plus @ 0e44e82, .SUBMODULES.json): `model.latent.sample_p` (:81-83, one
// categorical draw z ~ p(z | x) per sample) and the final `GMM2D.rsample()` of p_y_xz (:85-86).  The arithmetic
// is restated from the published Trajectron++ GMM2D (model/components/gmm2d.py):
//     log_pis <- log_pis - logsumexp(log_pis);  sigma = exp(log_sigma);  L = [[sx, 0], [rho sy, sy sqrt(1-rho^2)]]
//     component g ~ Cat(exp(log_pis));  y = mu_g + L_g eps,  eps ~ N(0, I)
// per (vehicle, sample, timestep), with the GMM parameters indexed by the sample's latent z.  Output is what
// prediction.py:103-104 returns: z [V,N] and float32 predictions [V,N,T,2] - the ingest format of
// ccmpc_group_by_mode - so samples never leave the device between the predictor and the constraint build.
// Parity is statistical (no reference code to run, torch draws from its own generator); a supplied-noise
// mode makes the arithmetic itself checkable against the oracle (oracle/restate.py: gmm2d_sample).
//
// One thread per (sample, one or two timestep pairs): one Philox-4x32-10 block gives the four normals of a
// pair, the latent draw is shared by the thread's pairs, the float4 stores of consecutive threads are contiguous.  HBM: 8 B written per (sample, timestep).
#include "common.cuh"
#include "philox.cuh"

namespace ccmpc {

constexpr int GMM_THREADS = 256;

struct GmmNoise {            // all three or none
    const float* u_z;        // [V,N]      uniform in [0,1): latent draw
    const float* u_c;        // [V,N,T]    uniform in [0,1): component draw
    const float* eps;        // [V,N,T,2]  standard normals
};

__device__ __forceinline__ float u01(uint32_t r) {          // [0, 1), 24 bits
    return (float)(r >> 8) * 5.9604644775390625e-8f;
}

// first k with u < cum_k over the normalised weights w_k / sum w (last bin catches round-off)
__device__ __forceinline__ int pick_latent(const float* __restrict__ s_cum, int Z, float u) {
    int z = Z - 1;
    for (int k = Z - 2; k >= 0; --k)
        if (u < s_cum[k]) z = k;
    return z;
}

template <bool VEC, int PAIRS>
__global__ void __launch_bounds__(GMM_THREADS)
gmm2d_sample_kernel(const double* __restrict__ latent_pmf, const float* __restrict__ log_pis,
                    const float* __restrict__ mus, const float* __restrict__ log_sigmas,
                    const float* __restrict__ corrs, int64_t N, int Z, int T, int G, uint64_t seed,
                    const uint64_t* __restrict__ seed_dev, uint64_t sample_offset, GmmNoise noise,
                    int32_t* __restrict__ z_out, float* __restrict__ pred) {
    extern __shared__ float s_cum[];                          // [Z] cumulative latent pmf of this vehicle
    const int v = blockIdx.y;
    // cumulative pmf: loads in parallel, the running sum in the oracle's sequential order (bit-identical bins)
    double* s_w = reinterpret_cast<double*>(s_cum + ((Z + 1) & ~1));
    for (int k = threadIdx.x; k < Z; k += GMM_THREADS) s_w[k] = fmax(latent_pmf[(int64_t)v * Z + k], 0.0);
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot = 0.0;
        for (int k = 0; k < Z; ++k) tot += s_w[k];
        double acc = 0.0;
        for (int k = 0; k < Z; ++k) {
            acc += s_w[k];
            s_cum[k] = (float)(acc / tot);
        }
    }
    __syncthreads();
    if (seed_dev) seed += *seed_dev;              // per-step key read on the device: a captured graph replays with new draws
    const uint32_t seed_lo = (uint32_t)seed, seed_hi = (uint32_t)(seed >> 32);
    const int TP = (T + 1) >> 1;
    const int TQ = (TP + PAIRS - 1) / PAIRS;          // work items per sample: PAIRS timestep pairs each
    for (int64_t i = (int64_t)blockIdx.x * GMM_THREADS + threadIdx.x; i < N * TQ; i += (int64_t)gridDim.x * GMM_THREADS) {
    const int64_t n = (N * TQ < ((int64_t)1 << 31)) ? (int64_t)((uint32_t)i / (uint32_t)TQ) : i / TQ;
    const int tq = (int)(i - n * TQ);
    const uint64_t gi = sample_offset + (uint64_t)n;
    const uint32_t gi_lo = (uint32_t)gi, gi_hi = (uint32_t)(gi >> 32);
    const int64_t vn = (int64_t)v * N + n;

    // latent of the sample: every thread of the sample derives the same value from the same counter
    float uz;
    if (noise.u_z) {
        uz = noise.u_z[vn];
    } else {
        Philox r = philox4x32_10(gi_lo, gi_hi, (uint32_t)v, 0xFFFFFFFEu, seed_lo, seed_hi);
        uz = u01(r.c[0]);
    }
    const int z = pick_latent(s_cum, Z, uz);
    if (tq == 0 && z_out) z_out[vn] = z;

#pragma unroll
    for (int pp = 0; pp < PAIRS; ++pp) {
    const int tp = tq * PAIRS + pp;
    if (tp >= TP) continue;
    float e[4] = {0.f, 0.f, 0.f, 0.f}, uc[2] = {0.f, 0.f};
    const int t0 = 2 * tp;
    if (noise.eps) {
#pragma unroll
        for (int j = 0; j < 2; ++j)
            if (t0 + j < T) {
                e[2 * j] = noise.eps[(vn * T + t0 + j) * 2];
                e[2 * j + 1] = noise.eps[(vn * T + t0 + j) * 2 + 1];
                uc[j] = noise.u_c[vn * T + t0 + j];
            }
    } else {
        Philox r = philox4x32_10(gi_lo, gi_hi, (uint32_t)v, (uint32_t)tp, seed_lo, seed_hi);
        box_muller(r.c[0], r.c[1], e[0], e[1]);
        box_muller(r.c[2], r.c[3], e[2], e[3]);
        if (G > 1) {
            Philox rc = philox4x32_10(gi_lo, gi_hi, (uint32_t)v, 0x80000000u | (uint32_t)tp, seed_lo, seed_hi);
            uc[0] = u01(rc.c[0]);
            uc[1] = u01(rc.c[1]);
        }
    }
    float out[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const int t = t0 + j;
        if (t >= T) continue;
        const int64_t row = (((int64_t)v * Z + z) * T + t) * G;
        int g = 0;
        if (G > 1) {
            // Cat(softmax(log_pis)): normalise as logsumexp does, then invert the CDF
            float mx = __ldg(log_pis + row);
            for (int k = 1; k < G; ++k) mx = fmaxf(mx, __ldg(log_pis + row + k));
            float tot = 0.f;
            for (int k = 0; k < G; ++k) tot += expf(__ldg(log_pis + row + k) - mx);
            const float target = uc[j] * tot;
            float acc = 0.f;
            g = G - 1;
            for (int k = 0; k < G - 1; ++k) {
                acc += expf(__ldg(log_pis + row + k) - mx);
                if (target < acc) { g = k; break; }
            }
        }
        const float2 mu = __ldg(reinterpret_cast<const float2*>(mus) + row + g);
        const float2 ls = __ldg(reinterpret_cast<const float2*>(log_sigmas) + row + g);
        const float rho = __ldg(corrs + row + g);
        const float sx = expf(ls.x), sy = expf(ls.y);
        const float e1 = e[2 * j], e2 = e[2 * j + 1];
        out[2 * j] = fmaf(sx, e1, mu.x);
        out[2 * j + 1] = fmaf(sy, fmaf(rho, e1, sqrtf(fmaxf(1.f - rho * rho, 0.f)) * e2), mu.y);
    }
    float* dst = pred + (vn * T + t0) * 2;
    if (VEC) {
        __stcs(reinterpret_cast<float4*>(dst), make_float4(out[0], out[1], out[2], out[3]));
    } else {
        dst[0] = out[0];
        dst[1] = out[1];
        if (t0 + 1 < T) { dst[2] = out[2]; dst[3] = out[3]; }
    }
    }   // pairs of this work item
    }   // grid-stride loop
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" int ccmpc_gmm2d_sample(const double* latent_pmf, const float* log_pis, const float* mus,
                                  const float* log_sigmas, const float* corrs, int64_t V, int64_t N, int Z, int T,
                                  int G, uint64_t seed, const uint64_t* seed_dev, uint64_t sample_offset,
                                  const float* u_z, const float* u_c, const float* eps, int32_t* z_out,
                                  float* predictions, void* stream) {
    CCMPC_REQUIRE(V >= 0 && N >= 0 && Z >= 1 && T >= 1 && G >= 1, "ccmpc_gmm2d_sample: bad dims");
    CCMPC_REQUIRE(Z <= 4096 && V <= 65535, "ccmpc_gmm2d_sample: need Z <= 4096 and V <= 65535");
    if (V == 0 || N == 0) return CCMPC_OK;
    CCMPC_REQUIRE(latent_pmf && mus && log_sigmas && corrs && predictions, "ccmpc_gmm2d_sample: null pointer");
    CCMPC_REQUIRE(G == 1 || log_pis, "ccmpc_gmm2d_sample: G > 1 needs log_pis");
    const bool any = u_z || u_c || eps, all = u_z && u_c && eps;
    CCMPC_REQUIRE(!any || all, "ccmpc_gmm2d_sample: supply u_z, u_c and eps together or none of them");
    const int TP = (T + 1) / 2;
    const int pairs = (T % 4 == 0) ? 2 : 1;            // 4 timesteps (32 contiguous bytes) per thread when possible
    const int TQ = (TP + pairs - 1) / pairs;
    int64_t blocks = (N * TQ + GMM_THREADS - 1) / GMM_THREADS;
    const int64_t cap = (148 * 8 + V - 1) / V;        // ~8 resident CTAs per SM over all vehicles; grid-stride beyond
    if (blocks > cap) blocks = cap;
    dim3 grid((unsigned)blocks, (unsigned)V);
    GmmNoise noise{u_z, u_c, eps};
    const size_t smem = (size_t)((Z + 1) & ~1) * sizeof(float) + (size_t)Z * sizeof(double);
    cudaStream_t st = (cudaStream_t)stream;
#define CCMPC_GMM_LAUNCH(VEC, PAIRS)                                                                          \
    gmm2d_sample_kernel<VEC, PAIRS><<<grid, GMM_THREADS, smem, st>>>(latent_pmf, log_pis, mus, log_sigmas, corrs, N, \
                                                                    Z, T, G, seed, seed_dev, sample_offset, noise,  \
                                                                    z_out, predictions)
    if (pairs == 2) CCMPC_GMM_LAUNCH(true, 2);
    else if ((T & 1) == 0) CCMPC_GMM_LAUNCH(true, 1);
    else CCMPC_GMM_LAUNCH(false, 1);
#undef CCMPC_GMM_LAUNCH
    return check_launch("gmm2d_sample_kernel");
}
