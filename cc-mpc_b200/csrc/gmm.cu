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

// the two GMM2D draws of timestep pair tp (timesteps 2 tp, 2 tp + 1) of sample (v, n) with latent z:
// out = (y0_{2tp}, y1_{2tp}, y0_{2tp+1}, y1_{2tp+1}); draws beyond T stay 0
__device__ __forceinline__ void gmm_pair(const float* __restrict__ log_pis, const float* __restrict__ mus,
                                         const float* __restrict__ log_sigmas, const float* __restrict__ corrs,
                                         const GmmNoise& noise, int v, int64_t vn, int z, int Z, int T, int G, int tp,
                                         uint32_t gi_lo, uint32_t gi_hi, uint32_t seed_lo, uint32_t seed_hi,
                                         float out[4]) {
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
    out[0] = out[1] = out[2] = out[3] = 0.f;
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
    float out[4];
    const int t0 = 2 * tp;
    gmm_pair(log_pis, mus, log_sigmas, corrs, noise, v, vn, z, Z, T, G, tp, gi_lo, gi_hi, seed_lo, seed_hi, out);
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

// ---------------------------------------------------------------------------------------------------------
// Control samples -> positions: `dynamic.integrate_samples` of p_y_xz in PREDICT mode (the second value
// prediction.py:85-86 keeps).  Published Trajectron++ arithmetic (model/dynamics/unicycle.py and
// single_integrator.py; submodule absent from the reference tree), float32 like the model:
//   single integrator : p_t = p_0 + dt * sum_{s <= t} u_s
//   unicycle, u = (dphi, a), state (x, y, phi, v), per step with w = dphi if |dphi| > 1e-2 else 1:
//       phi' = phi + w dt;  ds = (sin phi' - sin phi) / w;  dc = (cos phi' - cos phi) / w
//       |dphi| >  1e-2 : x += (a/w) dc + v ds + (a/w) sin(phi') dt;  y += -v dc + (a/w) ds - (a/w) cos(phi') dt;
//                        phi = phi'
//       |dphi| <= 1e-2 : x += v cos(phi) dt + (a/2) cos(phi) dt^2;   y += v sin(phi) dt + (a/2) sin(phi) dt^2
//       v += a dt
// One thread per sample (the recursion over t is sequential); the control draws use the SAME Philox counters
// as gmm2d_sample_kernel, so integrating the controls that kernel returns reproduces this kernel's output.
// The per-thread rows go through shared memory so the global stores are contiguous 16 B per lane.
constexpr int DYN_THREADS = 128;
enum { DYN_NONE = 0, DYN_SINGLE_INTEGRATOR = 1, DYN_UNICYCLE = 2 };

__device__ __forceinline__ void unicycle_step(float& x, float& y, float& phi, float& vel, float dphi, float a, float dt) {
    const bool small = fabsf(dphi) <= 1e-2f;
    const float w = small ? 1.f : dphi;
    float s0, c0;
    sincosf(phi, &s0, &c0);
    if (small) {
        const float h = (a * 0.5f) * c0 * (dt * dt), k = (a * 0.5f) * s0 * (dt * dt);
        x = x + vel * c0 * dt + h;
        y = y + vel * s0 * dt + k;
    } else {
        const float phin = phi + w * dt;
        float s1, c1;
        sincosf(phin, &s1, &c1);
        const float ds = (s1 - s0) / w, dc = (c1 - c0) / w, aw = a / w;
        x = x + aw * dc + vel * ds + aw * s1 * dt;
        y = y - vel * dc + aw * ds - aw * c1 * dt;
        phi = phi + dphi * dt;
    }
    vel = vel + a * dt;
}

template <int DYN>
__global__ void __launch_bounds__(DYN_THREADS)
gmm2d_dynamics_kernel(const double* __restrict__ latent_pmf, const float* __restrict__ log_pis,
                      const float* __restrict__ mus, const float* __restrict__ log_sigmas,
                      const float* __restrict__ corrs, const float* __restrict__ x0, float dt, int64_t N, int Z,
                      int T, int G, uint64_t seed, const uint64_t* __restrict__ seed_dev, uint64_t sample_offset,
                      GmmNoise noise, int32_t* __restrict__ z_out, float* __restrict__ pred) {
    extern __shared__ float s_cum[];
    const int v = blockIdx.y;
    const int zpad = (Z + 3) & ~3;                                        // keeps s_w and s_row 16-byte aligned
    double* s_w = reinterpret_cast<double*>(s_cum + zpad);
    float4* s_row = reinterpret_cast<float4*>(s_w + Z + (Z & 1));         // [DYN_THREADS][TP + 1 pad] pairs
    for (int k = threadIdx.x; k < Z; k += DYN_THREADS) s_w[k] = fmax(latent_pmf[(int64_t)v * Z + k], 0.0);
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
    if (seed_dev) seed += *seed_dev;
    const uint32_t seed_lo = (uint32_t)seed, seed_hi = (uint32_t)(seed >> 32);
    const int TP = (T + 1) >> 1;
    const int ldr = TP | 1;                                               // odd row pitch in float4: no bank conflicts
    const float4 st0 = __ldg(reinterpret_cast<const float4*>(x0) + v);   // (x, y, phi, v)
    for (int64_t base = (int64_t)blockIdx.x * DYN_THREADS; base < N; base += (int64_t)gridDim.x * DYN_THREADS) {
        const int64_t n = base + threadIdx.x;
        if (n < N) {
            const uint64_t gi = sample_offset + (uint64_t)n;
            const uint32_t gi_lo = (uint32_t)gi, gi_hi = (uint32_t)(gi >> 32);
            const int64_t vn = (int64_t)v * N + n;
            float uz;
            if (noise.u_z) {
                uz = noise.u_z[vn];
            } else {
                Philox r = philox4x32_10(gi_lo, gi_hi, (uint32_t)v, 0xFFFFFFFEu, seed_lo, seed_hi);
                uz = u01(r.c[0]);
            }
            const int z = pick_latent(s_cum, Z, uz);
            if (z_out) z_out[vn] = z;
            float x = st0.x, y = st0.y, phi = st0.z, vel = st0.w;
            for (int tp = 0; tp < TP; ++tp) {
                float u[4];
                gmm_pair(log_pis, mus, log_sigmas, corrs, noise, v, vn, z, Z, T, G, tp, gi_lo, gi_hi, seed_lo, seed_hi, u);
                float4 o;
                if (DYN == DYN_UNICYCLE) {
                    unicycle_step(x, y, phi, vel, u[0], u[1], dt);
                    o.x = x; o.y = y;
                    if (2 * tp + 1 < T) unicycle_step(x, y, phi, vel, u[2], u[3], dt);
                    o.z = x; o.w = y;
                } else {
                    x = x + u[0] * dt; y = y + u[1] * dt;
                    o.x = x; o.y = y;
                    x = x + u[2] * dt; y = y + u[3] * dt;
                    o.z = x; o.w = y;
                }
                s_row[threadIdx.x * ldr + tp] = o;
            }
        }
        __syncthreads();
        // rows of this block are one contiguous [rows, T, 2] float range of the output
        const int rows = (int)((N - base < DYN_THREADS) ? (N - base) : DYN_THREADS);
        float* dst = pred + ((int64_t)v * N + base) * T * 2;
        if ((T & 1) == 0) {
            for (int i = threadIdx.x; i < rows * TP; i += DYN_THREADS) {
                const int r = i / TP, c = i - r * TP;
                __stcs(reinterpret_cast<float4*>(dst) + i, s_row[r * ldr + c]);
            }
        } else {
            const float* s_f = reinterpret_cast<const float*>(s_row);
            for (int i = threadIdx.x; i < rows * T * 2; i += DYN_THREADS) {
                const int r = i / (2 * T), c = i - r * 2 * T;
                dst[i] = s_f[(r * ldr) * 4 + c];
            }
        }
        __syncthreads();
    }
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
    CCMPC_REQUIRE(CCMPC_ALIGNED16(predictions), "ccmpc_gmm2d_sample: predictions must be 16-byte aligned");
    CCMPC_REQUIRE(G == 1 || log_pis, "ccmpc_gmm2d_sample: G > 1 needs log_pis");
    const bool any = u_z || u_c || eps, all = u_z && u_c && eps;
    CCMPC_REQUIRE(!any || all, "ccmpc_gmm2d_sample: supply u_z, u_c and eps together or none of them");
    const int TP = (T + 1) / 2;
    const int pairs = (T % 4 == 0) ? 2 : 1;            // 4 timesteps (32 contiguous bytes) per thread when possible
    const int TQ = (TP + pairs - 1) / pairs;
    int64_t blocks = (N * TQ + GMM_THREADS - 1) / GMM_THREADS;
    const int64_t cap = ((int64_t)sm_count() * 8 + V - 1) / V;        // ~8 resident CTAs per SM over all vehicles; grid-stride beyond
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

extern "C" int ccmpc_gmm2d_sample_dynamics(const double* latent_pmf, const float* log_pis, const float* mus,
                                           const float* log_sigmas, const float* corrs, int dynamics, const float* x0,
                                           double dt, int64_t V, int64_t N, int Z, int T, int G, uint64_t seed,
                                           const uint64_t* seed_dev, uint64_t sample_offset, const float* u_z,
                                           const float* u_c, const float* eps, int32_t* z_out, float* predictions,
                                           void* stream) {
    if (dynamics == DYN_NONE)
        return ccmpc_gmm2d_sample(latent_pmf, log_pis, mus, log_sigmas, corrs, V, N, Z, T, G, seed, seed_dev,
                                  sample_offset, u_z, u_c, eps, z_out, predictions, stream);
    CCMPC_REQUIRE(dynamics == DYN_SINGLE_INTEGRATOR || dynamics == DYN_UNICYCLE,
                  "ccmpc_gmm2d_sample_dynamics: dynamics must be 0 (none), 1 (single integrator) or 2 (unicycle)");
    CCMPC_REQUIRE(V >= 0 && N >= 0 && Z >= 1 && T >= 1 && G >= 1, "ccmpc_gmm2d_sample_dynamics: bad dims");
    CCMPC_REQUIRE(Z <= 2048 && V <= 65535 && T <= 256, "ccmpc_gmm2d_sample_dynamics: need Z <= 2048, V <= 65535, T <= 256");
    if (V == 0 || N == 0) return CCMPC_OK;
    CCMPC_REQUIRE(latent_pmf && mus && log_sigmas && corrs && predictions && x0, "ccmpc_gmm2d_sample_dynamics: null pointer");
    CCMPC_REQUIRE(((uintptr_t)x0 & 15) == 0 && ((uintptr_t)predictions & 15) == 0,
                  "ccmpc_gmm2d_sample_dynamics: x0 and predictions must be 16-byte aligned");
    CCMPC_REQUIRE(G == 1 || log_pis, "ccmpc_gmm2d_sample_dynamics: G > 1 needs log_pis");
    const bool any = u_z || u_c || eps, all = u_z && u_c && eps;
    CCMPC_REQUIRE(!any || all, "ccmpc_gmm2d_sample_dynamics: supply u_z, u_c and eps together or none of them");
    const int TP = (T + 1) / 2;
    int64_t blocks = (N + DYN_THREADS - 1) / DYN_THREADS;
    const int64_t cap = ((int64_t)sm_count() * 8 + V - 1) / V;
    if (blocks > cap) blocks = cap;
    dim3 grid((unsigned)blocks, (unsigned)V);
    GmmNoise noise{u_z, u_c, eps};
    const size_t smem = (size_t)((Z + 3) & ~3) * sizeof(float) + (size_t)(Z + (Z & 1)) * sizeof(double) +
                        (size_t)DYN_THREADS * (TP | 1) * sizeof(float4);
    cudaStream_t st = (cudaStream_t)stream;
    if (dynamics == DYN_UNICYCLE) {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(gmm2d_dynamics_kernel<DYN_UNICYCLE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        gmm2d_dynamics_kernel<DYN_UNICYCLE><<<grid, DYN_THREADS, smem, st>>>(latent_pmf, log_pis, mus, log_sigmas, corrs, x0,
                                                                            (float)dt, N, Z, T, G, seed, seed_dev,
                                                                            sample_offset, noise, z_out, predictions);
    } else {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(gmm2d_dynamics_kernel<DYN_SINGLE_INTEGRATOR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        gmm2d_dynamics_kernel<DYN_SINGLE_INTEGRATOR><<<grid, DYN_THREADS, smem, st>>>(
            latent_pmf, log_pis, mus, log_sigmas, corrs, x0, (float)dt, N, Z, T, G, seed, seed_dev, sample_offset, noise,
            z_out, predictions);
    }
    return check_launch("gmm2d_dynamics_kernel");
}
