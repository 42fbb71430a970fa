// fp64_mix_bench.cu - what does a vector FP64 instruction cost INSIDE the in-order stream of the warp that issues the
// DMMAs (the consumer of csrc/fused.cu: 10 DMMAs + 4 pivot shifts + 4 S1 adds per group of 4 samples)?
//   one CTA per SM, warp w on sub-partition w % 4; warps 0-3 run the "consumer" stream MODE, warps 4.. run dependent
//   DFMA chains (nf per sub-partition: the rollout producers).  Every warp times itself with clock64().
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_mix_bench fp64_mix_bench.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma16(double (&c)[4], const double (&a)[2], double b) {      // m16n8k4 (sm_90+)
    asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(b));
}
__device__ __forceinline__ double dadd_(double a, double b) {
    double r;
    asm volatile("add.rn.f64 %0, %1, %2;" : "=d"(r) : "d"(a), "d"(b));
    return r;
}
__device__ __forceinline__ int iadd_(int a, int b) {
    int r;
    asm volatile("add.s32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

// MODE 0: 20 DMMA | 1: + 16 independent DADD, one after each DMMA | 2: + 16 independent DADD in one batch
//      3: + 16 integer adds interleaved | 4: the consumer's pattern: 8 shifts feed the DMMAs, 8 S1 adds (interleaved)
//      5: as 4, shifts + adds batched before the 20 DMMAs | 6: 10 x m16n8k4 (same flops as 20 x m8n8k4)
//      7: + 4 LDS.128 interleaved
template <int MODE>
__global__ void __launch_bounds__(1024) mix_kernel(long long* cyc, double* sink, int nf, int iters, double a, double b,
                                                   int kind) {
    __shared__ double2 sm[1024];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double acc[20], s[8], f[8];
    double c16[5][4];
    int ia[8];
    sm[threadIdx.x] = make_double2(a, b);
#pragma unroll
    for (int i = 0; i < 20; ++i) acc[i] = threadIdx.x + i;
#pragma unroll
    for (int i = 0; i < 8; ++i) { s[i] = i; f[i] = a + i; ia[i] = i; }
#pragma unroll
    for (int i = 0; i < 5; ++i) c16[i][0] = c16[i][1] = c16[i][2] = c16[i][3] = i;
    __syncthreads();
    const long long t0 = clock64();
    if (warp < 4) {
        for (int it = 0; it < iters && MODE != 8; ++it) {
            if (MODE == 6) {
                double aa[2] = {a, b};
#pragma unroll
                for (int i = 0; i < 10; ++i) dmma16(c16[i % 5], aa, b);
                continue;
            }
            if (MODE == 2) {
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] = dadd_(s[i], a); f[i] = dadd_(f[i], b); }
            }
            if (MODE == 5) {
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = dadd_(sm[(lane + i + it) & 1023].x, -b);
#pragma unroll
                for (int i = 0; i < 8; ++i) s[i] = dadd_(s[i], f[i]);
            }
#pragma unroll
            for (int i = 0; i < 20; ++i) {
                if (MODE == 4 && i < 16 && (i & 1) == 0) f[i / 2] = dadd_(sm[(lane + i + it) & 1023].x, -b);
                if (MODE == 4 || MODE == 5) dmma(acc[2 * (i % 10)], acc[2 * (i % 10) + 1], f[i % 8], f[(i + 3) % 8]);
                else dmma(acc[2 * (i % 10)], acc[2 * (i % 10) + 1], a, b);
                if (MODE == 1 && i < 16) { if (i & 1) s[i / 2] = dadd_(s[i / 2], a); else f[i / 2] = dadd_(f[i / 2], b); }
                if (MODE == 3 && i < 16) ia[i / 2] = iadd_(ia[i / 2], it);
                if (MODE == 4 && i < 16 && (i & 1)) s[i / 2] = dadd_(s[i / 2], f[i / 2]);
                if (MODE == 7 && i < 4) { double2 v = sm[(lane + i + it) & 1023]; s[i] = v.x; s[i + 4] = v.y; }
            }
        }
    } else {
        // the other warps: kind 0 one dependent DFMA chain | 1 integer multiply-adds, 5 independent chains (Philox-like)
        //                  2 FFMA, 5 independent chains | 3 MUFU.EX2, 5 independent chains
        double x = threadIdx.x;
        unsigned u[5];
        float g[5];
#pragma unroll
        for (int i = 0; i < 5; ++i) { u[i] = threadIdx.x + i; g[i] = 0.001f * (threadIdx.x + i); }
        const unsigned ua = (unsigned)iters | 1u;
        const float fa = (float)a, fb = (float)b;
        if (kind == 0) {
            for (int it = 0; it < iters; ++it) {
#pragma unroll
                for (int i = 0; i < 10; ++i) x = fma(x, a, b);
            }
        } else if (kind == 1) {
            for (int it = 0; it < iters; ++it) {
#pragma unroll
                for (int i = 0; i < 10; ++i) u[i % 5] = u[i % 5] * ua + 12345u;
            }
        } else if (kind == 2) {
            for (int it = 0; it < iters; ++it) {
#pragma unroll
                for (int i = 0; i < 10; ++i) g[i % 5] = fmaf(g[i % 5], fa, fb);
            }
        } else {
            for (int it = 0; it < iters; ++it) {
#pragma unroll
                for (int i = 0; i < 10; ++i) g[i % 5] = exp2f(g[i % 5]);
            }
        }
#pragma unroll
        for (int i = 0; i < 5; ++i) x += u[i] + g[i];
        acc[0] = x;
    }
    const long long t1 = clock64();
    double r = 0;
#pragma unroll
    for (int i = 0; i < 20; ++i) r += acc[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) r += s[i] + f[i] + ia[i];
#pragma unroll
    for (int i = 0; i < 5; ++i) r += c16[i][0] + c16[i][1] + c16[i][2] + c16[i][3];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (lane == 0) cyc[blockIdx.x * 32 + warp] = t1 - t0;
}

template <int MODE>
static void run(const char* what, int nf, long long* cyc, double* sink, int kind = 0) {
    const int blocks = 148, iters = 2000, threads = 128 * (1 + nf);
    static long long h[148 * 32];
    for (int rep = 0; rep < 2; ++rep) mix_kernel<MODE><<<blocks, threads>>>(cyc, sink, nf, iters, 1.0000001, 1e-9, kind);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return; }
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double sd = 0, sf = 0;
    for (int b = 0; b < blocks; ++b)
        for (int w = 0; w < 4 * (1 + nf); ++w) (w < 4 ? sd : sf) += (double)h[b * 32 + w];
    printf("mode %d F=%d kind %d  %-58s consumer %8.1f clk/iteration   other warps %6.2f clk/instr\n", MODE, nf, kind, what,
           sd / (blocks * 4) / iters, nf ? sf / (blocks * 4 * nf) / (10.0 * iters) : 0.0);
}

int main() {
    long long* cyc;
    double* sink;
    cudaMalloc(&cyc, sizeof(long long) * 148 * 32);
    cudaMalloc(&sink, sizeof(double) * 148 * 1024);
    printf("# one iteration = the FP64 work of TWO groups of 4 samples in fused.cu's consumer (20 DMMA m8n8k4)\n");
    for (int nf : {0, 5}) {
        run<0>("20 DMMA", nf, cyc, sink);
        run<1>("20 DMMA + 16 independent DADD interleaved", nf, cyc, sink);
        run<2>("20 DMMA + 16 independent DADD in one batch", nf, cyc, sink);
        run<3>("20 DMMA + 16 integer adds interleaved", nf, cyc, sink);
        run<4>("consumer pattern: LDS, 8 shifts feed DMMAs, 8 S1 adds", nf, cyc, sink);
        run<5>("consumer pattern, shifts + adds batched before the DMMAs", nf, cyc, sink);
        run<6>("10 DMMA m16n8k4", nf, cyc, sink);
        run<7>("20 DMMA + 4 LDS.128", nf, cyc, sink);
    }
    printf("# other warps (5 per sub-partition) of another kind, with an idle consumer (mode 8) and beside 20 DMMA (mode 0)\n");
    for (int kind : {0, 1, 2, 3}) {
        run<8>("idle consumer", 5, cyc, sink, kind);
        run<0>("20 DMMA", 5, cyc, sink, kind);
        run<7>("20 DMMA + 4 LDS.128", 5, cyc, sink, kind);
    }
    return 0;
}
