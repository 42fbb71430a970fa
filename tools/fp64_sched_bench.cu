// fp64_sched_bench.cu - how do a DMMA stream and dependent DFMA chains share the FP64 pipe of one SM sub-partition?
// (design input for the warp-specialised rollout + sums kernel, csrc/fused.cu)
//   one CTA per SM; warp w sits on sub-partition w % 4.  Roles per sub-partition: D warps issue mma.sync.m8n8k4.f64
//   over 10 independent accumulators (the Gram update), F warps run ONE dependent DFMA chain each (the rollout chain).
//   Every warp times itself with clock64(); the host prints cycles per instruction per role.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_sched_bench fp64_sched_bench.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// warps [0, 4 nd): DMMA role; warps [4 nd, 4 (nd + nf)): chain role (depth = ilp independent chains)
__global__ void __launch_bounds__(1024) sched_kernel(long long* cyc, double* sink, int nd, int nf, int ilp, int iters,
                                                     double a, double b) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double acc[20];
#pragma unroll
    for (int i = 0; i < 20; ++i) acc[i] = threadIdx.x + i;
    __syncthreads();
    const long long t0 = clock64();
    if (warp < 4 * nd) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 10; ++i) dmma(acc[2 * i], acc[2 * i + 1], a, b);
        }
    } else if (ilp == 1) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 10; ++i) acc[0] = fma(acc[0], a, b);
        }
    } else if (ilp == 2) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 5; ++i) { acc[0] = fma(acc[0], a, b); acc[1] = fma(acc[1], a, b); }
        }
    } else {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 10; ++i) acc[i] = fma(acc[i], a, b);
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < 20; ++i) s += acc[i];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (lane == 0) cyc[blockIdx.x * 32 + warp] = t1 - t0;
}

int main() {
    const int blocks = 148, iters = 2000;
    long long* cyc;
    double* sink;
    cudaMalloc(&cyc, sizeof(long long) * blocks * 32);
    cudaMalloc(&sink, sizeof(double) * blocks * 1024);
    long long h[148 * 32];
    printf("# D = DMMA warps per sub-partition, F = chain warps per sub-partition, ilp = independent DFMA chains per F warp\n");
    printf("# cycles per warp-instruction (10 per iteration) seen by each role\n");
    const int cfg[][3] = {{1, 0, 1}, {2, 0, 1}, {3, 0, 1}, {4, 0, 1}, {6, 0, 1}, {8, 0, 1}, {0, 1, 1}, {0, 4, 1}, {0, 1, 2}, {0, 1, 10}, {0, 4, 10},
                          {1, 1, 1}, {1, 2, 1}, {1, 4, 1}, {2, 4, 1}, {1, 4, 2}, {2, 4, 2}, {1, 4, 10}, {2, 4, 10},
                          {1, 6, 1}, {2, 6, 2}};
    for (auto& c : cfg) {
        const int nd = c[0], nf = c[1], ilp = c[2];
        const int threads = 128 * (nd + nf);
        for (int rep = 0; rep < 2; ++rep)
            sched_kernel<<<blocks, threads>>>(cyc, sink, nd, nf, ilp, iters, 1.0000001, 1e-9);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double sd = 0, sf = 0;
        for (int b = 0; b < blocks; ++b)
            for (int w = 0; w < 4 * (nd + nf); ++w) (w < 4 * nd ? sd : sf) += (double)h[b * 32 + w];
        const double per = 10.0 * iters;
        printf("D=%d F=%d ilp=%2d :  DMMA %7.2f clk/instr   DFMA %7.2f clk/instr\n", nd, nf, ilp,
               nd ? sd / (blocks * 4 * nd) / per : 0.0, nf ? sf / (blocks * 4 * nf) / per : 0.0);
    }
    return 0;
}
