// fp64_pipe_bench.cu - what bounds the 2T x 2T second-moment update of sample_sums_kernel on B200?
// Measures (1) vector DFMA issue rate, (2) mma.sync m8n8k4 f64 (DMMA) rate, (3) both interleaved.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_pipe_bench fp64_pipe_bench.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) dfma_kernel(double* out, int iters, double a, double b) {
    double acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(256) dmma_kernel(double* out, int iters, double a, double b) {
    double c[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) c[i] = threadIdx.x + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) dmma(c[2 * i], c[2 * i + 1], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) mixed_kernel(double* out, int iters, double a, double b) {
    double c[16], acc[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) c[i] = threadIdx.x + i;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = threadIdx.x - i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            dmma(c[2 * i], c[2 * i + 1], a, b);
            acc[i] = fma(acc[i], a, b);
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += c[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float time_ms(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    launch();
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int i = 0; i < 5; ++i) launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms / 5;
}

int main() {
    const int blocks = 148 * 8, threads = 256, iters = 4096;
    double* out;
    cudaMalloc(&out, sizeof(double) * blocks * threads);
    float ms = time_ms([&] { dfma_kernel<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); });
    double fl = 2.0 * 16 * iters * (double)blocks * threads;
    printf("DFMA  : %.3f ms  %.2f TFLOP/s\n", ms, fl / ms * 1e-9);
    ms = time_ms([&] { dmma_kernel<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); });
    fl = 2.0 * 8 * 256 * iters * (double)blocks * threads / 32;
    printf("DMMA  : %.3f ms  %.2f TFLOP/s\n", ms, fl / ms * 1e-9);
    ms = time_ms([&] { mixed_kernel<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); });
    fl = (2.0 * 8 * 256 / 32 + 2.0 * 8) * iters * (double)blocks * threads;
    printf("MIXED : %.3f ms  %.2f TFLOP/s (dmma share %.2f)\n", ms, fl / ms * 1e-9, 64.0 / 72.0);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
