// l4.cu - K5: box corners and the 4-face oriented outer approximation.
//
// Reference call sites (paths relative to the reference root):
//   python-utility/utility/utility/npu.py:308-350       vertices_of_bboxes
//   collect/in_simulation/midlevel/util.py:171-200      compute_L4_outerapproximation
//   collect/in_simulation/midlevel/v8ideal/__init__.py:627-640, :694-736   per (o, k, t) drivers
// The reference materialises (N_k, 4, 2) corners per cell and timestep and
// then takes max(At @ v) over samples and corners; l4_outer_kernel reads
// (x, y, yaw) once, evaluates the corner maximum in closed form and keeps only
// the four running maxima.
#include "common.cuh"

namespace ccmpc {

constexpr int L4_THREADS = 256;

// corners of one box, npu.py:343-347 written out (l-term + w-term)
__device__ __forceinline__ void box_corners(double cx, double cy, double th, double l, double w, double vx[4],
                                            double vy[4]) {
    double S, C;
    sincos(th, &S, &C);
    const double Cl = C * l, Sl = S * l, Cw = C * w, Sw = S * w;
    vx[0] = cx + 0.5 * (-Cl + Sw); vy[0] = cy + 0.5 * (-Sl - Cw);
    vx[1] = cx + 0.5 * (-Cl - Sw); vy[1] = cy + 0.5 * (-Sl + Cw);
    vx[2] = cx + 0.5 * (Cl - Sw);  vy[2] = cy + 0.5 * (Sl + Cw);
    vx[3] = cx + 0.5 * (Cl + Sw);  vy[3] = cy + 0.5 * (Sl - Cw);
}

__global__ void vertices_kernel(const double* __restrict__ centers, const double* __restrict__ theta,
                                int64_t theta_is_scalar, double l, double w, int64_t N,
                                double* __restrict__ vertices) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    double vx[4], vy[4];
    box_corners(centers[2 * i], centers[2 * i + 1], theta_is_scalar ? theta[0] : theta[i], l, w, vx, vy);
    double2* o = reinterpret_cast<double2*>(vertices) + 4 * i;
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = make_double2(vx[k], vy[k]);
}

// grid (C, Y): block (c, y) strides over the rows of cell c; thread = (row lane, t).
__global__ void __launch_bounds__(L4_THREADS)
l4_outer_kernel(const double* __restrict__ samples, const double* __restrict__ yaws,
                const int64_t* __restrict__ offsets, int T, const double* __restrict__ yaw_mean,
                const double* __restrict__ lw_cell, double l, double w,
                double* __restrict__ At, double* __restrict__ b0) {
    __shared__ double s_max[L4_THREADS * 4];
    const int64_t c = blockIdx.x;
    const int tid = threadIdx.x;
    const int rpp = L4_THREADS / T;
    const int rl = tid / T, t = tid - rl * T;
    const int64_t b = offsets[c], e = offsets[c + 1];
    const double ninf = __longlong_as_double(0xfff0000000000000LL);
    if (blockIdx.y > 0 && b + (int64_t)blockIdx.y * rpp >= e) return;      // no row for this block (CTA-uniform)
    double m0 = ninf, m1 = ninf, m2 = ninf, m3 = ninf;
    double ct = 1.0, stt = 0.0;
    if (lw_cell) { l = lw_cell[2 * c]; w = lw_cell[2 * c + 1]; }
    if (rl < rpp) {
        const double ym = yaw_mean[c * T + t];
        sincos(ym, &stt, &ct);
        const double2* g2 = reinterpret_cast<const double2*>(samples);
        const double hl = 0.5 * l, hw = 0.5 * w;
        const int64_t stride = (int64_t)gridDim.y * rpp;
        const int64_t row0 = b + (int64_t)blockIdx.y * rpp + rl;
        // 4 rows in flight per thread: the loads of a batch are issued before the first sincos
        for (int64_t rb = row0; rb < e; rb += 4 * stride) {
            double2 pp[4];
            double tt[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int64_t row = rb + k * stride;
                const bool ok = row < e;
                pp[k] = ok ? __ldcs(g2 + row * T + t) : make_double2(0.0, 0.0);
                tt[k] = ok ? __ldcs(yaws + row * T + t) : 0.0;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
            if (rb + k * stride >= e) break;
            const double2 p = pp[k];
            const double th = tt[k];
            // The corners are centre +- (l/2) e1 +- (w/2) e2 with e1 = (cos th, sin th), e2 = (-sin th, cos th)
            // (npu.py:343-347), and a row of At is a unit vector d, so max over the 4 corners of d.v is
            //   d.centre + (l/2) |d.e1| + (w/2) |d.e2|,
            // with d.e1 = cos(th - mean yaw), d.e2 = -sin(th - mean yaw) for the first row and the roles of
            // sin / cos swapped for the second (util.py:190-199): one sincos of the yaw DIFFERENCE per sample
            // instead of the four corners.
            double sd, cd;
            sincos(th - ym, &sd, &cd);
            const double ca = fabs(cd), sa = fabs(sd);
            const double eu = fma(hl, ca, hw * sa), ev = fma(hl, sa, hw * ca);
            const double u = fma(ct, p.x, stt * p.y);
            const double v = fma(-stt, p.x, ct * p.y);
            m0 = fmax(m0, u + eu);
            m1 = fmax(m1, v + ev);
            m2 = fmax(m2, eu - u);
            m3 = fmax(m3, ev - v);
            }
        }
    }
    s_max[tid * 4 + 0] = m0; s_max[tid * 4 + 1] = m1; s_max[tid * 4 + 2] = m2; s_max[tid * 4 + 3] = m3;
    __syncthreads();
    if (tid < T) {
        double r[4] = {ninf, ninf, ninf, ninf};
        for (int k = 0; k < rpp; ++k)
#pragma unroll
            for (int q = 0; q < 4; ++q) r[q] = fmax(r[q], s_max[(k * T + tid) * 4 + q]);
        double* bo = b0 + (c * T + tid) * 4;
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (r[q] > ninf) atomic_max_double(bo + q, r[q]);
        if (blockIdx.y == 0) {
            double* ao = At + (c * T + tid) * 8;
            ao[0] = ct;  ao[1] = stt;
            ao[2] = -stt; ao[3] = ct;
            ao[4] = -ct; ao[5] = -stt;
            ao[6] = stt; ao[7] = -ct;
        }
    }
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" int ccmpc_vertices_of_bboxes(const double* centers, const double* theta, int64_t theta_is_scalar,
                                        double l, double w, int64_t N, double* vertices, void* stream) {
    CCMPC_REQUIRE(N >= 0 && centers && theta && vertices, "ccmpc_vertices_of_bboxes: bad arguments");
    if (N == 0) return CCMPC_OK;
    vertices_kernel<<<(unsigned)((N + 127) / 128), 128, 0, (cudaStream_t)stream>>>(centers, theta, theta_is_scalar, l,
                                                                                  w, N, vertices);
    return check_launch("vertices_kernel");
}

extern "C" int ccmpc_l4_outer(const double* samples, const double* yaws, const int64_t* offsets, int64_t n_total,
                              int64_t C, int T, const double* yaw_mean, const double* lw_cell, double l, double w,
                              double* At, double* b0, void* stream) {
    CCMPC_REQUIRE(T >= 1 && T <= CCMPC_MAX_T && C >= 0 && n_total >= 0, "ccmpc_l4_outer: bad dims");
    CCMPC_REQUIRE(samples && yaws && offsets && yaw_mean && At && b0, "ccmpc_l4_outer: null pointer");
    CCMPC_REQUIRE(CCMPC_ALIGNED16(samples), "ccmpc_l4_outer: samples must be 16-byte aligned");
    if (C == 0) return CCMPC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    // 0xFF bytes = the identity of atomic_max_double (see common.cuh)
    cudaError_t e = cudaMemsetAsync(b0, 0xFF, (size_t)C * T * 4 * sizeof(double), st);
    if (e != cudaSuccess) { set_error("ccmpc_l4_outer: memset: %s", cudaGetErrorString(e)); return CCMPC_E_LAUNCH; }
    // Cell sizes live on the device (offsets), and ingest layouts leave most cells empty (cell = v * S + j), so
    // grid.y is sized for the case that ONE cell holds every row, 8 rows per thread: blocks whose first row lies
    // beyond their cell's end exit after reading two offsets, which is cheaper than the serial tail a small
    // grid.y leaves in the few populated cells.
    const int rpp = L4_THREADS / T;
    int64_t y = (n_total + (int64_t)rpp * 8 - 1) / ((int64_t)rpp * 8);
    if (y < 1) y = 1;
    if (y > 32) y = 32;
    dim3 grid((unsigned)C, (unsigned)y);
    l4_outer_kernel<<<grid, L4_THREADS, 0, st>>>(samples, yaws, offsets, T, yaw_mean, lw_cell, l, w, At, b0);
    return check_launch("l4_outer_kernel");
}
