// block.cu - solver-ready inequality blocks of libccmpc (the data-parallel half).
//
// The reference turns every emitted half-plane into ONE Python-level cvxpy row,
//   n_star.T @ cp.vstack([temp_x[t][0], temp_x[t][1]]) >= d_star     (v8ideal:927-930, :1503-1508)
//   ...                                                <= d_star     (v8ideal:934-937, :1509-1515)
// i.e. 2 800 expression objects per MPC step at config c2.  temp_x[t] are columns (nx t, nx t + 1) of the
// stacked state x = Gamma (u - u_bar) + x_bar, X = x.reshape(T, nx) (v8ideal:2873-2894, 2931).  This kernel
// writes the same rows as ONE dense block in that layout, with the side already folded in:
//     A_x [rows, nx T] x <= b_x [rows]      A_x[r, nx t + (0,1)] = -side n,   b_x[r] = -side rhs
// and, when the planner hands over its LTV model (Gamma [nx T, ldg], x_bar [nx T], u_bar [nu T];
// bicycle_v2.py:261-308 builds them on the host), the block in the control variable it actually solves for:
//     A_u [rows, nu T] u <= b_u [rows]      A_u = A_x Gamma[:, :nu T],   b_u = b_x - A_x (x_bar - Gamma u_bar).
// A_x has two non-zeros per row, so A_u is two scaled rows of Gamma: no GEMM, one coalesced write pass.
#include "common.cuh"

namespace ccmpc {

// One thread per output element of the wider of the two blocks; grid-stride over rows.
__global__ void halfplanes_block_kernel(const double* __restrict__ nd, int nd_stride, const double* __restrict__ rhs,
                                        int rhs_stride, const int8_t* __restrict__ idx_side,
                                        const int32_t* __restrict__ t_of_row, int64_t rows, int T, int nx, int nu,
                                        const double* __restrict__ gamma, int64_t ldg,
                                        const double* __restrict__ x_bar, const double* __restrict__ u_bar,
                                        double* __restrict__ A_x, double* __restrict__ b_x,
                                        double* __restrict__ A_u, double* __restrict__ b_u) {
    const int wx = nx * T, wu = nu * T;
    const int width = max(wx, gamma ? wu : 0);
    for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
        const int t = t_of_row[r];
        // side: bit1 of idx_side set = the row is n.x <= rhs (A = n, b = rhs); clear = n.x >= rhs (A = -n, b = -rhs)
        const double s = (idx_side[r] & 2) ? 1.0 : -1.0;
        const double a0 = s * nd[r * nd_stride], a1 = s * nd[r * nd_stride + 1];
        const double b = s * rhs[r * rhs_stride];
        const int c0 = nx * t;
        for (int c = threadIdx.x; c < width; c += blockDim.x) {
            if (A_x && c < wx) A_x[r * wx + c] = (c == c0) ? a0 : ((c == c0 + 1) ? a1 : 0.0);
            if (A_u && c < wu)
                A_u[r * wu + c] = a0 * gamma[(int64_t)c0 * ldg + c] + a1 * gamma[(int64_t)(c0 + 1) * ldg + c];
        }
        if (threadIdx.x == 0 && b_x) b_x[r] = b;
        if (b_u && threadIdx.x < 32) {
            // b_u = b - a . (x_bar - Gamma u_bar) on rows c0, c0 + 1: warp reduction over the nu T columns
            double acc0 = 0.0, acc1 = 0.0;
            for (int c = threadIdx.x; c < wu; c += 32) {
                acc0 = fma(gamma[(int64_t)c0 * ldg + c], u_bar[c], acc0);
                acc1 = fma(gamma[(int64_t)(c0 + 1) * ldg + c], u_bar[c], acc1);
            }
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                acc0 += __shfl_xor_sync(0xffffffffu, acc0, o);
                acc1 += __shfl_xor_sync(0xffffffffu, acc1, o);
            }
            if (threadIdx.x == 0) b_u[r] = b - (a0 * (x_bar[c0] - acc0) + a1 * (x_bar[c0 + 1] - acc1));
        }
    }
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" int ccmpc_halfplanes_to_block(const double* nd, int nd_stride, const double* rhs, int rhs_stride,
                                         const int8_t* idx_side, const int32_t* t_of_row, int64_t rows, int T, int nx,
                                         int nu, const double* gamma, int64_t ldg, const double* x_bar,
                                         const double* u_bar, double* A_x, double* b_x, double* A_u, double* b_u,
                                         void* stream) {
    CCMPC_REQUIRE(rows >= 0 && T >= 1 && T <= CCMPC_MAX_T && nx >= 2 && nx <= 16, "ccmpc_halfplanes_to_block: bad sizes");
    CCMPC_REQUIRE(nd && rhs && idx_side && t_of_row && nd_stride >= 2 && rhs_stride >= 1,
                  "ccmpc_halfplanes_to_block: null input");
    CCMPC_REQUIRE((A_x != nullptr) == (b_x != nullptr) && (A_u != nullptr) == (b_u != nullptr) && (A_x || A_u),
                  "ccmpc_halfplanes_to_block: outputs come in (A, b) pairs");
    if (A_u) {
        CCMPC_REQUIRE(gamma && x_bar && u_bar && nu >= 1 && nu <= 8 && ldg >= (int64_t)nu * T,
                      "ccmpc_halfplanes_to_block: the control-space block needs gamma [nx T, ldg >= nu T], x_bar, u_bar");
    }
    if (rows == 0) return CCMPC_OK;
    const unsigned grid = (unsigned)(rows < (int64_t)sm_count() * 16 ? rows : (int64_t)sm_count() * 16);
    halfplanes_block_kernel<<<grid, 64, 0, (cudaStream_t)stream>>>(nd, nd_stride, rhs, rhs_stride, idx_side, t_of_row,
                                                                   rows, T, nx, nu, A_u ? gamma : nullptr, ldg, x_bar,
                                                                   u_bar, A_x, b_x, A_u, b_u);
    return check_launch("halfplanes_block_kernel");
}
