// cost.cu - closed-loop episode cost of libccmpc.
//
// Replaces getData_singleEpisode.cost_final (dataForCost_ref.py:198-232) for a
// batch of E episodes: one thread per episode (the offline aggregator
// montecarlo_cost_ref.py:55-77 evaluates <= 100 episodes of <= 50 frames).
//   costX = 6 |x_F - g_x|^2 + 6 |y_F - g_y|^2                          (:216-219)
//   costU = U[H-2]^T R1 U[H-2]    (the loop at :223-224 OVERWRITES, so only tau = H-2 survives)
//         + sum_{tau < H-2} (U[tau+1]-U[tau])^T R2 (U[tau+1]-U[tau])   (:226-227)
// with R1 = [[.5,.2],[.2,1]], R2 = [[.5,.1],[.1,2]] (:214-215).  H < 2: NaN
// (the reference raises UnboundLocalError; the Python shim raises the same).
#include "common.cuh"

namespace ccmpc {

// u^T R u evaluated as NumPy does: (u^T R) first, then the dot with u
__device__ __forceinline__ double quad(double u0, double u1, double r00, double r01, double r10, double r11) {
    const double v0 = u0 * r00 + u1 * r10;
    const double v1 = u0 * r01 + u1 * r11;
    return v0 * u0 + v1 * u1;
}

__global__ void episode_cost_kernel(const double* __restrict__ x_last, const double* __restrict__ U,
                                    const int32_t* __restrict__ lengths, int64_t E, int Hmax, double gx, double gy,
                                    double* __restrict__ cost, double* __restrict__ costU_out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    const int H = lengths[e];
    const double* u = U + e * (int64_t)Hmax * 2;
    const double ax = fabs(x_last[2 * e] - gx), ay = fabs(x_last[2 * e + 1] - gy);
    const double costX = 6.0 * (ax * ax) + 6.0 * (ay * ay);
    double costU = __longlong_as_double(0x7ff8000000000000LL);
    if (H >= 2 && H <= Hmax) {
        costU = quad(u[2 * (H - 2)], u[2 * (H - 2) + 1], 0.5, 0.2, 0.2, 1.0);
        for (int tau = 0; tau < H - 2; ++tau) {
            const double d0 = u[2 * (tau + 1)] - u[2 * tau], d1 = u[2 * (tau + 1) + 1] - u[2 * tau + 1];
            costU = costU + quad(d0, d1, 0.5, 0.1, 0.1, 2.0);
        }
    }
    cost[e] = costX + costU;
    costU_out[e] = costU;
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" int ccmpc_episode_cost(const double* x_last, const double* U, const int32_t* lengths, int64_t E, int Hmax,
                                  double goal_x, double goal_y, double* cost, double* costU, void* stream) {
    CCMPC_REQUIRE(E >= 0 && Hmax >= 0 && x_last && U && lengths && cost && costU, "ccmpc_episode_cost: bad arguments");
    if (E == 0) return CCMPC_OK;
    episode_cost_kernel<<<(unsigned)((E + 127) / 128), 128, 0, (cudaStream_t)stream>>>(x_last, U, lengths, E, Hmax,
                                                                                      goal_x, goal_y, cost, costU);
    return check_launch("episode_cost_kernel");
}
