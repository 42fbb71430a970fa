// group.cu - stage-1 input preparation (A2) of libccmpc.
//
// Reference: collect/in_simulation/midlevel/ovehicle.py:72-76 (per-sample yaw from consecutive
// predicted positions): yaw[n, 0] = atan2 of (p[n, 0] - last observed position),
// yaw[n, t] = atan2 of (p[n, t] - p[n, t-1]).
#include "common.cuh"

namespace ccmpc {

// largest c with offsets[c] <= row  (cells are contiguous row ranges; empty cells are skipped)
__device__ __forceinline__ int64_t cell_of_row(const int64_t* __restrict__ offsets, int64_t C, int64_t row) {
    int64_t lo = 0, hi = C;
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (offsets[mid] <= row) lo = mid; else hi = mid;
    }
    return lo;
}

__global__ void yaws_kernel(const double* __restrict__ samples, const int64_t* __restrict__ offsets, int64_t C,
                            int64_t N, int T, const double* __restrict__ pos_last, double* __restrict__ yaws) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * T) return;
    const int64_t row = i / T;
    const int t = (int)(i - row * T);
    const double2* p2 = reinterpret_cast<const double2*>(samples);
    const double2 cur = p2[i];
    double2 prev;
    if (t == 0) {
        const int64_t c = cell_of_row(offsets, C, row);
        prev = make_double2(pos_last[2 * c], pos_last[2 * c + 1]);
    } else {
        prev = p2[i - 1];
    }
    yaws[i] = atan2(cur.y - prev.y, cur.x - prev.x);
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" int ccmpc_yaws_from_positions(const double* samples, const int64_t* offsets, int64_t n_total, int64_t C,
                                         int T, const double* pos_last, double* yaws, void* stream) {
    CCMPC_REQUIRE(T >= 1 && T <= CCMPC_MAX_T && C >= 1 && n_total >= 0, "ccmpc_yaws_from_positions: bad dims");
    CCMPC_REQUIRE(offsets && pos_last && yaws && (samples || n_total == 0), "ccmpc_yaws_from_positions: null pointer");
    if (n_total == 0) return CCMPC_OK;
    const int64_t tot = n_total * T;
    yaws_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>(samples, offsets, C, n_total, T,
                                                                                pos_last, yaws);
    return check_launch("yaws_kernel");
}
