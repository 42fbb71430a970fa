// abi.cu - version, last-error string and launch checking of libccmpc.
#include <stdarg.h>
#include <atomic>
#include <string.h>

#include "common.cuh"

namespace ccmpc {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return CCMPC_E_LAUNCH;
    }
    return CCMPC_OK;
}

int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
    }
    return n;
}

}  // namespace ccmpc

extern "C" int ccmpc_version(void) { return CCMPC_VERSION; }
extern "C" const char* ccmpc_last_error_string(void) { return ccmpc::g_err; }
extern "C" uint64_t ccmpc_launch_count(void) { return ccmpc::g_launches.load(std::memory_order_relaxed); }
#ifndef CCMPC_SRC_HASH
#define CCMPC_SRC_HASH 0ULL
#endif
extern "C" uint64_t ccmpc_source_hash(void) { return CCMPC_SRC_HASH; }
