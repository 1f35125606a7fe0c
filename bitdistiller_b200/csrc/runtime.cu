// Host-side runtime shared by all entry points: error reporting, launch accounting, device attributes.
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "bdq_common.cuh"

namespace bdq {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

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
        set_error("%s: CUDA error: %s", what, cudaGetErrorString(e));
        return BDQ_ERR_CUDA;
    }
    return BDQ_OK;
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

}  // namespace bdq

extern "C" int bdq_version(void) { return 100; }
extern "C" const char* bdq_last_error(void) { return bdq::g_err; }
extern "C" int64_t bdq_launch_count(void) { return bdq::g_launches.load(std::memory_order_relaxed); }
