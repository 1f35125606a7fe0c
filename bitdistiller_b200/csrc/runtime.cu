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

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per DEVICE: remember (device, function) pairs, so that a process
// that drives several GPUs (HF device_map inference over WQLinear, autoclip on cuda:1 after cuda:0) sets it on each.
bool ensure_dyn_smem(const void* fn, int bytes) {
    struct Key { int dev; const void* fn; };
    static Key done[512];
    static int ndone = 0;
    static std::atomic_flag lock = ATOMIC_FLAG_INIT;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    while (lock.test_and_set(std::memory_order_acquire)) {}
    bool found = false;
    for (int i = 0; i < ndone; ++i)
        if (done[i].dev == dev && done[i].fn == fn) { found = true; break; }
    bool ok = true;
    if (!found) {
        ok = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) == cudaSuccess;
        if (ok && ndone < 512) done[ndone++] = Key{dev, fn};
    }
    lock.clear(std::memory_order_release);
    return ok;
}

}  // namespace bdq

extern "C" int bdq_version(void) { return 100; }
extern "C" const char* bdq_last_error(void) { return bdq::g_err; }
extern "C" int64_t bdq_launch_count(void) { return bdq::g_launches.load(std::memory_order_relaxed); }
