// Library-wide plumbing of the C ABI: error reporting and device facts.
#include "common.cuh"
#include <stdarg.h>
#include <stdio.h>
#include <atomic>

static thread_local char g_err[512] = "";

void dc_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static std::atomic<long long> g_launches{0};

int dc_check_launch(const char* what) { return dc_check_launch_n(what, 1); }

int dc_check_launch_n(const char* what, int n) {
    g_launches += n;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return DC_OK;
    dc_set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
}

int dc_sm_count() {
    static thread_local int cached_dev = -1, cached = 148;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) cached = n;
        cached_dev = dev;
    }
    return cached;
}

extern "C" const char* dc_last_error(void) { return g_err; }
extern "C" int dc_abi_version(void) { return 2; }      // 2: round-2 signature changes
// kernels launched by this library since the last reset (bench.py reports it as gpu_launches)
extern "C" long long dc_launch_count(void) { return g_launches.load(); }
extern "C" void dc_reset_launch_count(void) { g_launches = 0; }
extern "C" int dc_device_sm_count(void) { return dc_sm_count(); }
