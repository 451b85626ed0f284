// Shared helpers for the deltaconv_b200 CUDA library (sm_100a).  No torch, no C++ types in the ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define DC_HD __host__ __device__ __forceinline__
#else
#define DC_HD inline
#endif

#define DC_OK 0
#define DC_EINVAL (-1)

// thread-local message behind dc_last_error()
void dc_set_error(const char* fmt, ...);
// returns 0 or the pending cudaError_t (and records its string)
int dc_check_launch(const char* what);
// same, and adds n to the library's launch counter (dc_launch_count)
int dc_check_launch_n(const char* what, int n);

#define DC_REQUIRE(cond, ...)            \
    do {                                 \
        if (!(cond)) {                   \
            dc_set_error(__VA_ARGS__);   \
            return DC_EINVAL;            \
        }                                \
    } while (0)

#define DC_CUDA(call)                                                    \
    do {                                                                 \
        cudaError_t _e = (call);                                         \
        if (_e != cudaSuccess) {                                         \
            dc_set_error("%s: %s", #call, cudaGetErrorString(_e));       \
            return (int)_e;                                              \
        }                                                                \
    } while (0)

static inline int dc_ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// tensor-core GEMM (gemm_tc.cu): returns 0 when it ran, 1 when the shape/layout is not eligible (caller falls
// back to the fp32 SIMT kernel), <0 on error; *launches = kernels launched
int dc_tc_gemm(const float* A, long long lda, int a_mn, const float* B, long long ldb, int b_mn, const float* bias,
               long long M, int N, long long K, float* C, long long ldc, void* ws, size_t ws_bytes, cudaStream_t st,
               int* launches);
size_t dc_tc_ws_bytes(long long M, long long N, long long K);

// cached SM count of the current device (148 on B200)
int dc_sm_count();

#if defined(__CUDACC__)
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL, v, o));
    return v;
}

// index of the cloud that owns point i: largest c with ptr[c] <= i   (ptr has B+1 entries)
__device__ __forceinline__ int cloud_of(const int* __restrict__ ptr, int B, int i) {
    int lo = 0, hi = B;  // invariant: ptr[lo] <= i < ptr[hi]
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (__ldg(ptr + mid) <= i) lo = mid; else hi = mid;
    }
    return lo;
}

// streaming (read-once / write-once) accesses that should not pollute L1
__device__ __forceinline__ float4 ld_stream4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w));
}
#endif
