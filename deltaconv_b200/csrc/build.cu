// K2/K3 — tangent frames and the fused WLS grad/div operator build, emitting fixed-k ELL.
//
// Reference path replaced (relative to /root/reference): deltaconv/geometry/grad_div_mls.py
//   :10-47 estimate_basis, :50-69 build_tangent_basis, :197-277 build_grad_div and its helpers —
//   ~70 ATen launches, [E,*] temporaries, two COO->CSR sorts become:
//     edge_mean_kernel  (pass A: per-point mean edge length)            -> ws.m[N]
//     cloud_mean_kernel (per-cloud mean of m, zero the cloud maxima)    -> ws.abar[B], ws.S[B]
//     fit_kernel   (pass B: the fit, ONCE; un-normalised entries, per-cloud max)  -> G[N,k,2], D[N,k,2], ws.S[B]
//     scale_kernel (pass C: entries x 1 / S_cloud, only if normalized; L2-resident) -> G, D
// One thread owns one point; its k slots live in shared memory slot-major (stride T+1, conflict
// free), which also turns the idx tile load and the G/D tile stores into fully coalesced
// streams.  The 6x6 normal equations are assembled from 15 moments and Cholesky-solved in registers.
#include "common.cuh"
#include "geom_core.cuh"

namespace {

__global__ void tangent_basis_kernel(const float* __restrict__ normal, int N, float* __restrict__ xb, float* __restrict__ yb) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    V3 n = v3_load(normal + 3 * (size_t)i), ex, ey;
    tangent_basis(n, ex, ey);
    xb[3 * (size_t)i] = ex.x; xb[3 * (size_t)i + 1] = ex.y; xb[3 * (size_t)i + 2] = ex.z;
    yb[3 * (size_t)i] = ey.x; yb[3 * (size_t)i + 1] = ey.y; yb[3 * (size_t)i + 2] = ey.z;
}

__global__ void estimate_basis_kernel(const float* __restrict__ pos, const int* __restrict__ idx, int N, int k,
                                      const float* __restrict__ orient, float* __restrict__ normal,
                                      float* __restrict__ xb, float* __restrict__ yb) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    V3 p = v3_load(pos + 3 * (size_t)i);
    double cov[6] = {0, 0, 0, 0, 0, 0};
    const int* row = idx + (size_t)i * k;
    for (int e = 0; e < k; ++e) {
        int j = __ldg(row + e);
        if (j < 0 || j >= N) j = 0;
        V3 d = v3_load(pos + 3 * (size_t)j) - p;
        cov[0] += (double)d.x * d.x; cov[1] += (double)d.x * d.y; cov[2] += (double)d.x * d.z;
        cov[3] += (double)d.y * d.y; cov[4] += (double)d.y * d.z; cov[5] += (double)d.z * d.z;
    }
    V3 n, ex, ey;
    frame_from_scatter(cov, orient ? orient + 3 * (size_t)i : nullptr, n, ex, ey);
    size_t o = 3 * (size_t)i;
    normal[o] = n.x; normal[o + 1] = n.y; normal[o + 2] = n.z;
    xb[o] = ex.x; xb[o + 1] = ex.y; xb[o + 2] = ex.z;
    yb[o] = ey.x; yb[o + 1] = ey.y; yb[o + 2] = ey.z;
}

// ---- gather records ----------------------------------------------------------------------------
// The fit gathers, per neighbour, its position (pass A) and its tangent frame (pass B).  From the [N,3] arrays that is 9
// scalar loads per neighbour, each touching 32 different sectors per warp: the kernel is then bound by the L1's
// sector-lookup rate, not by arithmetic or bandwidth (ncu, round 2: 15 % issue utilisation).  With a workspace that has room
// for them, positions are repacked as one 16-byte record and frames as one 32-byte record per point (one sector each,
// fetched with 1 + 2 LDG.128), which cuts the lookups per neighbour from 9 to 3.
__global__ void pack_kernel(const float* __restrict__ pos, const float* __restrict__ xb, const float* __restrict__ yb, int N,
                            float4* __restrict__ posp, float4* __restrict__ frm) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const V3 p = v3_load(pos + 3 * (size_t)i), x = v3_load(xb + 3 * (size_t)i), y = v3_load(yb + 3 * (size_t)i);
    posp[i] = make_float4(p.x, p.y, p.z, 0.f);
    frm[2 * (size_t)i] = make_float4(x.x, x.y, x.z, y.x);
    frm[2 * (size_t)i + 1] = make_float4(y.y, y.z, 0.f, 0.f);
}

template <bool PACKED>
__device__ __forceinline__ V3 gather_pos(const float* __restrict__ pos, const float4* __restrict__ posp, int j) {
    if (PACKED) { const float4 q = __ldg(posp + j); return v3(q.x, q.y, q.z); }
    return v3_load(pos + 3 * (size_t)j);
}

// ---- pass A: m_i = mean_e ||p_j - p_i||  (grad_div_mls.py:233, :112 inner mean) --------------
// One thread per point; the CTA's idx tile is loaded coalesced and transposed into shared memory (as in the fit), eight
// neighbour positions are in flight per thread.  (The first version spread a point over 8 lanes: 43 warp-instructions per
// point and 91 us per 2^20 points, more than the fit itself.)
constexpr int EDGE_T = 128;
template <bool PACKED>
__global__ void __launch_bounds__(EDGE_T)
edge_mean_kernel(const float* __restrict__ pos, const float4* __restrict__ posp, const int* __restrict__ idx, int N, int k,
                 float* __restrict__ m) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int STRIDE = EDGE_T + 1;
    int* s_idx = reinterpret_cast<int*>(smem_raw);                 // [k][STRIDE]
    const int base = blockIdx.x * EDGE_T;
    const int npts = min(EDGE_T, N - base);
    const int tile = npts * k;
    const int* gidx = idx + (size_t)base * k;
    for (int l = threadIdx.x, p = threadIdx.x / k, e = threadIdx.x % k; l < tile; l += EDGE_T) {
        s_idx[e * STRIDE + p] = __ldg(gidx + l);
        e += EDGE_T % k; p += EDGE_T / k;
        if (e >= k) { e -= k; ++p; }
    }
    __syncthreads();
    if (threadIdx.x >= npts) return;
    const int i = base + threadIdx.x;
    const int* jdx = s_idx + threadIdx.x;
    const V3 p = v3_load(pos + 3 * (size_t)i);
    float acc = 0.f;
#pragma unroll 1
    for (int e = 0; e < k; e += 8) {
        V3 pj[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
            if (e + u < k) {
                const int j = jdx[(e + u) * STRIDE];
                pj[u] = gather_pos<PACKED>(pos, posp, ((unsigned)j >= (unsigned)N) ? 0 : j);
            }
#pragma unroll
        for (int u = 0; u < 8; ++u)
            if (e + u < k) {
                const V3 d = pj[u] - p;
                const float d2 = dot(d, d);
                acc += d2 > 0.f ? d2 * dc_rsqrt(d2) : 0.f;
            }
    }
    m[i] = acc / (float)k;
}

// The same from the squared distances the kNN search already has (dc_knn / dc_knn_grid d2 output): no gather at all, one
// coalesced pass over [N,k] floats.  A padded slot (d2 = +inf) is measured against point 0 like the reference does.
__global__ void edge_mean_from_d2_kernel(const float* __restrict__ d2, const float* __restrict__ pos, int N, int k, float* __restrict__ m) {
    const int sub = threadIdx.x & 7;
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    float acc = 0.f;
    if (i < N) {
        for (int e = sub; e < k; e += 8) {
            float q = __ldg(d2 + (size_t)i * k + e);
            if (isinf(q)) { const V3 d = v3_load(pos) - v3_load(pos + 3 * (size_t)i); q = dot(d, d); }
            acc += q > 0.f ? q * dc_rsqrt(q) : 0.f;
        }
    }
    acc += __shfl_xor_sync(FULL, acc, 4);
    acc += __shfl_xor_sync(FULL, acc, 2);
    acc += __shfl_xor_sync(FULL, acc, 1);
    if (i < N && sub == 0) m[i] = acc / (float)k;
}

// ---- per-cloud mean of m (scatter_mean, grad_div_mls.py:112); fixed-order tree => deterministic
constexpr int CLOUD_MEAN_T = 1024;
__global__ void __launch_bounds__(CLOUD_MEAN_T) cloud_mean_kernel(const float* __restrict__ m, const int* __restrict__ cloud_ptr, int B,
                                                                  float* __restrict__ abar, unsigned* __restrict__ S) {
    __shared__ double part[CLOUD_MEAN_T];
    const int c = blockIdx.x;
    const int s = cloud_ptr[c], e = cloud_ptr[c + 1];
    // four independent partial sums per thread (a cloud of 2^20 points is 256 rounds instead of 4096 dependent adds)
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    int i = s + threadIdx.x;
    for (; i + 3 * CLOUD_MEAN_T < e; i += 4 * CLOUD_MEAN_T) {
        a0 += (double)m[i]; a1 += (double)m[i + CLOUD_MEAN_T]; a2 += (double)m[i + 2 * CLOUD_MEAN_T]; a3 += (double)m[i + 3 * CLOUD_MEAN_T];
    }
    for (; i < e; i += CLOUD_MEAN_T) a0 += (double)m[i];
    part[threadIdx.x] = (a0 + a1) + (a2 + a3);
    __syncthreads();
    for (int o = CLOUD_MEAN_T / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) part[threadIdx.x] += part[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        int cnt = e - s;
        abar[c] = (float)(part[0] / (double)(cnt > 0 ? cnt : 1));
        S[c] = 0u;
    }
}

// ---- the fit --------------------------------------------------------------------------------
// One thread per point, ONE pass (geom_core.cuh).  The k slots of the CTA's T points live slot-major in shared memory (stride
// T+1: conflict free), which also turns the idx tile load and the G/D tile stores into fully coalesced streams.  Gathers are
// issued four slots at a time before any dependent arithmetic (the compiler cannot hoist them across the shared-memory
// stores itself).  K is the compile-time neighbour count (0 = run-time k): with it the slot loops unroll and every
// shared-memory address is an immediate.  Entries are written WITHOUT the per-cloud normalisation; the per-cloud maxima go to
// S (one atomicMax per warp: s_i >= 0, so its bit pattern orders like the value and the integer max is exact and order
// independent, replacing scatter_max of grad_div_mls.py:259) and scale_kernel applies them.
constexpr int FIT_BATCH = 8;       // gathers in flight per thread (ncu: the kernel waits on its gathers at 16 warps per SM; 4 -> 8 measured below)
template <int T, int K, bool DEBUG, bool PACKED>
__global__ void __launch_bounds__(T)
fit_kernel(const float* __restrict__ pos, const float* __restrict__ normal, const float* __restrict__ xb,
           const float* __restrict__ yb, const float4* __restrict__ posp, const float4* __restrict__ frm,
           const int* __restrict__ idx, const int* __restrict__ cloud_ptr, int B,
           int N, int k_rt, int n0, int n1, float kernel_width, FitParams fp, int normalized, const float* __restrict__ abar,
           unsigned* __restrict__ S, float* __restrict__ G, float* __restrict__ D,
           float* __restrict__ dbg_coords, float* __restrict__ dbg_weights, float* __restrict__ dbg_mapping) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int STRIDE = T + 1;
    const int k = K ? K : k_rt;
    float4* s_dat = reinterpret_cast<float4*>(smem_raw);                                   // [k][STRIDE]
    int* s_idx = reinterpret_cast<int*>(smem_raw + (size_t)k * STRIDE * sizeof(float4));   // [k][STRIDE]

    const int base = n0 + blockIdx.x * T;
    const int npts = min(T, n1 - base);
    const int tile = npts * k;
    // coalesced idx tile load, transposed into slot-major shared memory
    const int* gidx = idx + (size_t)base * k;
    for (int l = threadIdx.x, p = threadIdx.x / k, e = threadIdx.x % k; l < tile; l += T) {
        s_idx[e * STRIDE + p] = __ldg(gidx + l);
        e += T % k; p += T / k;
        if (e >= k) { e -= k; ++p; }
    }
    __syncthreads();

    const int i = base + threadIdx.x;
    const bool active = threadIdx.x < npts;
    unsigned s_bits = 0u;
    int c = 0;
    if (active) {
        float4* dat = s_dat + threadIdx.x;
        const int* jdx = s_idx + threadIdx.x;
        const V3 p = v3_load(pos + 3 * (size_t)i), n = v3_load(normal + 3 * (size_t)i);
        const V3 ex = v3_load(xb + 3 * (size_t)i), ey = v3_load(yb + 3 * (size_t)i);
        c = cloud_of(cloud_ptr, B, i);
        const float sc = kernel_width * __ldg(abar + c);
        const float nl2 = -1.4426950408889634f / (sc * sc);
        // pass A, four gathers in flight
        FitSums sums;
        fit_sums_zero(sums);
#pragma unroll 1
        for (int e = 0; e < k; e += FIT_BATCH) {
            int j[FIT_BATCH];
            V3 pj[FIT_BATCH];
#pragma unroll
            for (int u = 0; u < FIT_BATCH; ++u)
                if (e + u < k) {
                    j[u] = jdx[(e + u) * STRIDE];
                    const int jj = ((unsigned)j[u] >= (unsigned)N) ? 0 : j[u];
                    pj[u] = gather_pos<PACKED>(pos, posp, jj);
                }
#pragma unroll
            for (int u = 0; u < FIT_BATCH; ++u)
                if (e + u < k) dat[(e + u) * STRIDE] = fit_slot_accumulate(pj[u], (unsigned)j[u] < (unsigned)N, p, n, ex, ey, nl2, sums);
        }
        float z1[6], z2[6], coef[6];
        const float inv_wnorm = fit_solve(sums, fp, z1, z2, coef);
        // pass B, neighbour frames of four slots in flight
        float* dc_ = DEBUG && dbg_coords ? dbg_coords + (size_t)i * k * 2 : nullptr;
        float* dw_ = DEBUG && dbg_weights ? dbg_weights + (size_t)i * k : nullptr;
        float* dm_ = DEBUG && dbg_mapping ? dbg_mapping + (size_t)i * k * 4 : nullptr;
        float ax = 0.f, ay = 0.f;
#pragma unroll 1
        for (int e = 0; e < k; e += FIT_BATCH) {
            V3 xj[FIT_BATCH], yj[FIT_BATCH];
#pragma unroll
            for (int u = 0; u < FIT_BATCH; ++u)
                if (e + u < k) {
                    const int j = jdx[(e + u) * STRIDE];
                    const int jj = ((unsigned)j >= (unsigned)N) ? 0 : j;
                    if (PACKED) {
                        const float4 a = __ldg(frm + 2 * (size_t)jj), b = __ldg(frm + 2 * (size_t)jj + 1);
                        xj[u] = v3(a.x, a.y, a.z); yj[u] = v3(a.w, b.x, b.y);
                    } else {
                        xj[u] = v3_load(xb + 3 * (size_t)jj); yj[u] = v3_load(yb + 3 * (size_t)jj);
                    }
                }
#pragma unroll
            for (int u = 0; u < FIT_BATCH; ++u)
                if (e + u < k) {
                    const float4 out = fit_slot_entry(xj[u], yj[u], dat[(e + u) * STRIDE], n, ex, ey, z1, z2, coef, inv_wnorm,
                                                      dc_ ? dc_ + 2 * (e + u) : nullptr, dw_ ? dw_ + (e + u) : nullptr,
                                                      dm_ ? dm_ + 4 * (e + u) : nullptr);
                    ax += fabsf(out.x); ay += fabsf(out.y);
                    dat[(e + u) * STRIDE] = out;
                }
        }
        s_bits = __float_as_uint(fit_row_norm(ax, ay));
    }
    if (normalized) {
        // one atomic per warp when the warp lies inside one cloud (points are sorted by cloud), else one per thread
        const int c_first = __shfl_sync(FULL, c, 0);
        const bool uniform = __all_sync(FULL, !active || c == c_first);
        if (uniform) {
            const unsigned m = __reduce_max_sync(FULL, s_bits);
            if ((threadIdx.x & 31) == 0 && active) atomicMax(S + c_first, m);
        } else if (active) {
            atomicMax(S + c, s_bits);
        }
    }
    __syncthreads();
    // coalesced G / D tile stores (float2 per slot)
    float2* gG = reinterpret_cast<float2*>(G) + (size_t)base * k;
    float2* gD = reinterpret_cast<float2*>(D) + (size_t)base * k;
    for (int l = threadIdx.x, p = threadIdx.x / k, e = threadIdx.x % k; l < tile; l += T) {
        float4 v = s_dat[e * STRIDE + p];
        gG[l] = make_float2(v.x, v.y);
        gD[l] = make_float2(v.z, v.w);
        e += T % k; p += T / k;
        if (e >= k) { e -= k; ++p; }
    }
}

// ---- the per-cloud normalisation (grad_div_mls.py:258-260, :271-272) -------------------------
// Multiplies the entries of the points [lo, hi) by 1 / S_cloud, where lo / hi are the starts of the clouds that contain
// points n0 / n1 (hi = N when n1 >= N): after the fit of the points [.., n1) exactly the clouds that END at or before n1 have
// their final S.  With one launch per chunk of points this runs while the chunk's entries are still in L2.
constexpr int SCALE_T = 256;
constexpr int SCALE_PTS = 128;
__device__ __forceinline__ float2 scaled(float2 v, float f) { return make_float2(v.x * f, v.y * f); }
__device__ __forceinline__ float4 scaled(float4 v, float f) { return make_float4(v.x * f, v.y * f, v.z * f, v.w * f); }

// V = float2: one slot per element, u = k elements per point; V = float4 (k even): two slots per element, u = k / 2
template <class V>
__global__ void __launch_bounds__(SCALE_T)
scale_kernel(V* __restrict__ G, V* __restrict__ D, const int* __restrict__ cloud_ptr, int B, const unsigned* __restrict__ S,
             int N, int u, int n0, int n1) {
    __shared__ float s_scale[SCALE_PTS];
    const int lo = n0 <= 0 ? 0 : __ldg(cloud_ptr + cloud_of(cloud_ptr, B, n0));
    const int hi = n1 >= N ? N : __ldg(cloud_ptr + cloud_of(cloud_ptr, B, n1));
    for (int base = lo + blockIdx.x * SCALE_PTS; base < hi; base += gridDim.x * SCALE_PTS) {
        const int npts = min(SCALE_PTS, hi - base);
        __syncthreads();
        if (threadIdx.x < npts) s_scale[threadIdx.x] = fit_scale(__uint_as_float(__ldg(S + cloud_of(cloud_ptr, B, base + threadIdx.x))));
        __syncthreads();
        V* gG = G + (size_t)base * u;
        V* gD = D + (size_t)base * u;
        const int tile = npts * u;
        for (int l = threadIdx.x, p = threadIdx.x / u, e = threadIdx.x % u; l < tile; l += SCALE_T) {
            const float f = s_scale[p];
            gG[l] = scaled(gG[l], f);
            gD[l] = scaled(gD[l], f);
            e += SCALE_T % u; p += SCALE_T / u;
            if (e >= u) { e -= u; ++p; }
        }
    }
}

// deferred normalisation: the factor of every point, for consumers that apply it to their result rows
__global__ void point_scale_kernel(const int* __restrict__ cloud_ptr, int B, const unsigned* __restrict__ S, int N, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) out[i] = fit_scale(__uint_as_float(__ldg(S + cloud_of(cloud_ptr, B, i))));
}

size_t fit_smem_bytes(int T, int k) { return (size_t)k * (T + 1) * (sizeof(float4) + sizeof(int)); }

struct FitArgs {
    const float *pos, *normal, *xb, *yb; const float4 *posp, *frm; const int *idx, *cloud_ptr; int B, N, k; float kernel_width; FitParams fp; int normalized;
    const float* abar; unsigned* S; float *G, *D, *dc, *dw, *dm;
};

template <int T, int K, bool DEBUG, bool PACKED>
int launch_fit_one(const FitArgs& a, int n0, int n1, cudaStream_t st) {
    const size_t smem = fit_smem_bytes(T, a.k);
    DC_CUDA(cudaFuncSetAttribute(fit_kernel<T, K, DEBUG, PACKED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    fit_kernel<T, K, DEBUG, PACKED><<<dc_ceil_div(n1 - n0, T), T, smem, st>>>(a.pos, a.normal, a.xb, a.yb, a.posp, a.frm, a.idx, a.cloud_ptr, a.B,
                                                                               a.N, a.k, n0, n1, a.kernel_width, a.fp, a.normalized, a.abar, a.S,
                                                                               a.G, a.D, a.dc, a.dw, a.dm);
    return dc_check_launch("dc_build_grad_div/fit");
}

template <int T, bool PACKED>
int launch_fit_k(const FitArgs& a, int n0, int n1, cudaStream_t st) {
    if (a.dc || a.dw || a.dm) return launch_fit_one<T, 0, true, PACKED>(a, n0, n1, st);
    switch (a.k) {
        case 10: return launch_fit_one<T, 10, false, PACKED>(a, n0, n1, st);
        case 20: return launch_fit_one<T, 20, false, PACKED>(a, n0, n1, st);
        case 30: return launch_fit_one<T, 30, false, PACKED>(a, n0, n1, st);
        case 40: return launch_fit_one<T, 40, false, PACKED>(a, n0, n1, st);
        default: return launch_fit_one<T, 0, false, PACKED>(a, n0, n1, st);
    }
}

template <int T>
int launch_fit(const FitArgs& a, int n0, int n1, cudaStream_t st) {
    return a.posp ? launch_fit_k<T, true>(a, n0, n1, st) : launch_fit_k<T, false>(a, n0, n1, st);
}

int g_build_chunk_points = 0;      // tuning switch (dc_set_build_chunk): points per fit/scale chunk, 0 = all points in one launch

}  // namespace

extern "C" int dc_tangent_basis(const float* normal, int N, float* xb, float* yb, void* stream) {
    if (N <= 0) return DC_OK;
    tangent_basis_kernel<<<dc_ceil_div(N, 256), 256, 0, (cudaStream_t)stream>>>(normal, N, xb, yb);
    return dc_check_launch("dc_tangent_basis");
}

extern "C" int dc_estimate_basis(const float* pos, const int32_t* idx, int N, int k, const float* orient_or_null,
                                 float* normal, float* xb, float* yb, void* stream) {
    DC_REQUIRE(k >= 1, "dc_estimate_basis: k must be >= 1");
    if (N <= 0) return DC_OK;
    estimate_basis_kernel<<<dc_ceil_div(N, 128), 128, 0, (cudaStream_t)stream>>>(pos, idx, N, k, orient_or_null, normal, xb, yb);
    return dc_check_launch("dc_estimate_basis");
}

// minimum: m[N] floats, abar[B] floats, S[B] uints; each region 256-byte aligned
static size_t build_ws_min_bytes(int N, int B) {
    size_t a = ((size_t)N * 4 + 255) / 256 * 256, b = ((size_t)B * 4 + 255) / 256 * 256;
    return a + 2 * b;
}
// What callers should provide: the minimum plus the 48-byte gather records per point (16 position + 32 frame).  A
// workspace of at least build_ws_min_bytes() but less than this still works, through the slower scalar gathers.
extern "C" size_t dc_build_ws_bytes(int N, int B) { return build_ws_min_bytes(N, B) + (((size_t)N * 48 + 255) / 256 * 256); }
extern "C" size_t dc_build_ws_min_bytes(int N, int B) { return build_ws_min_bytes(N, B); }

extern "C" int dc_build_grad_div(const float* pos, const float* normal, const float* xb, const float* yb,
                                 const int32_t* idx, const float* d2_or_null, const int32_t* cloud_ptr, int B, int N, int k,
                                 float kernel_width, float lam, float lam_shape, int use_shape, int normalized,
                                 float* G, float* D, float* point_scale_or_null, float* dbg_coords, float* dbg_weights, float* dbg_mapping,
                                 void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(normalized >= 0 && normalized <= 2, "dc_build_grad_div: normalized must be 0, 1 or 2 (deferred)");
    DC_REQUIRE(normalized != 2 || point_scale_or_null != nullptr, "dc_build_grad_div: normalized = 2 (deferred) needs the point_scale output");
    DC_REQUIRE(k >= 1 && k <= 256, "dc_build_grad_div: k must be in [1,256] (got %d)", k);
    DC_REQUIRE(B >= 1, "dc_build_grad_div: B must be >= 1");
    DC_REQUIRE(ws_bytes >= build_ws_min_bytes(N, B), "dc_build_grad_div: workspace too small (%zu < %zu)", ws_bytes,
               build_ws_min_bytes(N, B));
    if (N <= 0) return DC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t m_bytes = ((size_t)N * 4 + 255) / 256 * 256, b_bytes = ((size_t)B * 4 + 255) / 256 * 256;
    float* m = (float*)ws;
    float* abar = (float*)((char*)ws + m_bytes);
    unsigned* S = (unsigned*)((char*)ws + m_bytes + b_bytes);
    FitParams fp;
    fp.lam = lam; fp.lam_shape = lam_shape; fp.use_shape = use_shape;

    const bool packed = ws_bytes >= dc_build_ws_bytes(N, B);
    float4* posp = packed ? (float4*)((char*)ws + build_ws_min_bytes(N, B)) : nullptr;
    float4* frm = packed ? posp + N : nullptr;
    int rc;
    const size_t edge_smem = (size_t)k * (EDGE_T + 1) * 4;
    if (edge_smem > 48 * 1024) {
        DC_CUDA(cudaFuncSetAttribute(edge_mean_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)edge_smem));
        DC_CUDA(cudaFuncSetAttribute(edge_mean_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)edge_smem));
    }
    if (packed) {
        pack_kernel<<<dc_ceil_div(N, 256), 256, 0, st>>>(pos, xb, yb, N, posp, frm);
        rc = dc_check_launch_n("dc_build_grad_div/pack", 1);
        if (rc) return rc;
    }
    if (d2_or_null) edge_mean_from_d2_kernel<<<dc_ceil_div((long long)N * 8, 256), 256, 0, st>>>(d2_or_null, pos, N, k, m);
    else if (packed) edge_mean_kernel<true><<<dc_ceil_div(N, EDGE_T), EDGE_T, edge_smem, st>>>(pos, posp, idx, N, k, m);
    else edge_mean_kernel<false><<<dc_ceil_div(N, EDGE_T), EDGE_T, edge_smem, st>>>(pos, nullptr, idx, N, k, m);
    cloud_mean_kernel<<<B, CLOUD_MEAN_T, 0, st>>>(m, cloud_ptr, B, abar, S);
    rc = dc_check_launch_n("dc_build_grad_div/mean", 2);
    if (rc) return rc;
    FitArgs a = {pos, normal, xb, yb, posp, frm, idx, cloud_ptr, B, N, k, kernel_width, fp, normalized, abar, S, G, D, dbg_coords, dbg_weights, dbg_mapping};
    // threads per CTA: as many as keep two CTAs resident per SM (<= ~110 KB of shared memory each)
    const size_t budget = 110 * 1024;
    const int T = fit_smem_bytes(128, k) <= budget ? 128 : (fit_smem_bytes(64, k) <= budget ? 64 : 32);
    // Optional chunks of points (dc_set_build_chunk): the fit of a chunk followed at once by the normalisation of the clouds
    // it completed, while their 16 k bytes per point are still in L2.  Measured on 2^20 points (scripts/exp_build_chunk.py):
    // every chunk size LOSES against one launch each (0.57 ms) — 0.78 ms at 32 MB chunks, 0.66 ms at 80 MB — because each
    // short fit launch ends in a partially filled wave; so the default is one chunk.
    long long chunk = g_build_chunk_points > 0 ? g_build_chunk_points : N;
    chunk = (chunk + T - 1) / T * T;
    if (normalized != 1 || chunk > N) chunk = N;
    for (long long n0 = 0; n0 < N; n0 += chunk) {
        const int n1 = (int)(n0 + chunk < N ? n0 + chunk : N);
        if (T == 128) rc = launch_fit<128>(a, (int)n0, n1, st);
        else if (T == 64) rc = launch_fit<64>(a, (int)n0, n1, st);
        else rc = launch_fit<32>(a, (int)n0, n1, st);
        if (rc) return rc;
        if (normalized == 1) {
            const int tiles = dc_ceil_div(n1 - (int)n0, SCALE_PTS) + 1;
            const int cap = dc_sm_count() * 8;
            const int grid = tiles < cap ? tiles : cap;
            const bool wide = k % 2 == 0 && ((uintptr_t)G & 15u) == 0 && ((uintptr_t)D & 15u) == 0;
            if (wide) scale_kernel<float4><<<grid, SCALE_T, 0, st>>>((float4*)G, (float4*)D, cloud_ptr, B, S, N, k / 2, (int)n0, n1);
            else scale_kernel<float2><<<grid, SCALE_T, 0, st>>>((float2*)G, (float2*)D, cloud_ptr, B, S, N, k, (int)n0, n1);
            rc = dc_check_launch("dc_build_grad_div/scale");
            if (rc) return rc;
        }
    }
    if (normalized == 2) {
        point_scale_kernel<<<dc_ceil_div(N, 256), 256, 0, st>>>(cloud_ptr, B, S, N, point_scale_or_null);
        rc = dc_check_launch("dc_build_grad_div/point_scale");
        if (rc) return rc;
    }
    return DC_OK;
}

// ================================================================================================
// dc_build_grad_div_static — the operator build for callers that may not allocate device memory
// ================================================================================================
// Some hosts of this library cannot hand it a workspace: an inference server whose device memory is
// carved up once at start-up and whose plugins are forbidden to allocate afterwards, or a
// latency-critical loop in which cudaMalloc (which synchronises the device) is not acceptable and
// no spare buffer was provisioned.  For them the scratch of the build (per-point mean edge length,
// per-cloud mean, per-cloud infinity norm) lives in the fixed array below, which is part of the
// library image: every GPU that loads the library holds these ~1.03 MB until the process exits.
//
// Limits: N <= DC_STATIC_MAX_POINTS points and B <= DC_STATIC_MAX_CLOUDS clouds per call; larger
// batches are refused with an error (use dc_build_grad_div with a caller-provided workspace).
//
// Sharing: all calls on one GPU use the same array and it keeps whatever the previous call left.
// Nothing needs to be cleared by the caller: the per-cloud maxima are reset by the kernels of each
// call before they are accumulated, everything else is overwritten before it is read.  Because the
// array is shared, two calls whose kernels can overlap on the GPU (different streams, no
// dependency between them) race on it — serialise such calls.
//
// First call on each GPU: the array's device address is looked up once (cudaGetSymbolAddress, with
// that GPU current) and cached.  Make that first call outside of any stream capture.
//
// Which GPU: the one that is current when the call is made; its array is used.  The buffers and the stream must belong to
// that GPU (checked on the output buffer).
#include <mutex>

#define DC_STATIC_MAX_POINTS 262144
#define DC_STATIC_MAX_CLOUDS 4096
#define DC_STATIC_ALIGN(x) ((((size_t)(x)) + 255) / 256 * 256)
#define DC_STATIC_SCRATCH_BYTES (DC_STATIC_ALIGN((size_t)DC_STATIC_MAX_POINTS * 4) + 2 * DC_STATIC_ALIGN((size_t)DC_STATIC_MAX_CLOUDS * 4))

// In host code this name is only a handle for cudaGetSymbolAddress, never a device pointer.
static __device__ __align__(256) unsigned char g_build_scratch[DC_STATIC_SCRATCH_BYTES];

// tuning switch: points per fit/normalise chunk (0 = automatic: 32 MB of entries)
extern "C" int dc_set_build_chunk(int points) { g_build_chunk_points = points; return DC_OK; }

extern "C" int dc_build_static_max_points(void) { return DC_STATIC_MAX_POINTS; }
extern "C" int dc_build_static_max_clouds(void) { return DC_STATIC_MAX_CLOUDS; }

extern "C" int dc_build_grad_div_static(const float* pos, const float* normal, const float* xb, const float* yb,
                                        const int32_t* idx, const int32_t* cloud_ptr, int B, int N, int k,
                                        float kernel_width, float lam, float lam_shape, int use_shape, int normalized,
                                        float* G, float* D, void* stream) {
    DC_REQUIRE(N <= DC_STATIC_MAX_POINTS,
               "dc_build_grad_div_static: %d points exceed the built-in scratch (max %d); use dc_build_grad_div with a workspace",
               N, DC_STATIC_MAX_POINTS);
    DC_REQUIRE(B <= DC_STATIC_MAX_CLOUDS,
               "dc_build_grad_div_static: %d clouds exceed the built-in scratch (max %d); use dc_build_grad_div with a workspace",
               B, DC_STATIC_MAX_CLOUDS);
    DC_REQUIRE(build_ws_min_bytes(N, B) <= DC_STATIC_SCRATCH_BYTES, "dc_build_grad_div_static: internal size mismatch");
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static void* scratch_of_device[kMaxDevices] = {nullptr};
    int dev = -1;
    DC_CUDA(cudaGetDevice(&dev));
    DC_REQUIRE(dev >= 0 && dev < kMaxDevices, "dc_build_grad_div_static: device ordinal %d not supported", dev);
    // The built-in array used is the CURRENT device's.  A caller whose buffers (and stream) belong to another GPU would get
    // that GPU's kernels pointed at this GPU's array: refuse instead.
    cudaPointerAttributes where;
    DC_CUDA(cudaPointerGetAttributes(&where, G));
    DC_REQUIRE(where.type != cudaMemoryTypeDevice || where.device == dev,
               "dc_build_grad_div_static: the output buffer lives on device %d but device %d is current; make the buffers' "
               "device current before the call", where.device, dev);
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (scratch_of_device[dev] == nullptr) {       // set-up, once per device, with that device current
            void* p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, g_build_scratch);
            if (e != cudaSuccess || p == nullptr) {
                dc_set_error("dc_build_grad_div_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
                return e != cudaSuccess ? (int)e : DC_EINVAL;
            }
            scratch_of_device[dev] = p;
        }
        scratch = scratch_of_device[dev];
    }
    DC_REQUIRE(normalized == 0 || normalized == 1, "dc_build_grad_div_static: normalized must be 0 or 1");
    return dc_build_grad_div(pos, normal, xb, yb, idx, nullptr, cloud_ptr, B, N, k, kernel_width, lam, lam_shape, use_shape,
                             normalized, G, D, nullptr, nullptr, nullptr, nullptr, scratch, DC_STATIC_SCRATCH_BYTES, stream);
}
