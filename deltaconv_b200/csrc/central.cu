// Fused "centralized" first DeltaConv layer (reference: /root/reference/deltaconv/nn/deltaconv.py:50-52):
//
//     x_edge = x[col] - x[row]                      [E, Cin]      E = N*k edge rows
//     x_max  = scatter_max(s_mlp_max(x_edge), row)  Linear(Cin->Cout, no bias) -> BatchNorm over the E rows
//                                                   -> LeakyReLU -> max over each point's k edges
//
// The reference (and the unfused path in nnops.cu) materialises [E,Cin] and [E,Cout] (168 MB at 32x1024x20,
// Cout=64) several times.  With Cin <= 4 the edge feature t[e,c] = w_c . d_e costs Cin FMAs, so it is
// recomputed wherever it is needed and nothing of size E is ever stored:
//   * BatchNorm statistics follow exactly from the first two moments of the edge offsets d_e:
//         mean_c = w_c . m,   var_c = w_c^T (S - m m^T) w_c,      m = E[d], S = E[d d^T]       (one pass over edges, fp64)
//   * forward: one warp per point recomputes t for its k edges, normalises, activates, takes the max (+ arg slot)
//   * backward: only the arg-max edge of each (point, channel) carries dz; the BatchNorm mean terms that touch
//     every edge reduce to the same moments, so dW, dgamma, dbeta need one reduction over the N points:
//         S1_c = sum dz, S2_c = sum dz*xhat, A_c = sum dz * d_argmax
//         dW_c = gamma_c invstd_c ( A_c - S1_c m - S2_c invstd_c (S w_c - mean_c m) )
//   * the gradient w.r.t. x (rarely needed: x is the input point cloud) is formed per edge and reduced over the
//     incoming-edge CSR.
#include "common.cuh"
#include "colreduce.cuh"

extern "C" int dc_edge_diff_bwd(const int32_t* in_ptr, const int32_t* in_edge, int N, int k, int C, const float* d_edge,
                                float* dx, int64_t ld_dx, void* stream);

namespace {

constexpr int CIN_MAX = 4;
constexpr int NMOM = CIN_MAX + CIN_MAX * CIN_MAX;     // m[4], S[4][4]

// ---- pass 1: first and second moments of the edge offsets -------------------------------------------
__global__ void __launch_bounds__(256) edge_moments_kernel(const int* __restrict__ idx, int N, int k, int Cin,
                                                           const float* __restrict__ x, long long ldx,
                                                           double* __restrict__ partial) {
    __shared__ double sh[8][NMOM];
    double acc[NMOM];
#pragma unroll
    for (int q = 0; q < NMOM; ++q) acc[q] = 0.0;
    const long long E = (long long)N * k;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < E; e += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(e / k);
        int j = __ldg(idx + e);
        if ((unsigned)j >= (unsigned)N) j = i;
        float d[CIN_MAX];
#pragma unroll
        for (int q = 0; q < CIN_MAX; ++q) d[q] = q < Cin ? __ldg(x + (size_t)j * ldx + q) - __ldg(x + (size_t)i * ldx + q) : 0.f;
#pragma unroll
        for (int q = 0; q < CIN_MAX; ++q) {
            acc[q] += (double)d[q];
#pragma unroll
            for (int p = 0; p < CIN_MAX; ++p) acc[CIN_MAX + q * CIN_MAX + p] += (double)d[q] * (double)d[p];
        }
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int q = 0; q < NMOM; ++q) {
        double v = warp_sum(acc[q]);
        if (lane == 0) sh[w][q] = v;
    }
    __syncthreads();
    if (threadIdx.x < NMOM) {
        double s = 0.0;
        for (int q = 0; q < 8; ++q) s += sh[q][threadIdx.x];
        partial[(size_t)blockIdx.x * NMOM + threadIdx.x] = s;
    }
}

// moments (means over the E edges, fp64) and the per-channel BatchNorm statistics derived from them
__global__ void central_stats_finalize_kernel(const double* __restrict__ partial, int chunks, long long E, int Cin, int Cout,
                                              const float* __restrict__ W, float eps, float momentum,
                                              double* __restrict__ moments, float* __restrict__ mean, float* __restrict__ invstd,
                                              float* __restrict__ running_mean, float* __restrict__ running_var,
                                              long long* __restrict__ num_batches_tracked) {
    __shared__ double mom[NMOM];
    if (num_batches_tracked && threadIdx.x == 0) *num_batches_tracked += 1;
    {   // one warp per moment (blockDim.x = 32 * NMOM), lanes stride over the per-CTA partials, fixed-order shuffle tree
        const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
        double s = 0.0;
        for (int q = lane; q < chunks; q += 32) s += partial[(size_t)q * NMOM + w];
        s = warp_sum(s) / (double)E;
        if (lane == 0) { mom[w] = s; moments[w] = s; }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < Cout; c += blockDim.x) {
        double mu = 0.0, e2 = 0.0;
        for (int q = 0; q < Cin; ++q) {
            double wq = (double)W[(size_t)c * Cin + q];
            mu += wq * mom[q];
            for (int p = 0; p < Cin; ++p) e2 += wq * (double)W[(size_t)c * Cin + p] * mom[CIN_MAX + q * CIN_MAX + p];
        }
        double var = e2 - mu * mu;
        if (var < 0.0) var = 0.0;
        mean[c] = (float)mu;
        invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
        if (running_mean) {
            double unbiased = E > 1 ? var * (double)E / (double)(E - 1) : var;
            running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mu;
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
        }
    }
}

// edge offsets of point i held by the warp: slot e lives in lane e&31, register e>>5
__device__ __forceinline__ void load_offsets(const int* __restrict__ idx, int N, int k, int Cin, const float* __restrict__ x,
                                             long long ldx, int i, int lane, float (&d)[2][CIN_MAX]) {
    float xi[CIN_MAX];
#pragma unroll
    for (int q = 0; q < CIN_MAX; ++q) xi[q] = q < Cin ? __ldg(x + (size_t)i * ldx + q) : 0.f;
#pragma unroll
    for (int s = 0; s < 2; ++s) {
        const int e = lane + 32 * s;
        int j = i;
        if (e < k) { j = __ldg(idx + (size_t)i * k + e); if ((unsigned)j >= (unsigned)N) j = i; }
#pragma unroll
        for (int q = 0; q < CIN_MAX; ++q) d[s][q] = (e < k && q < Cin) ? __ldg(x + (size_t)j * ldx + q) - xi[q] : 0.f;
    }
}

// ---- forward: recompute, normalise, activate, max over the k edges ---------------------------------------
__global__ void __launch_bounds__(256) central_apply_kernel(const int* __restrict__ idx, int N, int k, int Cin, int Cout,
                                                            const float* __restrict__ x, long long ldx, const float* __restrict__ W,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            const float* __restrict__ mean, const float* __restrict__ invstd, float slope,
                                                            float* __restrict__ out, long long ldo, unsigned char* __restrict__ arg) {
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= N) return;
    const int i = (int)warp;
    float d[2][CIN_MAX];
    load_offsets(idx, N, k, Cin, x, ldx, i, lane, d);
    for (int c0 = 0; c0 < Cout; c0 += 32) {
        const int c = c0 + lane;
        const bool ok = c < Cout;
        float w[CIN_MAX];
#pragma unroll
        for (int q = 0; q < CIN_MAX; ++q) w[q] = (ok && q < Cin) ? __ldg(W + (size_t)c * Cin + q) : 0.f;
        const float mu = ok ? __ldg(mean + c) : 0.f, is = ok ? __ldg(invstd + c) : 0.f;
        const float g = (ok && gamma) ? __ldg(gamma + c) : 1.f, b = (ok && beta) ? __ldg(beta + c) : 0.f;
        float best = -INFINITY;
        int slot = 0;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const int cnt = min(32, k - 32 * s);
            for (int e = 0; e < cnt; ++e) {
                float t = 0.f;
#pragma unroll
                for (int q = 0; q < CIN_MAX; ++q) t = fmaf(w[q], __shfl_sync(FULL, d[s][q], e), t);
                float z = g * ((t - mu) * is) + b;
                z = z > 0.f ? z : slope * z;
                if (z > best) { best = z; slot = e + 32 * s; }
            }
        }
        if (ok) { out[(size_t)i * ldo + c] = best; arg[(size_t)i * Cout + c] = (unsigned char)slot; }
    }
}

// ---- backward: one reduction over the points ----------------------------------------------------------
template <int CIN>
struct CentralBwdVal {
    const int* idx; int N, k, Cout;
    const float* x; long long ldx; const float* W; const float* gamma; const float* beta; const float* mean; const float* invstd;
    float slope; const unsigned char* arg; const float* dout; long long ldd;
    __device__ void operator()(long long r, int c, float (&o)[2 + CIN]) const {
        const int i = (int)r;
        const int e = arg[(size_t)i * Cout + c];
        int j = __ldg(idx + (size_t)i * k + e);
        if ((unsigned)j >= (unsigned)N) j = i;
        float t = 0.f, d[CIN];
#pragma unroll
        for (int q = 0; q < CIN; ++q) {
            d[q] = __ldg(x + (size_t)j * ldx + q) - __ldg(x + (size_t)i * ldx + q);
            t = fmaf(__ldg(W + (size_t)c * CIN + q), d[q], t);
        }
        const float xh = (t - __ldg(mean + c)) * __ldg(invstd + c);
        const float g = gamma ? __ldg(gamma + c) : 1.f, b = beta ? __ldg(beta + c) : 0.f;
        const float z = g * xh + b;
        const float dz = __ldg(dout + (size_t)i * ldd + c) * (z > 0.f ? 1.f : slope);
        o[0] = dz; o[1] = dz * xh;
#pragma unroll
        for (int q = 0; q < CIN; ++q) o[2 + q] = dz * d[q];
    }
};

template <int CIN>
__global__ void __launch_bounds__(FIN_T) central_bwd_finalize_kernel(const double* __restrict__ partial, int chunks, int Cout,
                                                                     const float* __restrict__ W, const float* __restrict__ gamma,
                                                                     const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                     const double* __restrict__ moments, int training,
                                                                     float* __restrict__ dW, float* __restrict__ dgamma, float* __restrict__ dbeta) {
    double r[2 + CIN];
    if (!finalize_sums<2 + CIN>(partial, chunks, Cout, r)) return;
    const int c = finalize_column();
    dbeta[c] = (float)r[0];
    dgamma[c] = (float)r[1];
    const double is = (double)invstd[c], a = (gamma ? (double)gamma[c] : 1.0) * is, mu = (double)mean[c];
#pragma unroll
    for (int q = 0; q < CIN; ++q) {
        double v = r[2 + q];
        if (training) {
            double Sw = 0.0;
#pragma unroll
            for (int p = 0; p < CIN; ++p) Sw += moments[CIN_MAX + q * CIN_MAX + p] * (double)W[(size_t)c * CIN + p];
            v -= r[0] * moments[q] + r[1] * is * (Sw - mu * moments[q]);
        }
        dW[(size_t)c * CIN + q] = (float)(a * v);
    }
}

// per-edge gradient w.r.t. the offset d_e (only when the caller needs d/dx): g[e,:] = sum_c dt[e,c] W[c,:]
__global__ void __launch_bounds__(256) central_bwd_edge_kernel(const int* __restrict__ idx, int N, int k, int Cin, int Cout,
                                                               const float* __restrict__ x, long long ldx, const float* __restrict__ W,
                                                               const float* __restrict__ gamma, const float* __restrict__ beta,
                                                               const float* __restrict__ mean, const float* __restrict__ invstd, float slope,
                                                               const unsigned char* __restrict__ arg, const float* __restrict__ dout, long long ldd,
                                                               const float* __restrict__ S1, const float* __restrict__ S2, int training,
                                                               float* __restrict__ g_out) {
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= N) return;
    const int i = (int)warp;
    float d[2][CIN_MAX];
    load_offsets(idx, N, k, Cin, x, ldx, i, lane, d);
    const float invE = 1.f / ((float)N * (float)k);
    for (int e = 0; e < k; ++e) {
        float de[CIN_MAX];
#pragma unroll
        for (int q = 0; q < CIN_MAX; ++q) de[q] = __shfl_sync(FULL, d[e >> 5][q], e & 31);
        float acc[CIN_MAX] = {0.f, 0.f, 0.f, 0.f};
        for (int c = lane; c < Cout; c += 32) {
            float w[CIN_MAX], t = 0.f;
#pragma unroll
            for (int q = 0; q < CIN_MAX; ++q) { w[q] = q < Cin ? __ldg(W + (size_t)c * Cin + q) : 0.f; t = fmaf(w[q], de[q], t); }
            const float is = __ldg(invstd + c), xh = (t - __ldg(mean + c)) * is;
            const float gm = gamma ? __ldg(gamma + c) : 1.f, b = beta ? __ldg(beta + c) : 0.f;
            float dt = 0.f;
            if (arg[(size_t)i * Cout + c] == (unsigned char)e) {
                const float z = gm * xh + b;
                dt = __ldg(dout + (size_t)i * ldd + c) * (z > 0.f ? 1.f : slope);
            }
            if (training) dt -= (__ldg(S1 + c) + xh * __ldg(S2 + c)) * invE;
            dt *= gm * is;
#pragma unroll
            for (int q = 0; q < CIN_MAX; ++q) acc[q] = fmaf(dt, w[q], acc[q]);
        }
#pragma unroll
        for (int q = 0; q < CIN_MAX; ++q) acc[q] = warp_sum(acc[q]);
        if (lane == 0)
            for (int q = 0; q < Cin; ++q) g_out[((size_t)i * k + e) * Cin + q] = acc[q];
    }
}

// ---- deeper edge MLPs (s_mlp_max of depth >= 2: the ShapeNet configuration) --------------------------------
// Only the FIRST block can be recomputed from the offsets; its activation h[e,c] = act(bn(w_c . d_e)) feeds a Linear over the
// E rows and has to exist.  It is written once, straight from the point coordinates: the [E,Cin] offsets, the [E,Cout]
// pre-activation, the statistics pass over it and the normalise pass (write, read, read + write of E x Cout floats) never
// happen — the statistics are the moment formula above.
__global__ void __launch_bounds__(256) central_edge_fwd_kernel(const int* __restrict__ idx, int N, int k, int Cin, int Cout,
                                                               const float* __restrict__ x, long long ldx, const float* __restrict__ W,
                                                               const float* __restrict__ gamma, const float* __restrict__ beta,
                                                               const float* __restrict__ mean, const float* __restrict__ invstd, float slope,
                                                               float* __restrict__ out) {
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= N) return;
    const int i = (int)warp;
    float d[2][CIN_MAX];
    load_offsets(idx, N, k, Cin, x, ldx, i, lane, d);
    float* row = out + (size_t)i * k * Cout;
    for (int c0 = 0; c0 < Cout; c0 += 32) {
        const int c = c0 + lane;
        const bool ok = c < Cout;
        float w[CIN_MAX];
#pragma unroll
        for (int q = 0; q < CIN_MAX; ++q) w[q] = (ok && q < Cin) ? __ldg(W + (size_t)c * Cin + q) : 0.f;
        const float mu = ok ? __ldg(mean + c) : 0.f, is = ok ? __ldg(invstd + c) : 0.f;
        const float g = (ok && gamma) ? __ldg(gamma + c) : 1.f, b = (ok && beta) ? __ldg(beta + c) : 0.f;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const int cnt = min(32, k - 32 * s);
            for (int e = 0; e < cnt; ++e) {
                float t = 0.f;
#pragma unroll
                for (int q = 0; q < CIN_MAX; ++q) t = fmaf(w[q], __shfl_sync(FULL, d[s][q], e), t);
                float z = g * ((t - mu) * is) + b;
                z = z > 0.f ? z : slope * z;
                if (ok) __stcs(row + (size_t)(e + 32 * s) * Cout + c, z);       // 128 contiguous bytes per warp store
            }
        }
    }
}

// Backward of that block from the DENSE gradient dh [E,Cout] of its activation: the same three sums as CentralBwdVal, taken
// over all E edge rows in one pass (dh is read once; offsets and pre-activation are recomputed), and the same finalize.
// Replaces the BatchNorm backward (two reads of y and dh, one write) and the weight-gradient product over E rows.
// A warp owns a point (its k offsets live in the lanes, as in the forward kernel), a lane owns a channel: every load of dh is
// 128 contiguous bytes, six edge rows are in flight per warp, a point's k terms are summed in fp32 and the points of a warp in
// fp64.  partial[block][v][c] has the layout of the column reductions (fixed grid and order: deterministic).
template <int CIN>
__global__ void __launch_bounds__(256) central_dense_bwd_kernel(const int* __restrict__ idx, int N, int k, int Cout, const float* __restrict__ x,
                                                                long long ldx, const float* __restrict__ W, const float* __restrict__ gamma,
                                                                const float* __restrict__ beta, const float* __restrict__ mean,
                                                                const float* __restrict__ invstd, float slope, const float* __restrict__ dh,
                                                                long long ldd, double* __restrict__ partial) {
    constexpr int NV = 2 + CIN;
    __shared__ double sh[8][NV][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int first = blockIdx.x * 8 + warp, stride = gridDim.x * 8;
    for (int c0 = 0; c0 < Cout; c0 += 32) {
        const int c = c0 + lane;
        const bool ok = c < Cout;
        float w[CIN];
#pragma unroll
        for (int q = 0; q < CIN; ++q) w[q] = ok ? __ldg(W + (size_t)c * CIN + q) : 0.f;
        const float mu = ok ? __ldg(mean + c) : 0.f, is = ok ? __ldg(invstd + c) : 0.f;
        const float g = (ok && gamma) ? __ldg(gamma + c) : 1.f, b = (ok && beta) ? __ldg(beta + c) : 0.f;
        double acc[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) acc[v] = 0.0;
        for (int i = first; i < N; i += stride) {
            float d[2][CIN_MAX];
            load_offsets(idx, N, k, CIN, x, ldx, i, lane, d);
            const float* row = dh + (size_t)i * k * ldd + (ok ? c : 0);
            float s[NV];
#pragma unroll
            for (int v = 0; v < NV; ++v) s[v] = 0.f;
            for (int e0 = 0; e0 < k; e0 += 6) {
                float dv[6];
#pragma unroll
                for (int u = 0; u < 6; ++u) dv[u] = (ok && e0 + u < k) ? __ldcs(row + (size_t)(e0 + u) * ldd) : 0.f;
#pragma unroll
                for (int u = 0; u < 6; ++u) {
                    const int e = e0 + u;
                    if (e < k) {                                       // warp-uniform
                        float de[CIN], t = 0.f;
#pragma unroll
                        for (int q = 0; q < CIN; ++q) { de[q] = __shfl_sync(FULL, e < 32 ? d[0][q] : d[1][q], e & 31); t = fmaf(w[q], de[q], t); }
                        const float xh = (t - mu) * is;
                        const float z = g * xh + b;
                        const float dz = dv[u] * (z > 0.f ? 1.f : slope);
                        s[0] += dz; s[1] = fmaf(dz, xh, s[1]);
#pragma unroll
                        for (int q = 0; q < CIN; ++q) s[2 + q] = fmaf(dz, de[q], s[2 + q]);
                    }
                }
            }
#pragma unroll
            for (int v = 0; v < NV; ++v) acc[v] += (double)s[v];
        }
#pragma unroll
        for (int v = 0; v < NV; ++v) sh[warp][v][lane] = acc[v];
        __syncthreads();
        if (warp < NV && ok) {                                         // warp v adds the 8 warp sums of value v in a fixed order
            double t = 0.0;
#pragma unroll
            for (int q = 0; q < 8; ++q) t += sh[q][warp][lane];
            partial[((size_t)blockIdx.x * NV + warp) * Cout + c] = t;
        }
        __syncthreads();
    }
}

int moments_chunks(long long E) {
    long long g = (E + 255) / 256, cap = (long long)dc_sm_count() * 4;
    return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

template <int CIN>
int launch_bwd_reduce(const CentralBwdVal<CIN>& f, int N, int Cout, const float* W, const float* gamma, const float* mean,
                      const float* invstd, const double* moments, int training, float* dW, float* dgamma, float* dbeta,
                      double* partial, cudaStream_t st) {
    ColLayout L = col_layout(N, Cout);
    dim3 grid(L.chunks, L.col_blocks);
    col_reduce_kernel<2 + CIN><<<grid, RED_T, 0, st>>>(f, (long long)N, Cout, L.col_lanes, L.rows_per_chunk, partial);
    central_bwd_finalize_kernel<CIN><<<dc_ceil_div(Cout, FIN_COLS), FIN_T, 0, st>>>(partial, L.chunks, Cout, W, gamma, mean, invstd, moments,
                                                                             training, dW, dgamma, dbeta);
    return DC_OK;
}

}  // namespace

extern "C" int dc_central_max_cin(void) { return CIN_MAX; }

extern "C" size_t dc_central_ws_bytes(int N, int k, int Cin, int Cout) {
    size_t a = (size_t)RED_MAX_CHUNKS * (2 + CIN_MAX) * (size_t)Cout * sizeof(double);     // backward reduction partials
    size_t b = (size_t)dc_sm_count() * 4 * NMOM * sizeof(double);                          // moment partials
    size_t c = (size_t)N * k * Cin * sizeof(float);                                        // per-edge gradient (d/dx only)
    return (a > b ? a : b) + c + 256;
}

// Batch statistics of the edge MLP's pre-activation (training mode).  moments [20] doubles (m[4], S[4][4]) are kept
// for the backward pass; mean/invstd [Cout]; running stats updated like torch.nn.BatchNorm1d.
extern "C" int dc_central_stats(const int32_t* idx, int N, int k, int Cin, int Cout, const float* x, int64_t ldx, const float* W,
                                float eps, float momentum, double* moments, float* mean, float* invstd,
                                float* running_mean_or_null, float* running_var_or_null, int64_t* num_batches_tracked_or_null,
                                void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(Cin >= 1 && Cin <= CIN_MAX, "dc_central_stats: Cin must be in [1,%d]", CIN_MAX);
    DC_REQUIRE(N >= 1 && k >= 1 && Cout >= 1, "dc_central_stats: empty input");
    DC_REQUIRE(ws_bytes >= dc_central_ws_bytes(N, k, Cin, Cout), "dc_central_stats: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const long long E = (long long)N * k;
    const int chunks = moments_chunks(E);
    edge_moments_kernel<<<chunks, 256, 0, st>>>(idx, N, k, Cin, x, ldx, (double*)ws);
    central_stats_finalize_kernel<<<1, 32 * NMOM, 0, st>>>((const double*)ws, chunks, E, Cin, Cout, W, eps, momentum, moments, mean, invstd,
                                                     running_mean_or_null, running_var_or_null, (long long*)num_batches_tracked_or_null);
    return dc_check_launch_n("dc_central_stats", 2);
}

extern "C" int dc_central_fwd(const int32_t* idx, int N, int k, int Cin, int Cout, const float* x, int64_t ldx, const float* W,
                              const float* gamma_or_null, const float* beta_or_null, const float* mean, const float* invstd, float slope,
                              float* out, int64_t ldo, uint8_t* arg, void* stream) {
    DC_REQUIRE(Cin >= 1 && Cin <= CIN_MAX, "dc_central_fwd: Cin must be in [1,%d]", CIN_MAX);
    DC_REQUIRE(k >= 1 && k <= 64, "dc_central_fwd: k must be in [1,64]");
    if (N <= 0 || Cout <= 0) return DC_OK;
    central_apply_kernel<<<dc_ceil_div((long long)N * 32, 256), 256, 0, (cudaStream_t)stream>>>(idx, N, k, Cin, Cout, x, ldx, W, gamma_or_null,
                                                                                                 beta_or_null, mean, invstd, slope, out, ldo, arg);
    return dc_check_launch("dc_central_fwd");
}

extern "C" int dc_central_bwd(const int32_t* idx, const int32_t* in_ptr_or_null, const int32_t* in_edge_or_null, int N, int k, int Cin, int Cout,
                              const float* x, int64_t ldx, const float* W, const float* gamma_or_null, const float* beta_or_null,
                              const float* mean, const float* invstd, const double* moments, float slope, int training,
                              const uint8_t* arg, const float* dout, int64_t ldd, float* dW, float* dgamma, float* dbeta,
                              float* dx_or_null, int64_t ld_dx, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(Cin >= 1 && Cin <= CIN_MAX, "dc_central_bwd: Cin must be in [1,%d]", CIN_MAX);
    DC_REQUIRE(k >= 1 && k <= 64, "dc_central_bwd: k must be in [1,64]");
    DC_REQUIRE(ws_bytes >= dc_central_ws_bytes(N, k, Cin, Cout), "dc_central_bwd: workspace too small");
    DC_REQUIRE(dx_or_null == nullptr || (in_ptr_or_null && in_edge_or_null), "dc_central_bwd: d/dx needs the incoming-edge structure");
    if (N <= 0 || Cout <= 0) return DC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    double* partial = (double*)ws;
#define DC_CENTRAL_CASE(CIN)                                                                                                      \
    case CIN: {                                                                                                                   \
        CentralBwdVal<CIN> f = {idx, N, k, Cout, x, ldx, W, gamma_or_null, beta_or_null, mean, invstd, slope, arg, dout, ldd};   \
        launch_bwd_reduce<CIN>(f, N, Cout, W, gamma_or_null, mean, invstd, moments, training, dW, dgamma, dbeta, partial, st);   \
        break;                                                                                                                    \
    }
    switch (Cin) { DC_CENTRAL_CASE(1) DC_CENTRAL_CASE(2) DC_CENTRAL_CASE(3) DC_CENTRAL_CASE(4) }
#undef DC_CENTRAL_CASE
    int launches = 2;
    if (dx_or_null) {
        size_t off = ((size_t)RED_MAX_CHUNKS * (2 + CIN_MAX) * (size_t)Cout * sizeof(double) + 255) / 256 * 256;
        size_t off2 = (size_t)dc_sm_count() * 4 * NMOM * sizeof(double);
        if (off2 > off) off = (off2 + 255) / 256 * 256;
        float* g = (float*)((char*)ws + off);
        central_bwd_edge_kernel<<<dc_ceil_div((long long)N * 32, 256), 256, 0, st>>>(idx, N, k, Cin, Cout, x, ldx, W, gamma_or_null, beta_or_null,
                                                                                      mean, invstd, slope, arg, dout, ldd, dbeta, dgamma, training, g);
        int rc = dc_check_launch_n("dc_central_bwd", launches + 1);
        if (rc) return rc;
        return dc_edge_diff_bwd(in_ptr_or_null, in_edge_or_null, N, k, Cin, g, dx_or_null, ld_dx, stream);
    }
    return dc_check_launch_n("dc_central_bwd", launches);
}

// First block of a deeper edge MLP: h[e, :] = act(bn(W (x_j - x_i))) for every edge slot, [N*k, Cout] row-major (the layout
// dc_edge_diff_fwd + Linear + dc_bn_act_fwd would produce).  mean / invstd come from dc_central_stats (training) or the
// running statistics (evaluation).
extern "C" int dc_central_edge_fwd(const int32_t* idx, int N, int k, int Cin, int Cout, const float* x, int64_t ldx, const float* W,
                                   const float* gamma_or_null, const float* beta_or_null, const float* mean, const float* invstd, float slope,
                                   float* out, void* stream) {
    DC_REQUIRE(Cin >= 1 && Cin <= CIN_MAX, "dc_central_edge_fwd: Cin must be in [1,%d]", CIN_MAX);
    DC_REQUIRE(k >= 1 && k <= 64, "dc_central_edge_fwd: k must be in [1,64]");
    if (N <= 0 || Cout <= 0) return DC_OK;
    central_edge_fwd_kernel<<<dc_ceil_div((long long)N * 32, 256), 256, 0, (cudaStream_t)stream>>>(idx, N, k, Cin, Cout, x, ldx, W, gamma_or_null,
                                                                                                    beta_or_null, mean, invstd, slope, out);
    return dc_check_launch("dc_central_edge_fwd");
}

// Its backward from dh [N*k, Cout] (row pitch ldd): dW [Cout,Cin], dgamma, dbeta.  No gradient w.r.t. x (callers that need it
// use the unfused sequence).  Workspace: dc_central_ws_bytes.
extern "C" int dc_central_edge_bwd(const int32_t* idx, int N, int k, int Cin, int Cout, const float* x, int64_t ldx, const float* W,
                                   const float* gamma_or_null, const float* beta_or_null, const float* mean, const float* invstd,
                                   const double* moments, float slope, int training, const float* dh, int64_t ldd, float* dW, float* dgamma,
                                   float* dbeta, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(Cin >= 1 && Cin <= CIN_MAX, "dc_central_edge_bwd: Cin must be in [1,%d]", CIN_MAX);
    DC_REQUIRE(k >= 1 && k <= 64, "dc_central_edge_bwd: k must be in [1,64]");
    DC_REQUIRE((long long)N * k < (1LL << 31), "dc_central_edge_bwd: more than 2^31 edge rows");
    DC_REQUIRE(ws_bytes >= dc_central_ws_bytes(N, k, Cin, Cout), "dc_central_edge_bwd: workspace too small");
    if (N <= 0 || Cout <= 0) return DC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    double* partial = (double*)ws;
    int chunks = dc_ceil_div(N, 8);                            // a warp per point, 8 warps per CTA, up to 8 CTAs per SM
    if (chunks > RED_MAX_CHUNKS) chunks = RED_MAX_CHUNKS;
#define DC_CENTRAL_CASE(CIN)                                                                                                      \
    case CIN:                                                                                                                     \
        central_dense_bwd_kernel<CIN><<<chunks, 256, 0, st>>>(idx, N, k, Cout, x, ldx, W, gamma_or_null, beta_or_null, mean, invstd, slope,  \
                                                              dh, ldd, partial);                                                 \
        central_bwd_finalize_kernel<CIN><<<dc_ceil_div(Cout, FIN_COLS), FIN_T, 0, st>>>(partial, chunks, Cout, W, gamma_or_null, mean, invstd,  \
                                                                                 moments, training, dW, dgamma, dbeta);          \
        break;
    switch (Cin) { DC_CENTRAL_CASE(1) DC_CENTRAL_CASE(2) DC_CENTRAL_CASE(3) DC_CENTRAL_CASE(4) }
#undef DC_CENTRAL_CASE
    return dc_check_launch_n("dc_central_edge_bwd", 2);
}
