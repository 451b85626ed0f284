// K9 (first cut) — fp32 SIMT GEMM for the per-point channel mixing: torch.nn.Linear inside
// MLP / VectorMLP (/root/reference/deltaconv/nn/mlp.py:9,15) and the heads
// (deltanet_classification.py:33-36, deltanet_segmentation.py:40-51), forward and both backward
// products, exact fp32 (FFMA) so that the logits parity bound of 1e-4 holds with margin.
// The tensor-core (tcgen05) kernel for the large layers lives in gemm_tc.cu and is selected by
// dc_linear_* when the shape allows; this kernel remains the generic path (any M, N, K, strides).
//
//   C[i,j] = sum_p A(i,p) * B(p,j) (+ bias[j]),   A(i,p) = A[i*sa_i + p*sa_p],  B(p,j) = B[p*sb_p + j*sb_j]
#include "common.cuh"

namespace {

constexpr int BM = 128, BN = 64, BK = 16, GT = 256;

struct GemmArgs {
    const float* A; long long sa_i, sa_p;
    const float* B; long long sb_p, sb_j;
    float* C; long long ldc;
    const float* bias;
    int M, N, K;
    int k_per_split;        // reduction elements per grid.z slice
    float* partial;         // [splits][M][N] when splits > 1
};

template <bool A_PC, bool B_JC>
__global__ void __launch_bounds__(GT) sgemm_kernel(GemmArgs g) {
    __shared__ __align__(16) float As[BK][BM + 4];
    __shared__ __align__(16) float Bs[BK][BN + 4];
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    const int i0 = blockIdx.y * BM, j0 = blockIdx.x * BN;
    const int k_begin = blockIdx.z * g.k_per_split;
    const int k_end = min(g.K, k_begin + g.k_per_split);
    float acc[8][4];
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = 0.f;

    for (int k0 = k_begin; k0 < k_end; k0 += BK) {
#pragma unroll
        for (int q = 0; q < (BM * BK) / GT; ++q) {
            int e = t + GT * q;
            int p, i;
            if (A_PC) { p = e % BK; i = e / BK; } else { i = e % BM; p = e / BM; }
            float v = 0.f;
            if (i0 + i < g.M && k0 + p < k_end) v = __ldg(g.A + (long long)(i0 + i) * g.sa_i + (long long)(k0 + p) * g.sa_p);
            As[p][i] = v;
        }
#pragma unroll
        for (int q = 0; q < (BK * BN) / GT; ++q) {
            int e = t + GT * q;
            int p, j;
            if (B_JC) { j = e % BN; p = e / BN; } else { p = e % BK; j = e / BK; }
            float v = 0.f;
            if (j0 + j < g.N && k0 + p < k_end) v = __ldg(g.B + (long long)(k0 + p) * g.sb_p + (long long)(j0 + j) * g.sb_j);
            Bs[p][j] = v;
        }
        __syncthreads();
#pragma unroll
        for (int p = 0; p < BK; ++p) {
            float4 a0 = *reinterpret_cast<const float4*>(&As[p][ty * 8]);
            float4 a1 = *reinterpret_cast<const float4*>(&As[p][ty * 8 + 4]);
            float4 b = *reinterpret_cast<const float4*>(&Bs[p][tx * 4]);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(a[r], bb[c], acc[r][c]);
        }
        __syncthreads();
    }
    float* out = g.partial ? g.partial + (size_t)blockIdx.z * g.M * g.N : g.C;
    const long long ldo = g.partial ? g.N : g.ldc;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        int i = i0 + ty * 8 + r;
        if (i >= g.M) continue;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            int j = j0 + tx * 4 + c;
            if (j >= g.N) continue;
            float v = acc[r][c];
            if (!g.partial && g.bias) v += __ldg(g.bias + j);
            out[(long long)i * ldo + j] = v;
        }
    }
}

__global__ void splitk_reduce_kernel(const float* __restrict__ partial, int splits, long long MN, int N, float* __restrict__ C,
                                     long long ldc, const float* __restrict__ bias) {
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (t >= MN) return;
    float s = 0.f;
    for (int z = 0; z < splits; ++z) s += partial[(size_t)z * MN + t];     // fixed order: deterministic
    long long i = t / N; int j = (int)(t - i * N);
    if (bias) s += __ldg(bias + j);
    C[i * ldc + j] = s;
}

int choose_splits(int M, int N, int K) {
    long long tiles = (long long)dc_ceil_div(M, BM) * dc_ceil_div(N, BN);
    long long target = 2LL * dc_sm_count();
    if (tiles >= target || K <= 4 * BK) return 1;
    long long s = (target + tiles - 1) / tiles;
    long long maxs = K / (4 * BK);
    if (s > maxs) s = maxs;
    if (s > 128) s = 128;
    return (int)(s < 1 ? 1 : s);
}

int run_gemm(GemmArgs g, bool a_pc, bool b_jc, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (g.M <= 0 || g.N <= 0) return DC_OK;
    int splits = choose_splits(g.M, g.N, g.K);
    while (splits > 1 && (size_t)splits * g.M * g.N * sizeof(float) > ws_bytes) --splits;
    int kps = dc_ceil_div(dc_ceil_div(g.K, splits), BK) * BK;
    if (kps < BK) kps = BK;
    splits = g.K > 0 ? dc_ceil_div(g.K, kps) : 1;
    g.k_per_split = kps;
    g.partial = splits > 1 ? (float*)ws : nullptr;
    dim3 grid(dc_ceil_div(g.N, BN), dc_ceil_div(g.M, BM), splits);
    DC_REQUIRE(grid.y <= 65535u, "gemm: M too large for this launch shape");
    if (a_pc && b_jc) sgemm_kernel<true, true><<<grid, GT, 0, st>>>(g);
    else if (a_pc && !b_jc) sgemm_kernel<true, false><<<grid, GT, 0, st>>>(g);
    else if (!a_pc && b_jc) sgemm_kernel<false, true><<<grid, GT, 0, st>>>(g);
    else sgemm_kernel<false, false><<<grid, GT, 0, st>>>(g);
    if (splits > 1) {
        long long MN = (long long)g.M * g.N;
        splitk_reduce_kernel<<<dc_ceil_div(MN, 256), 256, 0, st>>>(g.partial, splits, MN, g.N, g.C, g.ldc, g.bias);
    }
    return dc_check_launch_n("gemm", splits > 1 ? 2 : 1);
}


// ---- weight gradient of a thin layer: dW[n, k] = sum_m dY[m, n] X[m, k] with K <= 16 (the first layer: 4 x 3 input channels) ----
// The tiled kernel spends a 128 x 64 tile on 64 x 12 outputs and needs ~100 split-K slices.  Here a CTA walks a chunk of rows in
// 64-row tiles staged in shared memory; thread (n, g) owns dW[n, g*KG .. g*KG+KG) in registers; per-CTA partials are added in a
// fixed order by a second small kernel (deterministic).
constexpr int THIN_ROWS = 64;
template <int KG>            // k values per thread; 256 / 64 = 4 groups cover K <= 4 * KG
__global__ void __launch_bounds__(256) thin_dw_kernel(const float* __restrict__ dY, long long ld_dy, const float* __restrict__ X, long long ldx,
                                                      int M, int K, int N, int rows_per_cta, float* __restrict__ partial) {
    __shared__ float dys[THIN_ROWS][64 + 1];
    __shared__ float xs[THIN_ROWS][4 * KG];
    const int n_local = threadIdx.x & 63, g = threadIdx.x >> 6;
    const int n0 = blockIdx.y * 64;
    const int r_begin = blockIdx.x * rows_per_cta, r_end = min(M, r_begin + rows_per_cta);
    const bool vec4 = (ld_dy & 3) == 0 && (N & 3) == 0 && (reinterpret_cast<uintptr_t>(dY) & 15u) == 0;
    float acc[KG];
#pragma unroll
    for (int u = 0; u < KG; ++u) acc[u] = 0.f;
    for (int r0 = r_begin; r0 < r_end; r0 += THIN_ROWS) {
        __syncthreads();
        if (vec4) {                 // 16-byte loads: 4 per thread and tile instead of 16
            for (int e = threadIdx.x; e < THIN_ROWS * 16; e += 256) {
                const int m = e >> 4, c = (e & 15) * 4;
                float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r0 + m < r_end && n0 + c < N) q = __ldg(reinterpret_cast<const float4*>(dY + (long long)(r0 + m) * ld_dy + n0 + c));
                dys[m][c] = q.x; dys[m][c + 1] = q.y; dys[m][c + 2] = q.z; dys[m][c + 3] = q.w;
            }
        } else {
            for (int e = threadIdx.x; e < THIN_ROWS * 64; e += 256) {
                const int m = e >> 6, c = e & 63;
                dys[m][c] = (r0 + m < r_end && n0 + c < N) ? __ldg(dY + (long long)(r0 + m) * ld_dy + n0 + c) : 0.f;
            }
        }
        for (int e = threadIdx.x; e < THIN_ROWS * 4 * KG; e += 256) {
            const int m = e / (4 * KG), c = e - m * 4 * KG;
            xs[m][c] = (r0 + m < r_end && c < K) ? __ldg(X + (long long)(r0 + m) * ldx + c) : 0.f;
        }
        __syncthreads();
#pragma unroll 8
        for (int m = 0; m < THIN_ROWS; ++m) {
            const float d = dys[m][n_local];
#pragma unroll
            for (int u = 0; u < KG; ++u) acc[u] = fmaf(d, xs[m][g * KG + u], acc[u]);
        }
    }
    if (n0 + n_local < N)
#pragma unroll
        for (int u = 0; u < KG; ++u)
            if (g * KG + u < K) partial[((size_t)blockIdx.x * N + n0 + n_local) * K + g * KG + u] = acc[u];
}

// one warp per output element: lanes take every 32nd partial (all of a lane's loads in flight), lane sums are combined by a
// shuffle tree — fixed order, deterministic (a thread per output walking ~600 partials one after the other took ~30 us)
__global__ void __launch_bounds__(256) thin_dw_reduce_kernel(const float* __restrict__ partial, int chunks, int NK, int K,
                                                             float* __restrict__ dW, long long ld_dw) {
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (t >= NK) return;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int q = lane;
    for (; q + 96 < chunks; q += 128) {
        s0 += partial[(size_t)q * NK + t]; s1 += partial[(size_t)(q + 32) * NK + t];
        s2 += partial[(size_t)(q + 64) * NK + t]; s3 += partial[(size_t)(q + 96) * NK + t];
    }
    for (; q < chunks; q += 32) s0 += partial[(size_t)q * NK + t];
    const float s = warp_sum((s0 + s1) + (s2 + s3));
    if (lane == 0) {
        const int n = t / K, k = t - n * K;
        dW[(long long)n * ld_dw + k] = s;
    }
}

int thin_dw_chunks(long long M) {
    long long c = (M + 4 * THIN_ROWS - 1) / (4 * THIN_ROWS);       // at least four tiles per CTA
    const long long cap = 6LL * dc_sm_count();                     // each CTA is one latency chain of tile loads: several per SM
    if (c > cap) c = cap;
    return (int)(c < 1 ? 1 : c);
}

// W [Cout, 2K] = [W1 | W2]  ->  W' [2Cout, 2K] = [[W1, -W2], [W2, W1]]
// so that VectorMLP's Linear on I_J(v) (deltaconv.py:68, operators.py:19-21) becomes one plain
// GEMM on the [N, 2K] view of the interleaved vector field: [v_x | v_y] W'^T = [out_x | out_y].
// Kp >= K is the padded half-width of W' ([2Cout, 2Kp], columns [K, Kp) and [Kp+K, 2Kp) zero): the vector concatenation it
// multiplies is stored with a row pitch that is a multiple of 4 floats so that every kernel touching it takes its 16-byte path.
// `col` (nullable) maps source column p of a half to its column in the padded half: the concatenation's blocks each start
// at a multiple of 4 floats.  W' must be zero-filled beforehand when Kp > K (the pad columns are not touched here).
__global__ void expand_vec_weight_kernel(const float* __restrict__ W, int Cout, int K, int Kp, const int* __restrict__ col,
                                         float* __restrict__ We) {
    long long total = (long long)Cout * K;
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (t >= total) return;
    int o = (int)(t / K), p = (int)(t - (long long)o * K);
    const float w1 = W[(size_t)o * 2 * K + p], w2 = W[(size_t)o * 2 * K + K + p];
    const int q = col ? __ldg(col + p) : p;
    We[(size_t)o * 2 * Kp + q] = w1;
    We[(size_t)o * 2 * Kp + Kp + q] = -w2;
    We[(size_t)(Cout + o) * 2 * Kp + q] = w2;
    We[(size_t)(Cout + o) * 2 * Kp + Kp + q] = w1;
}

__global__ void fold_vec_weight_grad_kernel(const float* __restrict__ dWe, int Cout, int K, int Kp, const int* __restrict__ col,
                                            float* __restrict__ dW) {
    long long total = (long long)Cout * K;
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (t >= total) return;
    int o = (int)(t / K), p = (int)(t - (long long)o * K);
    const int q = col ? __ldg(col + p) : p;
    dW[(size_t)o * 2 * K + p] = dWe[(size_t)o * 2 * Kp + q] + dWe[(size_t)(Cout + o) * 2 * Kp + Kp + q];
    dW[(size_t)o * 2 * K + K + p] = dWe[(size_t)(Cout + o) * 2 * Kp + q] - dWe[(size_t)o * 2 * Kp + Kp + q];
}

}  // namespace

// split-K workspace for a product with output M x N and reduction length K (0 when no split is used)
extern "C" size_t dc_gemm_ws_bytes(int64_t M, int64_t N, int64_t K) {
    int s = choose_splits((int)M, (int)N, (int)K);
    size_t simt = s > 1 ? (size_t)s * (size_t)M * (size_t)N * sizeof(float) : 0;
    size_t tc = dc_tc_ws_bytes(M, N, K);
    size_t need = simt > tc ? simt : tc;
    // thin weight gradient (dc_linear_bwd_weight passes output rows as M, reduction length as K): per-CTA partials
    if (N >= 1 && N <= 16 && M >= 1 && M <= 512 && K >= 1024) {
        const size_t thin = (size_t)thin_dw_chunks(K) * (size_t)M * (size_t)N * sizeof(float);
        if (thin > need) need = thin;
    }
    return need;
}

// Y[M,Nout] = X[M,K] W[Nout,K]^T (+ bias)              — torch.nn.Linear forward
extern "C" int dc_linear_fwd(const float* X, int64_t ldx, const float* W, int64_t ldw, const float* bias_or_null,
                             int64_t M, int K, int Nout, float* Y, int64_t ldy, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(M < 2147483647LL, "dc_linear_fwd: M must fit in int32");
    {
        int n = 0;
        int rc = dc_tc_gemm(X, ldx, 0, W, ldw, 0, bias_or_null, M, Nout, K, Y, ldy, ws, ws_bytes, (cudaStream_t)stream, &n);
        if (rc <= 0) return rc < 0 ? rc : dc_check_launch_n("dc_linear_fwd/tc", n);
    }
    GemmArgs g = {};
    g.A = X; g.sa_i = ldx; g.sa_p = 1;
    g.B = W; g.sb_p = 1; g.sb_j = ldw;
    g.C = Y; g.ldc = ldy; g.bias = bias_or_null; g.M = (int)M; g.N = Nout; g.K = K;
    return run_gemm(g, true, false, ws, ws_bytes, (cudaStream_t)stream);
}

// dX[M,K] = dY[M,Nout] W[Nout,K]                       — Linear backward w.r.t. the input
extern "C" int dc_linear_bwd_input(const float* dY, int64_t ld_dy, const float* W, int64_t ldw, int64_t M, int K, int Nout,
                                   float* dX, int64_t ld_dx, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(M < 2147483647LL, "dc_linear_bwd_input: M must fit in int32");
    {   // A = dY (reduction over Nout, contiguous); B = W stored [Nout, K] i.e. reduction rows, output index contiguous
        int n = 0;
        int rc = dc_tc_gemm(dY, ld_dy, 0, W, ldw, 1, nullptr, M, K, Nout, dX, ld_dx, ws, ws_bytes, (cudaStream_t)stream, &n);
        if (rc <= 0) return rc < 0 ? rc : dc_check_launch_n("dc_linear_bwd_input/tc", n);
    }
    GemmArgs g = {};
    g.A = dY; g.sa_i = ld_dy; g.sa_p = 1;
    g.B = W; g.sb_p = ldw; g.sb_j = 1;
    g.C = dX; g.ldc = ld_dx; g.M = (int)M; g.N = K; g.K = Nout;
    return run_gemm(g, true, true, ws, ws_bytes, (cudaStream_t)stream);
}

// dW[Nout,K] = dY[M,Nout]^T X[M,K]                     — Linear backward w.r.t. the weight
extern "C" int dc_linear_bwd_weight(const float* dY, int64_t ld_dy, const float* X, int64_t ldx, int64_t M, int K, int Nout,
                                    float* dW, int64_t ld_dw, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(M < 2147483647LL, "dc_linear_bwd_weight: M must fit in int32");
    {   // reduction over the M rows: both operands have their output index contiguous (MN-major)
        int n = 0;
        int rc = dc_tc_gemm(dY, ld_dy, 1, X, ldx, 1, nullptr, Nout, K, M, dW, ld_dw, ws, ws_bytes, (cudaStream_t)stream, &n);
        if (rc <= 0) return rc < 0 ? rc : dc_check_launch_n("dc_linear_bwd_weight/tc", n);
    }
    if (K >= 1 && K <= 16 && Nout >= 1 && Nout <= 512 && M >= 1024) {      // thin layer: see thin_dw_kernel
        const int chunks = thin_dw_chunks(M);
        const size_t need = (size_t)chunks * Nout * K * sizeof(float);
        if (ws_bytes >= need) {
            const int rows_per_cta = (int)((M + chunks - 1) / chunks);
            dim3 grid(chunks, dc_ceil_div(Nout, 64));
            if (K <= 4) thin_dw_kernel<1><<<grid, 256, 0, (cudaStream_t)stream>>>(dY, ld_dy, X, ldx, (int)M, K, Nout, rows_per_cta, (float*)ws);
            else if (K <= 8) thin_dw_kernel<2><<<grid, 256, 0, (cudaStream_t)stream>>>(dY, ld_dy, X, ldx, (int)M, K, Nout, rows_per_cta, (float*)ws);
            else if (K <= 12) thin_dw_kernel<3><<<grid, 256, 0, (cudaStream_t)stream>>>(dY, ld_dy, X, ldx, (int)M, K, Nout, rows_per_cta, (float*)ws);
            else thin_dw_kernel<4><<<grid, 256, 0, (cudaStream_t)stream>>>(dY, ld_dy, X, ldx, (int)M, K, Nout, rows_per_cta, (float*)ws);
            thin_dw_reduce_kernel<<<dc_ceil_div((long long)Nout * K * 32, 256), 256, 0, (cudaStream_t)stream>>>((const float*)ws, chunks, Nout * K, K, dW, ld_dw);
            return dc_check_launch_n("dc_linear_bwd_weight/thin", 2);
        }
    }
    GemmArgs g = {};
    g.A = dY; g.sa_i = 1; g.sa_p = ld_dy;
    g.B = X; g.sb_p = ldx; g.sb_j = 1;
    g.C = dW; g.ldc = ld_dw; g.M = Nout; g.N = K; g.K = (int)M;
    return run_gemm(g, false, true, ws, ws_bytes, (cudaStream_t)stream);
}

extern "C" int dc_expand_vec_weight(const float* W, int Cout, int K, int K_padded, const int32_t* col_map_or_null,
                                    float* W_expanded, void* stream) {
    if (Cout <= 0 || K <= 0) return DC_OK;
    DC_REQUIRE(K_padded >= K, "dc_expand_vec_weight: padded width %d < width %d", K_padded, K);
    if (K_padded > K) DC_CUDA(cudaMemsetAsync(W_expanded, 0, (size_t)4 * Cout * K_padded * sizeof(float), (cudaStream_t)stream));
    expand_vec_weight_kernel<<<dc_ceil_div((long long)Cout * K, 256), 256, 0, (cudaStream_t)stream>>>(W, Cout, K, K_padded, col_map_or_null, W_expanded);
    return dc_check_launch("dc_expand_vec_weight");
}

extern "C" int dc_fold_vec_weight_grad(const float* dW_expanded, int Cout, int K, int K_padded, const int32_t* col_map_or_null,
                                       float* dW, void* stream) {
    if (Cout <= 0 || K <= 0) return DC_OK;
    DC_REQUIRE(K_padded >= K, "dc_fold_vec_weight_grad: padded width %d < width %d", K_padded, K);
    fold_vec_weight_grad_kernel<<<dc_ceil_div((long long)Cout * K, 256), 256, 0, (cudaStream_t)stream>>>(dW_expanded, Cout, K, K_padded, col_map_or_null, dW);
    return dc_check_launch("dc_fold_vec_weight_grad");
}

