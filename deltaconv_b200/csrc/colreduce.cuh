// Deterministic column reductions shared by the BatchNorm-type kernels (nnops.cu, central.cu):
// per-CTA fp64 partial sums over a row chunk (fixed grid), then a finalize step that adds the
// partials in a fixed order.  No atomics anywhere, so results are bit-reproducible.
#pragma once
#include "common.cuh"

namespace {

constexpr int RED_T = 256;       // threads per CTA in the column reductions
constexpr int RED_MAX_CHUNKS = 1184;   // 8 x 148

// ------------------------------------------------------------------------------------------
// generic column reduction: every element (r, c) contributes NV values; partial[chunk][v][c]
// ------------------------------------------------------------------------------------------
struct ColLayout {
    int col_lanes;     // threads along the columns
    int row_lanes;     // threads along the rows (col_lanes * row_lanes == RED_T)
    int col_blocks;    // grid.y : ceil(C / col_lanes)
    int chunks;        // grid.x
    long long rows_per_chunk;
};

ColLayout col_layout(long long M, int C) {
    ColLayout L;
    int cl = 1;
    while (cl < C && cl < 64) cl <<= 1;
    L.col_lanes = cl;
    L.row_lanes = RED_T / cl;
    L.col_blocks = (C + cl - 1) / cl;
    long long want = (long long)dc_sm_count() * 4 / L.col_blocks;     // measured: 1, 2, 6 and 8 chunks per SM are all slower
    if (want < 1) want = 1;
    long long by_rows = (M + L.row_lanes * 4 - 1) / (L.row_lanes * 4);   // at least 4 rows per thread
    if (by_rows < 1) by_rows = 1;
    long long chunks = want < by_rows ? want : by_rows;
    if (chunks > RED_MAX_CHUNKS) chunks = RED_MAX_CHUNKS;
    L.chunks = (int)chunks;
    L.rows_per_chunk = (M + chunks - 1) / chunks;
    return L;
}

template <int NV, class F>
__global__ void __launch_bounds__(RED_T) col_reduce_kernel(F f, long long M, int C, int col_lanes, long long rows_per_chunk,
                                                           double* __restrict__ partial) {
    __shared__ double sh[NV][RED_T];
    const int cl = threadIdx.x % col_lanes, rl = threadIdx.x / col_lanes, row_lanes = RED_T / col_lanes;
    const int c = blockIdx.y * col_lanes + cl;
    const long long r0 = blockIdx.x * rows_per_chunk;
    const long long r1 = min(M, r0 + rows_per_chunk);
    double acc[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) acc[v] = 0.0;
    if (c < C) {
        long long r = r0 + rl;
        // eight independent rows in flight per thread (the loads do not depend on the accumulators): these reductions are
        // bound by latency x bytes in flight, and fewer, longer chunks keep the finalize step short
        for (; r + 7LL * row_lanes < r1; r += 8LL * row_lanes) {
            float v0[NV], v1[NV], v2[NV], v3[NV], v4[NV], v5[NV], v6[NV], v7[NV];
            f(r, c, v0); f(r + row_lanes, c, v1); f(r + 2LL * row_lanes, c, v2); f(r + 3LL * row_lanes, c, v3);
            f(r + 4LL * row_lanes, c, v4); f(r + 5LL * row_lanes, c, v5); f(r + 6LL * row_lanes, c, v6); f(r + 7LL * row_lanes, c, v7);
#pragma unroll
            for (int v = 0; v < NV; ++v)
                acc[v] += (((double)v0[v] + (double)v1[v]) + ((double)v2[v] + (double)v3[v])) +
                          (((double)v4[v] + (double)v5[v]) + ((double)v6[v] + (double)v7[v]));
        }
        for (; r + 3LL * row_lanes < r1; r += 4LL * row_lanes) {
            float v0[NV], v1[NV], v2[NV], v3[NV];
            f(r, c, v0); f(r + row_lanes, c, v1); f(r + 2LL * row_lanes, c, v2); f(r + 3LL * row_lanes, c, v3);
#pragma unroll
            for (int v = 0; v < NV; ++v) acc[v] += ((double)v0[v] + (double)v1[v]) + ((double)v2[v] + (double)v3[v]);
        }
        for (; r < r1; r += row_lanes) {
            float vals[NV];
            f(r, c, vals);
#pragma unroll
            for (int v = 0; v < NV; ++v) acc[v] += (double)vals[v];
        }
    }
#pragma unroll
    for (int v = 0; v < NV; ++v) sh[v][threadIdx.x] = acc[v];
    __syncthreads();
    if (rl == 0 && c < C) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            double s = 0.0;
            for (int q = 0; q < row_lanes; ++q) s += sh[v][q * col_lanes + cl];
            partial[((size_t)blockIdx.x * NV + v) * C + c] = s;
        }
    }
}


// The same reduction with a thread owning 4 adjacent columns and fetching them with one 16-byte load per row (functors that
// have a `quad` method; C % 4 == 0, 16-byte aligned rows): a quarter of the load instructions and four times the bytes in
// flight per thread, which is what these latency-bound passes over narrow matrices (64-256 columns) need.
struct NoColumns {};
struct BnColumns { float mean[4], invstd[4], gamma[4], beta[4]; };

ColLayout col_layout_vec4(long long M, int C) {
    ColLayout L;
    const int c4 = C / 4;
    int cl = 1;
    while (cl < c4 && cl < 64) cl <<= 1;
    L.col_lanes = cl;
    L.row_lanes = RED_T / cl;
    L.col_blocks = (c4 + cl - 1) / cl;
    long long want = (long long)dc_sm_count() * 4 / L.col_blocks;
    if (want < 1) want = 1;
    long long by_rows = (M + L.row_lanes * 4 - 1) / (L.row_lanes * 4);
    if (by_rows < 1) by_rows = 1;
    long long chunks = want < by_rows ? want : by_rows;
    if (chunks > RED_MAX_CHUNKS) chunks = RED_MAX_CHUNKS;
    L.chunks = (int)chunks;
    L.rows_per_chunk = (M + chunks - 1) / chunks;
    return L;
}

template <int NV, class F>
__global__ void __launch_bounds__(RED_T) col_reduce_vec4_kernel(F f, long long M, int C, int col_lanes, long long rows_per_chunk,
                                                                double* __restrict__ partial) {
    __shared__ double sh[NV][4][RED_T];
    const int cl = threadIdx.x % col_lanes, rl = threadIdx.x / col_lanes, row_lanes = RED_T / col_lanes;
    const int c = (blockIdx.y * col_lanes + cl) * 4;
    const long long r0 = blockIdx.x * rows_per_chunk;
    const long long r1 = min(M, r0 + rows_per_chunk);
    double acc[4][NV];
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < NV; ++v) acc[u][v] = 0.0;
    if (c < C) {
        const auto cols = f.columns(c);                                 // per-column constants of the functor, loaded once
        long long r = r0 + rl;
        for (; r + 3LL * row_lanes < r1; r += 4LL * row_lanes) {       // four rows = 16 values in flight per thread
            float v0[4][NV], v1[4][NV], v2[4][NV], v3[4][NV];
            f.quad(r, c, cols, v0); f.quad(r + row_lanes, c, cols, v1); f.quad(r + 2LL * row_lanes, c, cols, v2);
            f.quad(r + 3LL * row_lanes, c, cols, v3);
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < NV; ++v) acc[u][v] += ((double)v0[u][v] + (double)v1[u][v]) + ((double)v2[u][v] + (double)v3[u][v]);
        }
        for (; r < r1; r += row_lanes) {
            float vals[4][NV];
            f.quad(r, c, cols, vals);
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < NV; ++v) acc[u][v] += (double)vals[u][v];
        }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < NV; ++v) sh[v][u][threadIdx.x] = acc[u][v];
    __syncthreads();
    // row_lanes partial sums per column; thread (rl, cl) with rl < 4 adds up column c + rl in a fixed order
    if (rl < 4 && c < C) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            double t = 0.0;
            for (int q = 0; q < row_lanes; ++q) t += sh[v][rl][q * col_lanes + cl];
            partial[((size_t)blockIdx.x * NV + v) * C + c + rl] = t;
        }
    }
}

// reduction of f over [M,C] into partial[chunks][NV][C]; returns the number of chunks.  vec4_ok: the caller checked the
// alignment of everything the functor's quad() loads
template <int NV, class F>
int col_reduce_launch(const F& f, long long M, int C, bool vec4_ok, double* partial, cudaStream_t st) {
    if (vec4_ok && C % 4 == 0 && RED_T / col_layout_vec4(M, C).col_lanes >= 4) {
        const ColLayout L = col_layout_vec4(M, C);
        col_reduce_vec4_kernel<NV><<<dim3(L.chunks, L.col_blocks), RED_T, 0, st>>>(f, M, C, L.col_lanes, L.rows_per_chunk, partial);
        return L.chunks;
    }
    const ColLayout L = col_layout(M, C);
    col_reduce_kernel<NV><<<dim3(L.chunks, L.col_blocks), RED_T, 0, st>>>(f, M, C, L.col_lanes, L.rows_per_chunk, partial);
    return L.chunks;
}

// Sum partial[q][v][c] over the chunks q for one column c and value v.  A CTA of 256 threads owns 2 columns, four warps
// per column: each warp takes every fourth group of 32 chunks (with ~600 chunks that is 4-5 independent loads per lane, all
// in flight at once: the step is pure latency), lane sums are combined by a shuffle tree and the four warp sums in a
// fixed order through shared memory — deterministic.  Every thread of the CTA must call it (it holds a barrier).
constexpr int FIN_T = 256;
constexpr int FIN_SPLIT = 4;                       // warps per column
constexpr int FIN_COLS = FIN_T / 32 / FIN_SPLIT;   // columns per CTA
__device__ __forceinline__ int finalize_column() { return blockIdx.x * FIN_COLS + (threadIdx.x >> 5) / FIN_SPLIT; }

template <int NV>
__device__ __forceinline__ bool finalize_sums(const double* __restrict__ partial, int chunks, int C, double (&out)[NV]) {
    __shared__ double comb[NV][FIN_T / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, sub = w % FIN_SPLIT;
    const int c = finalize_column();
    constexpr int STRIDE = 32 * FIN_SPLIT;
    double acc[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) acc[v] = 0.0;
    if (c < C) {
        int q = sub * 32 + lane;
        // four chunks in flight per lane: the loads of one round are independent, only the adds are ordered
        for (; q + 3 * STRIDE < chunks; q += 4 * STRIDE) {
            double t0[NV], t1[NV], t2[NV], t3[NV];
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                t0[v] = partial[((size_t)q * NV + v) * C + c];
                t1[v] = partial[((size_t)(q + STRIDE) * NV + v) * C + c];
                t2[v] = partial[((size_t)(q + 2 * STRIDE) * NV + v) * C + c];
                t3[v] = partial[((size_t)(q + 3 * STRIDE) * NV + v) * C + c];
            }
#pragma unroll
            for (int v = 0; v < NV; ++v) acc[v] += (t0[v] + t1[v]) + (t2[v] + t3[v]);
        }
        for (; q < chunks; q += STRIDE)
#pragma unroll
            for (int v = 0; v < NV; ++v) acc[v] += partial[((size_t)q * NV + v) * C + c];
    }
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const double t = warp_sum(acc[v]);
        if (lane == 0) comb[v][w] = t;
    }
    __syncthreads();
    if (lane != 0 || sub != 0 || c >= C) return false;
#pragma unroll
    for (int v = 0; v < NV; ++v) out[v] = (comb[v][w] + comb[v][w + 1]) + (comb[v][w + 2] + comb[v][w + 3]);
    return true;
}

}  // namespace
