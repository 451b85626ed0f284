// K4/K5/K6 (+ transposes) — gather-reduce SpMMs on the fixed-k ELL operators, and K11, the
// transposed (incoming-edge) structure that makes every backward pass a gather instead of a
// scatter with atomics.
//
// Reference call sites replaced (relative to /root/reference): every `SparseTensor @ dense`
// (deltaconv/models/deltanet_base.py:78; deltaconv/nn/deltaconv.py:57,66;
// deltaconv/geometry/operators.py:27,33,40,43) plus the J / I_J / cat copies around them
// (operators.py:9-21).  One kernel template covers all of them:
//
//   forward  (row i, slots e of the ELL row):      other = idx[i,e],  (c0,c1) = coef[i,e,:]
//   transpose(row j, incoming edges t of point j): other = edge/k,    (c0,c1) = coef[edge,:]
//
// A group of LPP lanes owns one row and VEC*LPP consecutive channels (float4 per lane when the
// layout allows), so every gathered row segment is one coalesced request; idx/coef are loaded
// once per group, coalesced, and broadcast with shuffles.  All pointers carry a row stride so
// that results land directly in column slices of the concatenated MLP inputs (no torch.cat).
#include "common.cuh"

namespace {

enum SpOp {
    OP_GRAD = 0,        // in0[o]            -> out0[2r], out0[2r+1]            acc0 += c0 s ; acc1 += c1 s
    OP_DIVCURL = 1,     // v[2o], v[2o+1]    -> out0[r]=div, out1[r]=curl, out2[r]=|v_r|
    OP_HODGE = 2,       // in0[o]=a, in1[o]=b-> out0[2r], out0[2r+1]            -(grad a + J grad b)
    OP_GRAD_T = 3,      // dV[2o], dV[2o+1]  -> out0[r]                         acc0 += c0 px + c1 py
    OP_DIVCURL_T = 4,   // dDiv[o], dCurl[o] -> out0[2r], out0[2r+1] (+ norm backward)
    OP_HODGE_T = 5,     // dO[2o], dO[2o+1]  -> out0[r]=da, out1[r]=db
};

struct SpArgs {
    const int* idx;        // forward: [N,k]
    const int* in_ptr;     // transpose: [N+1]
    const int* in_edge;    // transpose: [nnz] edge ids (i*k+e), ascending inside a row
    const float2* coef;    // G or D as float2 per edge
    int N, k, C;
    const float* in0; long long ld_in0;     // see SpOp for the meaning per operation
    const float* in1; long long ld_in1;
    const float* in2; long long ld_in2;     // OP_DIVCURL_T: dNorm [N,C]
    const float* in3; long long ld_in3;     // OP_DIVCURL_T: v [2N,C] (to rebuild the norm)
    float* out0; long long ld_out0;
    float* out1; long long ld_out1;
    float* out2; long long ld_out2;
    const float* cp_src; long long ld_cp_src;   // OP_DIVCURL: x [N,C], OP_HODGE: v [2N,C] — copied through into a column
    float* cp_dst; long long ld_cp_dst;         // slice of the concatenation buffer (saves a strided copy launch)
    const float* row_scale;                       // nullable: per-point factor of the operator (deferred per-cloud normalisation of
                                                  // build_grad_div, grad_div_mls.py:258-260).  All of a point's neighbours lie in its
                                                  // cloud, so for every one of the six operations the factor of the RESULT row applies
    const int* cloud_ptr; int B; int max_cloud;   // staged variant: clouds and the size of the largest one (0 = unknown)
};

template <int VEC> struct Vec;
template <> struct Vec<4> {
    float4 v;
    __device__ __forceinline__ static Vec zero() { Vec r; r.v = make_float4(0.f, 0.f, 0.f, 0.f); return r; }
    __device__ __forceinline__ static Vec load(const float* p) { Vec r; r.v = __ldg(reinterpret_cast<const float4*>(p)); return r; }
    __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = v; }
    __device__ __forceinline__ void fma(float a, const Vec& b) { v.x = fmaf(a, b.v.x, v.x); v.y = fmaf(a, b.v.y, v.y); v.z = fmaf(a, b.v.z, v.z); v.w = fmaf(a, b.v.w, v.w); }
    template <class F> __device__ __forceinline__ static Vec map2(const Vec& a, const Vec& b, F f) { Vec r; r.v.x = f(a.v.x, b.v.x); r.v.y = f(a.v.y, b.v.y); r.v.z = f(a.v.z, b.v.z); r.v.w = f(a.v.w, b.v.w); return r; }
};
template <> struct Vec<1> {
    float v;
    __device__ __forceinline__ static Vec zero() { Vec r; r.v = 0.f; return r; }
    __device__ __forceinline__ static Vec load(const float* p) { Vec r; r.v = __ldg(p); return r; }
    __device__ __forceinline__ void store(float* p) const { *p = v; }
    __device__ __forceinline__ void fma(float a, const Vec& b) { v = fmaf(a, b.v, v); }
    template <class F> __device__ __forceinline__ static Vec map2(const Vec& a, const Vec& b, F f) { Vec r; r.v = f(a.v, b.v); return r; }
};

template <int OP> struct OpTraits {
    static constexpr bool transposed = OP >= OP_GRAD_T;
    static constexpr bool two_inputs = OP != OP_GRAD;
};

// SCALED: multiply the result rows by a.row_scale (deferred per-cloud normalisation).  A template flag, not a run-time test:
// the factor costs the light kernels a register allocation step (grad^T: 32 -> 40 registers = 64 -> 48 resident warps, +12 %
// time), which the un-scaled instantiation must not pay; forcing the old allocation with launch bounds spills and is slower.
template <int OP, int VEC, bool SCALED>
__global__ void __launch_bounds__(256) spmm_kernel(SpArgs a, int lpp_log2) {
    typedef Vec<VEC> V;
    const int lpp = 1 << lpp_log2;
    const int lane = threadIdx.x & 31;
    const int sub = lane & (lpp - 1);
    const int rows_per_warp = 32 >> lpp_log2;
    const long long warp_global = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const long long row_ll = warp_global * rows_per_warp + (lane >> lpp_log2);
    const bool row_ok = row_ll < a.N;
    const int r = row_ok ? (int)row_ll : 0;
    const int ch = (blockIdx.y * lpp + sub) * VEC;
    const bool ch_ok = ch < a.C;
    // the row's normalisation factor, requested before the gather loop so that its latency is hidden behind it
    const float sc = (SCALED && row_ok) ? __ldg(a.row_scale + r) : 1.f;

    int t_begin, t_end;
    if (OpTraits<OP>::transposed) { t_begin = row_ok ? __ldg(a.in_ptr + r) : 0; t_end = row_ok ? __ldg(a.in_ptr + r + 1) : 0; }
    else { t_begin = 0; t_end = row_ok ? a.k : 0; }
    // the whole warp iterates to the longest row so that shuffles stay converged
    int len = t_end - t_begin, max_len = len;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) max_len = max(max_len, __shfl_xor_sync(FULL, max_len, o));

    V acc0 = V::zero(), acc1 = V::zero();
    for (int t0 = 0; t0 < max_len; t0 += lpp) {
        // cooperative, coalesced load of up to lpp (other, coef) pairs of this row
        int my_other = 0;
        float2 my_c = make_float2(0.f, 0.f);
        const int t = t0 + sub;
        if (t < len) {
            if (OpTraits<OP>::transposed) {
                int edge = __ldg(a.in_edge + t_begin + t);
                my_other = edge / a.k;
                my_c = __ldg(a.coef + edge);
            } else {
                size_t eo = (size_t)r * a.k + t;
                int j = __ldg(a.idx + eo);
                my_c = __ldg(a.coef + eo);
                if ((unsigned)j >= (unsigned)a.N) { j = 0; my_c = make_float2(0.f, 0.f); }   // padding sentinel
                my_other = j;
            }
        }
        const int cnt = min(lpp, max_len - t0);
#pragma unroll 4
        for (int q = 0; q < cnt; ++q) {
            const int o = __shfl_sync(FULL, my_other, q, lpp);
            const float c0 = __shfl_sync(FULL, my_c.x, q, lpp);
            const float c1 = __shfl_sync(FULL, my_c.y, q, lpp);
            if (!ch_ok || t0 + q >= len) continue;
            if (OP == OP_GRAD) {
                V s = V::load(a.in0 + (size_t)o * a.ld_in0 + ch);
                acc0.fma(c0, s); acc1.fma(c1, s);
            } else if (OP == OP_DIVCURL) {
                V vx = V::load(a.in0 + (size_t)(2 * o) * a.ld_in0 + ch), vy = V::load(a.in0 + (size_t)(2 * o + 1) * a.ld_in0 + ch);
                acc0.fma(c0, vx); acc0.fma(c1, vy);
                acc1.fma(c0, vy); acc1.fma(-c1, vx);
            } else if (OP == OP_HODGE) {
                V va = V::load(a.in0 + (size_t)o * a.ld_in0 + ch), vb = V::load(a.in1 + (size_t)o * a.ld_in1 + ch);
                acc0.fma(-c0, va); acc0.fma(c1, vb);
                acc1.fma(-c1, va); acc1.fma(-c0, vb);
            } else if (OP == OP_GRAD_T) {
                V px = V::load(a.in0 + (size_t)(2 * o) * a.ld_in0 + ch), py = V::load(a.in0 + (size_t)(2 * o + 1) * a.ld_in0 + ch);
                acc0.fma(c0, px); acc0.fma(c1, py);
            } else if (OP == OP_DIVCURL_T) {
                V dd = V::zero(), dc = V::zero();
                if (a.in0) dd = V::load(a.in0 + (size_t)o * a.ld_in0 + ch);
                if (a.in1) dc = V::load(a.in1 + (size_t)o * a.ld_in1 + ch);
                acc0.fma(c0, dd); acc0.fma(-c1, dc);
                acc1.fma(c1, dd); acc1.fma(c0, dc);
            } else {  // OP_HODGE_T
                V px = V::load(a.in0 + (size_t)(2 * o) * a.ld_in0 + ch), py = V::load(a.in0 + (size_t)(2 * o + 1) * a.ld_in0 + ch);
                acc0.fma(-c0, px); acc0.fma(-c1, py);
                acc1.fma(c1, px); acc1.fma(-c0, py);
            }
        }
    }
    if (!row_ok || !ch_ok) return;
    if (SCALED) {
        acc0 = V::map2(acc0, acc0, [sc](float p, float) { return p * sc; });
        acc1 = V::map2(acc1, acc1, [sc](float p, float) { return p * sc; });
    }
    if (OP == OP_DIVCURL && a.cp_src) V::load(a.cp_src + (size_t)r * a.ld_cp_src + ch).store(a.cp_dst + (size_t)r * a.ld_cp_dst + ch);
    if (OP == OP_HODGE && a.cp_src) {
        V::load(a.cp_src + (size_t)(2 * r) * a.ld_cp_src + ch).store(a.cp_dst + (size_t)(2 * r) * a.ld_cp_dst + ch);
        V::load(a.cp_src + (size_t)(2 * r + 1) * a.ld_cp_src + ch).store(a.cp_dst + (size_t)(2 * r + 1) * a.ld_cp_dst + ch);
    }
    if (OP == OP_GRAD || OP == OP_HODGE) {
        acc0.store(a.out0 + (size_t)(2 * r) * a.ld_out0 + ch);
        acc1.store(a.out0 + (size_t)(2 * r + 1) * a.ld_out0 + ch);
    } else if (OP == OP_DIVCURL) {
        if (a.out0) acc0.store(a.out0 + (size_t)r * a.ld_out0 + ch);
        if (a.out1) acc1.store(a.out1 + (size_t)r * a.ld_out1 + ch);
        if (a.out2) {
            V vx = V::load(a.in0 + (size_t)(2 * r) * a.ld_in0 + ch), vy = V::load(a.in0 + (size_t)(2 * r + 1) * a.ld_in0 + ch);
            V n = V::map2(vx, vy, [](float x, float y) { return sqrtf(x * x + y * y); });
            n.store(a.out2 + (size_t)r * a.ld_out2 + ch);
        }
    } else if (OP == OP_GRAD_T) {
        acc0.store(a.out0 + (size_t)r * a.ld_out0 + ch);
    } else if (OP == OP_DIVCURL_T) {
        if (a.in2) {   // d|v| : dv = dNorm * v / |v|   (0 where |v| = 0, torch's subgradient)
            V vx = V::load(a.in3 + (size_t)(2 * r) * a.ld_in3 + ch), vy = V::load(a.in3 + (size_t)(2 * r + 1) * a.ld_in3 + ch);
            V dn = V::load(a.in2 + (size_t)r * a.ld_in2 + ch);
            V inv = V::map2(vx, vy, [](float x, float y) { float n = sqrtf(x * x + y * y); return n > 0.f ? 1.f / n : 0.f; });
            V sx = V::map2(dn, inv, [](float d, float w) { return d * w; });
            acc0 = V::map2(acc0, V::map2(sx, vx, [](float s, float x) { return s * x; }), [](float p, float q) { return p + q; });
            acc1 = V::map2(acc1, V::map2(sx, vy, [](float s, float y) { return s * y; }), [](float p, float q) { return p + q; });
        }
        acc0.store(a.out0 + (size_t)(2 * r) * a.ld_out0 + ch);
        acc1.store(a.out0 + (size_t)(2 * r + 1) * a.ld_out0 + ch);
    } else {  // OP_HODGE_T
        acc0.store(a.out0 + (size_t)r * a.ld_out0 + ch);
        acc1.store(a.out1 + (size_t)r * a.ld_out1 + ch);
    }
}


// ---- shared-memory staged variant ---------------------------------------------------------------
// The gather formulation above reads every feature row k times from L2.  Clouds are small (a few
// thousand points), so one CTA stages the rows of ONE cloud for a chunk of CH channels in shared
// memory (one coalesced pass over HBM) and all k-fold gathers then hit shared memory.  Same
// arithmetic, same summation order per row as spmm_kernel — results are bitwise identical.
template <int OP> struct StagedTraits {
    // rows staged per point for input 0 / input 1 (0 = input not staged)
    static constexpr int m0 = (OP == OP_GRAD || OP == OP_HODGE || OP == OP_DIVCURL_T) ? 1 : 2;
    static constexpr int m1 = (OP == OP_HODGE || OP == OP_DIVCURL_T) ? 1 : 0;
};

// STAGE = false is the same cloud-major schedule WITHOUT the explicit copy: the CTA gathers straight from global memory and
// relies on the SM's L1 (no shared memory is requested, so nearly all of the 256 KB is cache) to hold the cloud's rows of
// this channel chunk — the k-fold re-reads then hit L1 instead of crossing to L2.  1024 threads per CTA in that mode.
template <int OP, bool STAGE>
__global__ void __launch_bounds__(STAGE ? 256 : 1024) spmm_staged_kernel(SpArgs a, int CH, int lpp_log2) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* t0 = reinterpret_cast<float*>(smem_raw);
    const int b = blockIdx.y;
    const int s = __ldg(a.cloud_ptr + b), e = __ldg(a.cloud_ptr + b + 1), P = e - s;
    if (P <= 0) return;
    const int ch0 = blockIdx.x * CH;
    const int cw = min(CH, a.C - ch0);                      // channels of this chunk (multiple of 4)
    constexpr int M0 = StagedTraits<OP>::m0, M1 = StagedTraits<OP>::m1;
    float* t1 = t0 + (size_t)M0 * a.max_cloud * CH;
    const bool has0 = a.in0 != nullptr, has1 = M1 > 0 && a.in1 != nullptr;
    // stage: rows [M*s, M*e) x channels [ch0, ch0+cw), float4 granularity, coalesced along channels
    const int q4 = cw >> 2;
    if (STAGE && has0)
        for (int t = threadIdx.x; t < M0 * P * q4; t += blockDim.x) {
            const int r = t / q4, c4 = t - r * q4;
            *reinterpret_cast<float4*>(t0 + (size_t)r * CH + 4 * c4) =
                __ldg(reinterpret_cast<const float4*>(a.in0 + ((size_t)M0 * s + r) * a.ld_in0 + ch0 + 4 * c4));
        }
    if (STAGE && has1)
        for (int t = threadIdx.x; t < M1 * P * q4; t += blockDim.x) {
            const int r = t / q4, c4 = t - r * q4;
            *reinterpret_cast<float4*>(t1 + (size_t)r * CH + 4 * c4) =
                __ldg(reinterpret_cast<const float4*>(a.in1 + ((size_t)M1 * s + r) * a.ld_in1 + ch0 + 4 * c4));
        }
    if (STAGE) __syncthreads();

    typedef Vec<4> V;
    const int lpp = 1 << lpp_log2, lane = threadIdx.x & 31, sub = lane & (lpp - 1);
    const int rows_per_warp = 32 >> lpp_log2, groups = (blockDim.x >> 5) * rows_per_warp;
    const int chl = sub * 4;                                // channel offset inside the chunk
    const bool ch_ok = chl < cw;
    const int ch = ch0 + chl;
    const float* g0 = a.in0 ? a.in0 + (size_t)M0 * s * a.ld_in0 + ch : nullptr;     // the cloud's first row, this lane's channels
    const float* g1 = (M1 > 0 && a.in1) ? a.in1 + (size_t)M1 * s * a.ld_in1 + ch : nullptr;
    auto ld0 = [&](int row) {
        V r;
        if (STAGE) r.v = *reinterpret_cast<const float4*>(t0 + (size_t)row * CH + chl);
        else r.v = __ldg(reinterpret_cast<const float4*>(g0 + (size_t)row * a.ld_in0));
        return r;
    };
    auto ld1 = [&](int row) {
        V r;
        if (STAGE) r.v = *reinterpret_cast<const float4*>(t1 + (size_t)row * CH + chl);
        else r.v = __ldg(reinterpret_cast<const float4*>(g1 + (size_t)row * a.ld_in1));
        return r;
    };

    for (int rbase = 0; rbase < P; rbase += groups) {
        const int rl = rbase + (threadIdx.x >> 5) * rows_per_warp + (lane >> lpp_log2);   // local row
        const bool row_ok = rl < P;
        const int r = s + (row_ok ? rl : 0);
        const float sc = (a.row_scale && row_ok) ? __ldg(a.row_scale + r) : 1.f;
        int t_begin, len;
        if (OpTraits<OP>::transposed) { t_begin = row_ok ? __ldg(a.in_ptr + r) : 0; len = row_ok ? __ldg(a.in_ptr + r + 1) - t_begin : 0; }
        else { t_begin = 0; len = row_ok ? a.k : 0; }
        int max_len = len;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) max_len = max(max_len, __shfl_xor_sync(FULL, max_len, o));
        V acc0 = V::zero(), acc1 = V::zero();
        for (int t0i = 0; t0i < max_len; t0i += lpp) {
            int my_other = 0;
            float2 my_c = make_float2(0.f, 0.f);
            const int t = t0i + sub;
            if (t < len) {
                if (OpTraits<OP>::transposed) {
                    int edge = __ldg(a.in_edge + t_begin + t);
                    my_other = edge / a.k - s;
                    my_c = __ldg(a.coef + edge);
                } else {
                    size_t eo = (size_t)r * a.k + t;
                    int j = __ldg(a.idx + eo);
                    my_c = __ldg(a.coef + eo);
                    if ((unsigned)j >= (unsigned)a.N) { j = s; my_c = make_float2(0.f, 0.f); }   // padding sentinel
                    my_other = j - s;
                }
            }
            const int cnt = min(lpp, max_len - t0i);
#pragma unroll 4
            for (int q = 0; q < cnt; ++q) {
                const int o = __shfl_sync(FULL, my_other, q, lpp);
                const float c0 = __shfl_sync(FULL, my_c.x, q, lpp);
                const float c1 = __shfl_sync(FULL, my_c.y, q, lpp);
                if (!ch_ok || t0i + q >= len) continue;
                if (OP == OP_GRAD) {
                    V sv = ld0(o);
                    acc0.fma(c0, sv); acc1.fma(c1, sv);
                } else if (OP == OP_DIVCURL) {
                    V vx = ld0(2 * o), vy = ld0(2 * o + 1);
                    acc0.fma(c0, vx); acc0.fma(c1, vy);
                    acc1.fma(c0, vy); acc1.fma(-c1, vx);
                } else if (OP == OP_HODGE) {
                    V va = ld0(o), vb = ld1(o);
                    acc0.fma(-c0, va); acc0.fma(c1, vb);
                    acc1.fma(-c1, va); acc1.fma(-c0, vb);
                } else if (OP == OP_GRAD_T) {
                    V px = ld0(2 * o), py = ld0(2 * o + 1);
                    acc0.fma(c0, px); acc0.fma(c1, py);
                } else if (OP == OP_DIVCURL_T) {
                    V dd = V::zero(), dc = V::zero();
                    if (has0) dd = ld0(o);
                    if (has1) dc = ld1(o);
                    acc0.fma(c0, dd); acc0.fma(-c1, dc);
                    acc1.fma(c1, dd); acc1.fma(c0, dc);
                } else {  // OP_HODGE_T
                    V px = ld0(2 * o), py = ld0(2 * o + 1);
                    acc0.fma(-c0, px); acc0.fma(-c1, py);
                    acc1.fma(c1, px); acc1.fma(-c0, py);
                }
            }
        }
        if (!row_ok || !ch_ok) continue;
        if (a.row_scale) {
            acc0 = V::map2(acc0, acc0, [sc](float p, float) { return p * sc; });
            acc1 = V::map2(acc1, acc1, [sc](float p, float) { return p * sc; });
        }
        if (OP == OP_DIVCURL && a.cp_src) V::load(a.cp_src + (size_t)r * a.ld_cp_src + ch).store(a.cp_dst + (size_t)r * a.ld_cp_dst + ch);
        if (OP == OP_HODGE && a.cp_src) {
            V::load(a.cp_src + (size_t)(2 * r) * a.ld_cp_src + ch).store(a.cp_dst + (size_t)(2 * r) * a.ld_cp_dst + ch);
            V::load(a.cp_src + (size_t)(2 * r + 1) * a.ld_cp_src + ch).store(a.cp_dst + (size_t)(2 * r + 1) * a.ld_cp_dst + ch);
        }
        if (OP == OP_GRAD || OP == OP_HODGE) {
            acc0.store(a.out0 + (size_t)(2 * r) * a.ld_out0 + ch);
            acc1.store(a.out0 + (size_t)(2 * r + 1) * a.ld_out0 + ch);
        } else if (OP == OP_DIVCURL) {
            if (a.out0) acc0.store(a.out0 + (size_t)r * a.ld_out0 + ch);
            if (a.out1) acc1.store(a.out1 + (size_t)r * a.ld_out1 + ch);
            if (a.out2) {
                V vx = ld0(2 * rl), vy = ld0(2 * rl + 1);
                V n = V::map2(vx, vy, [](float x, float y) { return sqrtf(x * x + y * y); });
                n.store(a.out2 + (size_t)r * a.ld_out2 + ch);
            }
        } else if (OP == OP_GRAD_T) {
            acc0.store(a.out0 + (size_t)r * a.ld_out0 + ch);
        } else if (OP == OP_DIVCURL_T) {
            if (a.in2) {   // d|v| : dv = dNorm * v / |v|   (0 where |v| = 0, torch's subgradient)
                V vx = V::load(a.in3 + (size_t)(2 * r) * a.ld_in3 + ch), vy = V::load(a.in3 + (size_t)(2 * r + 1) * a.ld_in3 + ch);
                V dn = V::load(a.in2 + (size_t)r * a.ld_in2 + ch);
                V inv = V::map2(vx, vy, [](float x, float y) { float n = sqrtf(x * x + y * y); return n > 0.f ? 1.f / n : 0.f; });
                V sx = V::map2(dn, inv, [](float d, float w) { return d * w; });
                acc0 = V::map2(acc0, V::map2(sx, vx, [](float s_, float x) { return s_ * x; }), [](float p, float q) { return p + q; });
                acc1 = V::map2(acc1, V::map2(sx, vy, [](float s_, float y) { return s_ * y; }), [](float p, float q) { return p + q; });
            }
            acc0.store(a.out0 + (size_t)(2 * r) * a.ld_out0 + ch);
            acc1.store(a.out0 + (size_t)(2 * r + 1) * a.ld_out0 + ch);
        } else {  // OP_HODGE_T
            acc0.store(a.out0 + (size_t)r * a.ld_out0 + ch);
            acc1.store(a.out1 + (size_t)r * a.ld_out1 + ch);
        }
    }
}

bool aligned16(const void* p) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int g_spmm_staged = 0;     // tuning switch (dc_set_spmm_staged); off: measured slower than the L2-gather kernel, see DESIGN.md

template <int OP>
int launch_spmm(const SpArgs& a, cudaStream_t st) {
    if (a.N <= 0 || a.C <= 0) return DC_OK;
    const bool vec4 = (a.C % 4 == 0) && a.ld_in0 % 4 == 0 && a.ld_in1 % 4 == 0 && a.ld_in2 % 4 == 0 && a.ld_in3 % 4 == 0 &&
                      a.ld_out0 % 4 == 0 && a.ld_out1 % 4 == 0 && a.ld_out2 % 4 == 0 && a.ld_cp_src % 4 == 0 && a.ld_cp_dst % 4 == 0 &&
                      aligned16(a.cp_src) && aligned16(a.cp_dst) && aligned16(a.in0) && aligned16(a.in1) &&
                      aligned16(a.in2) && aligned16(a.in3) && aligned16(a.out0) && aligned16(a.out1) && aligned16(a.out2);
    if (vec4 && a.cloud_ptr != nullptr && a.max_cloud > 0 && g_spmm_staged == 2) {
        // cloud-major through L1: a CTA = (cloud, channel chunk); the chunk is as wide as keeps ~one CTA per SM busy and the
        // cloud's rows of the chunk inside L1 (~200 KB)
        constexpr int M0 = StagedTraits<OP>::m0, M1 = StagedTraits<OP>::m1;
        const int rows = (M0 + ((M1 > 0 && a.in1) ? M1 : 0)) * a.max_cloud;
        int CH = 128;
        while (CH > 4 && (CH > a.C || (size_t)rows * CH * 4 > 200 * 1024)) CH >>= 1;
        while (CH > 8 && (long long)a.B * ((a.C + CH - 1) / CH) < dc_sm_count() * 3 / 4) CH >>= 1;
        int lpp_log2 = 0;
        while ((1 << lpp_log2) < CH / 4) ++lpp_log2;
        DC_CUDA(cudaFuncSetAttribute(spmm_staged_kernel<OP, false>, cudaFuncAttributePreferredSharedMemoryCarveout, 0));
        dim3 grid(dc_ceil_div(a.C, CH), a.B);
        spmm_staged_kernel<OP, false><<<grid, 1024, 0, st>>>(a, CH, lpp_log2);
        return dc_check_launch("spmm/cloud-major");
    }
    if (vec4 && a.cloud_ptr != nullptr && a.max_cloud > 0 && g_spmm_staged == 1) {
        // shared-memory staged variant: pick the channel chunk so that the staged rows fit (<= ~200 KB)
        constexpr int M0 = StagedTraits<OP>::m0, M1 = StagedTraits<OP>::m1;
        const int rows = (M0 + ((M1 > 0 && a.in1) ? M1 : 0)) * a.max_cloud;
        int CH = 32;
        while (CH > 4 && ((size_t)rows * CH * 4 > 100 * 1024 || CH > a.C)) CH >>= 1;
        // keep the machine full: at least ~1 CTA per SM
        while (CH > 8 && (long long)a.B * ((a.C + CH - 1) / CH) < dc_sm_count()) CH >>= 1;
        const size_t smem = (size_t)rows * CH * 4;
        if (smem <= 200 * 1024) {
            int lpp_log2 = 0;
            while ((1 << lpp_log2) < CH / 4) ++lpp_log2;
            DC_CUDA(cudaFuncSetAttribute(spmm_staged_kernel<OP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            dim3 grid(dc_ceil_div(a.C, CH), a.B);
            spmm_staged_kernel<OP, true><<<grid, 256, smem, st>>>(a, CH, lpp_log2);
            return dc_check_launch("spmm/staged");
        }
    }
    const int vec = vec4 ? 4 : 1;
    const int units = (a.C + vec - 1) / vec;            // lanes needed to cover the channels once
    int lpp_log2 = 0;
    while ((1 << lpp_log2) < units && lpp_log2 < 5) ++lpp_log2;
    const int lpp = 1 << lpp_log2;
    const int rows_per_cta = 8 * (32 >> lpp_log2);
    dim3 grid(dc_ceil_div(a.N, rows_per_cta), dc_ceil_div(units, lpp));
    if (a.row_scale) {
        if (vec4) spmm_kernel<OP, 4, true><<<grid, 256, 0, st>>>(a, lpp_log2);
        else spmm_kernel<OP, 1, true><<<grid, 256, 0, st>>>(a, lpp_log2);
    } else {
        if (vec4) spmm_kernel<OP, 4, false><<<grid, 256, 0, st>>>(a, lpp_log2);
        else spmm_kernel<OP, 1, false><<<grid, 256, 0, st>>>(a, lpp_log2);
    }
    return dc_check_launch("spmm");
}

// ---- K11: incoming-edge CSR of the kNN graph ---------------------------------------------------
__global__ void count_in_kernel(const int* __restrict__ idx, long long E, int N, int* __restrict__ cnt) {
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (t >= E) return;
    int j = idx[t];
    if ((unsigned)j < (unsigned)N) atomicAdd(cnt + j, 1);
}

constexpr int SCAN_T = 1024;
constexpr int SCAN_V = 4;            // elements per thread and round
// exclusive scan of cnt[0..n) into out[0..n], out[n] = total; single CTA, 4096 elements per round with a carry.
// n is at most a few million and the kernel runs once per batch, so one CTA is enough.
__global__ void __launch_bounds__(SCAN_T) scan_kernel(const int* __restrict__ cnt, int n, int* __restrict__ out) {
    __shared__ int warp_tot[32];
    __shared__ int carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < n; base += SCAN_T * SCAN_V) {
        const int i0 = base + threadIdx.x * SCAN_V;
        int v[SCAN_V], sum = 0;
#pragma unroll
        for (int u = 0; u < SCAN_V; ++u) { v[u] = (i0 + u < n) ? cnt[i0 + u] : 0; sum += v[u]; }
        int x = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(FULL, x, o); if (lane >= o) x += y; }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int w = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(FULL, w, o); if (lane >= o) w += y; }
            warp_tot[lane] = w;
        }
        __syncthreads();
        const int carry = carry_s;
        const int incl = x + (warp > 0 ? warp_tot[warp - 1] : 0);
        int run = carry + incl - sum;
#pragma unroll
        for (int u = 0; u < SCAN_V; ++u) { if (i0 + u < n) out[i0 + u] = run; run += v[u]; }
        __syncthreads();
        if (threadIdx.x == SCAN_T - 1) carry_s = carry + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) out[n] = carry_s;
}

__global__ void fill_in_kernel(const int* __restrict__ idx, long long E, int N, const int* __restrict__ in_ptr,
                               int* __restrict__ cursor, int* __restrict__ in_edge) {
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (t >= E) return;
    int j = idx[t];
    if ((unsigned)j >= (unsigned)N) return;
    int slot = atomicAdd(cursor + j, 1);
    in_edge[in_ptr[j] + slot] = (int)t;
}

// The atomic fill order is arbitrary; sorting each row makes the structure (and therefore every
// backward sum) bit-reproducible.  Rows are short (mean k): a CTA stages the entries of SORT_ROWS consecutive rows in
// shared memory (one coalesced pass each way), one thread insertion-sorts one row there; blocks of rows with more
// entries than the staging buffer (very uneven in-degrees) are sorted in place in global memory.
constexpr int SORT_ROWS = 128;
constexpr int SORT_CAP = 8192;       // staged entries per CTA (32 KB)
__global__ void __launch_bounds__(SORT_ROWS) sort_rows_kernel(const int* __restrict__ in_ptr, int N, int* __restrict__ in_edge) {
    __shared__ int buf[SORT_CAP];
    const int j0 = blockIdx.x * SORT_ROWS;
    const int j = j0 + threadIdx.x;
    const int lo = in_ptr[j0], hi = in_ptr[min(j0 + SORT_ROWS, N)];
    const bool staged = hi - lo <= SORT_CAP;
    if (staged) {
        for (int t = threadIdx.x; t < hi - lo; t += SORT_ROWS) buf[t] = in_edge[lo + t];
        __syncthreads();
    }
    if (j < N) {
        int* a = staged ? buf - lo : in_edge;
        const int b = in_ptr[j], e = in_ptr[j + 1];
        for (int p = b + 1; p < e; ++p) {
            int key = a[p], q = p - 1;
            while (q >= b && a[q] > key) { a[q + 1] = a[q]; --q; }
            a[q + 1] = key;
        }
    }
    if (staged) {
        __syncthreads();
        for (int t = threadIdx.x; t < hi - lo; t += SORT_ROWS) in_edge[lo + t] = buf[t];
    }
}

}  // namespace

#define SP_COMMON(a_, N_, k_, C_) a_.N = N_; a_.k = k_; a_.C = C_

extern "C" size_t dc_ell_transpose_ws_bytes(int N) { return ((size_t)(N + 1) * 4 + 255) / 256 * 256 * 2; }

extern "C" int dc_ell_transpose(const int32_t* idx, int N, int k, int32_t* in_ptr, int32_t* in_edge,
                                void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(k >= 1 && N >= 0, "dc_ell_transpose: bad sizes");
    DC_REQUIRE((long long)N * k < 2147483647LL, "dc_ell_transpose: N*k must fit in int32");
    DC_REQUIRE(ws_bytes >= dc_ell_transpose_ws_bytes(N), "dc_ell_transpose: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    if (N == 0) return DC_OK;
    const long long E = (long long)N * k;
    int* cnt = (int*)ws;
    int* cursor = (int*)((char*)ws + dc_ell_transpose_ws_bytes(N) / 2);
    DC_CUDA(cudaMemsetAsync(ws, 0, dc_ell_transpose_ws_bytes(N), st));
    count_in_kernel<<<dc_ceil_div(E, 256), 256, 0, st>>>(idx, E, N, cnt);
    scan_kernel<<<1, SCAN_T, 0, st>>>(cnt, N, in_ptr);
    fill_in_kernel<<<dc_ceil_div(E, 256), 256, 0, st>>>(idx, E, N, in_ptr, cursor, in_edge);
    sort_rows_kernel<<<dc_ceil_div(N, SORT_ROWS), SORT_ROWS, 0, st>>>(in_ptr, N, in_edge);
    return dc_check_launch_n("dc_ell_transpose", 4);
}

extern "C" int dc_grad_fwd(const int32_t* idx, const float* G, int N, int k, int C, const float* x, int64_t ldx,
                           float* out, int64_t ldo, const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream) {
    SpArgs a = {};
    SP_COMMON(a, N, k, C);
    a.idx = idx; a.coef = (const float2*)G; a.in0 = x; a.ld_in0 = ldx; a.out0 = out; a.ld_out0 = ldo;
    a.row_scale = row_scale_or_null; a.cloud_ptr = cloud_ptr_or_null; a.B = B; a.max_cloud = max_cloud_points;
    return launch_spmm<OP_GRAD>(a, (cudaStream_t)stream);
}

extern "C" int dc_grad_bwd(const int32_t* in_ptr, const int32_t* in_edge, const float* G, int N, int k, int C,
                           const float* dV, int64_t ld_dv, float* dX, int64_t ld_dx, const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream) {
    SpArgs a = {};
    SP_COMMON(a, N, k, C);
    a.in_ptr = in_ptr; a.in_edge = in_edge; a.coef = (const float2*)G; a.in0 = dV; a.ld_in0 = ld_dv; a.out0 = dX; a.ld_out0 = ld_dx;
    a.row_scale = row_scale_or_null; a.cloud_ptr = cloud_ptr_or_null; a.B = B; a.max_cloud = max_cloud_points;
    return launch_spmm<OP_GRAD_T>(a, (cudaStream_t)stream);
}

extern "C" int dc_divcurlnorm_fwd(const int32_t* idx, const float* D, int N, int k, int C, const float* v, int64_t ldv,
                                  float* div_or_null, int64_t ld_div, float* curl_or_null, int64_t ld_curl,
                                  float* norm_or_null, int64_t ld_norm, const float* x_or_null, int64_t ldx, float* x_copy, int64_t ld_copy,
                                  const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream) {
    SpArgs a = {};
    SP_COMMON(a, N, k, C);
    DC_REQUIRE(x_or_null == nullptr || x_copy != nullptr, "dc_divcurlnorm_fwd: x given without a destination");
    a.cp_src = x_or_null; a.ld_cp_src = ldx; a.cp_dst = x_copy; a.ld_cp_dst = ld_copy;
    a.idx = idx; a.coef = (const float2*)D; a.in0 = v; a.ld_in0 = ldv;
    a.out0 = div_or_null; a.ld_out0 = ld_div; a.out1 = curl_or_null; a.ld_out1 = ld_curl; a.out2 = norm_or_null; a.ld_out2 = ld_norm;
    a.row_scale = row_scale_or_null; a.cloud_ptr = cloud_ptr_or_null; a.B = B; a.max_cloud = max_cloud_points;
    return launch_spmm<OP_DIVCURL>(a, (cudaStream_t)stream);
}

extern "C" int dc_divcurlnorm_bwd(const int32_t* in_ptr, const int32_t* in_edge, const float* D, int N, int k, int C,
                                  const float* d_div_or_null, int64_t ld_ddiv, const float* d_curl_or_null, int64_t ld_dcurl,
                                  const float* d_norm_or_null, int64_t ld_dnorm, const float* v, int64_t ldv,
                                  float* dV, int64_t ld_dv, const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream) {
    SpArgs a = {};
    SP_COMMON(a, N, k, C);
    a.in_ptr = in_ptr; a.in_edge = in_edge; a.coef = (const float2*)D;
    a.in0 = d_div_or_null; a.ld_in0 = ld_ddiv; a.in1 = d_curl_or_null; a.ld_in1 = ld_dcurl;
    a.in2 = d_norm_or_null; a.ld_in2 = ld_dnorm; a.in3 = v; a.ld_in3 = ldv;
    a.out0 = dV; a.ld_out0 = ld_dv;
    a.row_scale = row_scale_or_null; a.cloud_ptr = cloud_ptr_or_null; a.B = B; a.max_cloud = max_cloud_points;
    return launch_spmm<OP_DIVCURL_T>(a, (cudaStream_t)stream);
}

extern "C" int dc_hodge_fwd(const int32_t* idx, const float* G, int N, int k, int C, const float* div_v, int64_t ld_a,
                            const float* curl_v, int64_t ld_b, float* out, int64_t ldo, const float* v_or_null, int64_t ldv,
                            float* v_copy, int64_t ld_copy, const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream) {
    SpArgs a = {};
    SP_COMMON(a, N, k, C);
    DC_REQUIRE(v_or_null == nullptr || v_copy != nullptr, "dc_hodge_fwd: v given without a destination");
    a.cp_src = v_or_null; a.ld_cp_src = ldv; a.cp_dst = v_copy; a.ld_cp_dst = ld_copy;
    a.idx = idx; a.coef = (const float2*)G; a.in0 = div_v; a.ld_in0 = ld_a; a.in1 = curl_v; a.ld_in1 = ld_b; a.out0 = out; a.ld_out0 = ldo;
    a.row_scale = row_scale_or_null; a.cloud_ptr = cloud_ptr_or_null; a.B = B; a.max_cloud = max_cloud_points;
    return launch_spmm<OP_HODGE>(a, (cudaStream_t)stream);
}

extern "C" int dc_hodge_bwd(const int32_t* in_ptr, const int32_t* in_edge, const float* G, int N, int k, int C,
                            const float* d_out, int64_t ld_do, float* d_div, int64_t ld_da, float* d_curl, int64_t ld_db,
                            const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream) {
    SpArgs a = {};
    SP_COMMON(a, N, k, C);
    a.in_ptr = in_ptr; a.in_edge = in_edge; a.coef = (const float2*)G; a.in0 = d_out; a.ld_in0 = ld_do;
    a.out0 = d_div; a.ld_out0 = ld_da; a.out1 = d_curl; a.ld_out1 = ld_db;
    a.row_scale = row_scale_or_null; a.cloud_ptr = cloud_ptr_or_null; a.B = B; a.max_cloud = max_cloud_points;
    return launch_spmm<OP_HODGE_T>(a, (cudaStream_t)stream);
}

// tuning switch: 0 (default) = row-major grid gathering from L2, 1 = a cloud's rows staged in shared memory, 2 = cloud-major grid
// gathering through L1 (both need the cloud layout)
extern "C" int dc_set_spmm_staged(int v) { g_spmm_staged = v; return DC_OK; }
