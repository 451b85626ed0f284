// K7/K8/K10 and the BatchNorm pieces — everything in DeltaConv / MLP / VectorMLP that is not an
// SpMM or a GEMM.  Reference code replaced (relative to /root/reference):
//   deltaconv/nn/deltaconv.py:50-54   torch_scatter.scatter(reduce='max') over the kNN graph
//   deltaconv/nn/nonlin.py:11-35      BatchNorm1d wrapper (+ mlp.py:7-11 LeakyReLU)
//   deltaconv/nn/nonlin.py:63-82      VectorNonLin
//   deltaconv/models/deltanet_classification.py:46-47, deltanet_segmentation.py:61  global pools
//
// BatchNorm in training mode needs a column reduction over all rows between the GEMM and the
// activation.  It is done as: per-CTA fp64 partial sums (one CTA per row chunk, fixed grid) ->
// one finalize kernel that adds the partials in a fixed order.  No atomics, bit-reproducible.
#include "common.cuh"
#include "colreduce.cuh"

namespace {

// Grid-stride iteration over an [M, W] index space: (row, column) advance by the constant (stride / W, stride % W) instead
// of a 64-bit division per element (which cost more instructions than the element's arithmetic).
struct RowCol {
    long long r, dr; int c, dc, W;
    __device__ __forceinline__ explicit RowCol(int W_) : W(W_) {
        const long long start = blockIdx.x * (long long)blockDim.x + threadIdx.x, stride = (long long)gridDim.x * blockDim.x;
        r = start / W; c = (int)(start - r * W);
        dr = stride / W; dc = (int)(stride - dr * W);
    }
    __device__ __forceinline__ void next() { r += dr; c += dc; if (c >= W) { c -= W; ++r; } }
};

struct SumVal {        // plain column sums of a [M,C] matrix
    const float* x; long long ld;
    __device__ void operator()(long long r, int c, float (&o)[1]) const { o[0] = __ldg(x + r * ld + c); }
};
__global__ void sum1_finalize_kernel(const double* __restrict__ partial, int chunks, int C, float* __restrict__ s0) {
    double r[1];
    if (!finalize_sums<1>(partial, chunks, C, r)) return;
    s0[finalize_column()] = (float)r[0];
}

// ---- BatchNorm statistics ------------------------------------------------------------------
// Sums are taken around a per-column pivot (the column's value in row 0): s = sum(x - p), s2 = sum((x - p)^2), so that
// mean = p + s / M and var = s2 / M - (s / M)^2 do not cancel when |mean| >> std (squares of the raw values, rounded to fp32
// before the fp64 accumulation, would lose the variance's digits there).
struct Pivot4 { float p[4]; };
struct PlainVal {      // (x - p) and (x - p)^2 of a [M,C] matrix
    const float* x; long long ld;
    __device__ float pivot(int c) const { return __ldg(x + c); }
    __device__ void operator()(long long r, int c, float (&o)[2]) const { float v = __ldg(x + r * ld + c) - pivot(c); o[0] = v; o[1] = v * v; }
    __device__ Pivot4 columns(int c) const {
        const float4 q = __ldg((const float4*)(x + c));
        Pivot4 k; k.p[0] = q.x; k.p[1] = q.y; k.p[2] = q.z; k.p[3] = q.w; return k;
    }
    __device__ void quad(long long r, int c, const Pivot4& k, float (&o)[4][2]) const {
        const float4 q = __ldg((const float4*)(x + r * ld + c));
        const float d[4] = {q.x - k.p[0], q.y - k.p[1], q.z - k.p[2], q.w - k.p[3]};
#pragma unroll
        for (int u = 0; u < 4; ++u) { o[u][0] = d[u]; o[u][1] = d[u] * d[u]; }
    }
};
struct NormVal {       // the same for |v| of a [2M,C] vector field (nonlin.py:68)
    const float* x; long long ld;
    __device__ float pivot(int c) const { float a = __ldg(x + c), b = __ldg(x + ld + c); return sqrtf(a * a + b * b); }
    __device__ void operator()(long long r, int c, float (&o)[2]) const {
        float a = __ldg(x + (2 * r) * ld + c), b = __ldg(x + (2 * r + 1) * ld + c);
        float d = sqrtf(a * a + b * b) - pivot(c);
        o[0] = d; o[1] = d * d;
    }
    __device__ Pivot4 columns(int c) const {
        Pivot4 k;
#pragma unroll
        for (int u = 0; u < 4; ++u) k.p[u] = pivot(c + u);
        return k;
    }
    __device__ void quad(long long r, int c, const Pivot4& k, float (&o)[4][2]) const {
        const float4 a = __ldg((const float4*)(x + (2 * r) * ld + c)), b = __ldg((const float4*)(x + (2 * r + 1) * ld + c));
        const float aa[4] = {a.x, a.y, a.z, a.w}, bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) { const float d = sqrtf(aa[u] * aa[u] + bb[u] * bb[u]) - k.p[u]; o[u][0] = d; o[u][1] = d * d; }
    }
};

// mean / biased variance -> (mean, invstd); running stats updated like torch.nn.BatchNorm1d
// (momentum, unbiased variance).  One thread per column, partials added in chunk order.
template <class F>
__global__ void bn_finalize_kernel(F f, const double* __restrict__ partial, int chunks, long long M, int C, float eps,
                                   float momentum, float* __restrict__ mean, float* __restrict__ invstd,
                                   float* __restrict__ running_mean, float* __restrict__ running_var,
                                   long long* __restrict__ num_batches_tracked) {
    if (num_batches_tracked && blockIdx.x == 0 && threadIdx.x == 0) *num_batches_tracked += 1;
    double r[2];
    if (!finalize_sums<2>(partial, chunks, C, r)) return;
    const int c = finalize_column();
    double s = r[0], s2 = r[1];
    double shifted = s / (double)M;                       // mean of (x - pivot)
    double var = s2 / (double)M - shifted * shifted;
    if (var < 0.0) var = 0.0;
    double mu = (double)f.pivot(c) + shifted;
    mean[c] = (float)mu;
    invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
    if (running_mean) {
        double unbiased = M > 1 ? var * (double)M / (double)(M - 1) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mu;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
}

// ---- fused BN + LeakyReLU (+ residual addend), forward -------------------------------------
// out = act(gamma * (y - mean) * invstd + beta) + addend        act(t) = t > 0 ? t : slope * t
__global__ void bn_act_fwd_kernel(const float* __restrict__ y, long long ldy, long long M, int C,
                                  const float* __restrict__ mean, const float* __restrict__ invstd,
                                  const float* __restrict__ gamma, const float* __restrict__ beta, float slope,
                                  const float* __restrict__ addend, long long ld_add, float* __restrict__ out, long long ldo) {
    const long long total = M * (long long)C;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        long long r = t / C; int c = (int)(t - r * C);
        float g = gamma ? __ldg(gamma + c) : 1.f, b = beta ? __ldg(beta + c) : 0.f;
        float z = g * ((__ldg(y + r * ldy + c) - __ldg(mean + c)) * __ldg(invstd + c)) + b;
        z = z > 0.f ? z : slope * z;
        if (addend) z += __ldg(addend + r * ld_add + c);
        out[r * ldo + c] = z;
    }
}

// float4 variant (C % 4 == 0, every leading dimension % 4 == 0, 16-byte aligned bases)
__global__ void bn_act_fwd_vec4_kernel(const float* __restrict__ y, long long ldy, long long M, int C4,
                                       const float* __restrict__ mean, const float* __restrict__ invstd,
                                       const float* __restrict__ gamma, const float* __restrict__ beta, float slope,
                                       const float* __restrict__ addend, long long ld_add, float* __restrict__ out, long long ldo) {
    const long long total = M * (long long)C4;
    for (RowCol rc(C4); rc.r < M; rc.next()) {            // grid-stride over [M, C4] without a 64-bit division per element
        const long long r = rc.r;
        const int c = rc.c * 4;
        const float4 v = ld_stream4(reinterpret_cast<const float4*>(y + r * ldy + c));
        const float4 mu = __ldg(reinterpret_cast<const float4*>(mean + c)), is = __ldg(reinterpret_cast<const float4*>(invstd + c));
        const float4 g = gamma ? __ldg(reinterpret_cast<const float4*>(gamma + c)) : make_float4(1.f, 1.f, 1.f, 1.f);
        const float4 b = beta ? __ldg(reinterpret_cast<const float4*>(beta + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 z;
        z.x = g.x * ((v.x - mu.x) * is.x) + b.x; z.y = g.y * ((v.y - mu.y) * is.y) + b.y;
        z.z = g.z * ((v.z - mu.z) * is.z) + b.z; z.w = g.w * ((v.w - mu.w) * is.w) + b.w;
        z.x = z.x > 0.f ? z.x : slope * z.x; z.y = z.y > 0.f ? z.y : slope * z.y;
        z.z = z.z > 0.f ? z.z : slope * z.z; z.w = z.w > 0.f ? z.w : slope * z.w;
        if (addend) {
            const float4 a = ld_stream4(reinterpret_cast<const float4*>(addend + r * ld_add + c));
            z.x += a.x; z.y += a.y; z.z += a.z; z.w += a.w;
        }
        *reinterpret_cast<float4*>(out + r * ldo + c) = z;
    }
}

// backward, reduction part: dz = dOut * act'(bn); sums of dz and dz * xhat per column
// mean / invstd / gamma / beta of columns c..c+3 (BatchNorm backward reductions)
__device__ __forceinline__ BnColumns bn_columns(const float* mean, const float* invstd, const float* gamma, const float* beta, int c) {
    BnColumns k;
    const float4 mu = __ldg((const float4*)(mean + c)), is = __ldg((const float4*)(invstd + c));
    const float4 g = gamma ? __ldg((const float4*)(gamma + c)) : make_float4(1.f, 1.f, 1.f, 1.f);
    const float4 b = beta ? __ldg((const float4*)(beta + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
    k.mean[0] = mu.x; k.mean[1] = mu.y; k.mean[2] = mu.z; k.mean[3] = mu.w;
    k.invstd[0] = is.x; k.invstd[1] = is.y; k.invstd[2] = is.z; k.invstd[3] = is.w;
    k.gamma[0] = g.x; k.gamma[1] = g.y; k.gamma[2] = g.z; k.gamma[3] = g.w;
    k.beta[0] = b.x; k.beta[1] = b.y; k.beta[2] = b.z; k.beta[3] = b.w;
    return k;
}

struct BnBwdVal {
    const float* y; long long ldy; const float* dout; long long ldd;
    const float* mean; const float* invstd; const float* gamma; const float* beta; float slope;
    __device__ void operator()(long long r, int c, float (&o)[2]) const {
        float xh = (__ldg(y + r * ldy + c) - __ldg(mean + c)) * __ldg(invstd + c);
        float g = gamma ? __ldg(gamma + c) : 1.f, b = beta ? __ldg(beta + c) : 0.f;
        float z = g * xh + b;
        float dz = __ldg(dout + r * ldd + c) * (z > 0.f ? 1.f : slope);
        o[0] = dz; o[1] = dz * xh;
    }
    __device__ BnColumns columns(int c) const { return bn_columns(mean, invstd, gamma, beta, c); }
    __device__ void quad(long long r, int c, const BnColumns& k, float (&o)[4][2]) const {
        const float4 yv = __ldg((const float4*)(y + r * ldy + c)), dv = __ldg((const float4*)(dout + r * ldd + c));
        const float yy[4] = {yv.x, yv.y, yv.z, yv.w}, dd[4] = {dv.x, dv.y, dv.z, dv.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float xh = (yy[u] - k.mean[u]) * k.invstd[u];
            const float z = k.gamma[u] * xh + k.beta[u];
            const float dz = dd[u] * (z > 0.f ? 1.f : slope);
            o[u][0] = dz; o[u][1] = dz * xh;
        }
    }
};

__global__ void sum2_finalize_kernel(const double* __restrict__ partial, int chunks, int C, float* __restrict__ s0, float* __restrict__ s1) {
    double r[2];
    if (!finalize_sums<2>(partial, chunks, C, r)) return;
    const int c = finalize_column();
    s0[c] = (float)r[0]; s1[c] = (float)r[1];
}

// backward, elementwise part: dy = gamma * invstd * (dz - sum_dz/M - xhat * sum_dz_xhat/M)   (training)
//                              dy = gamma * invstd * dz                                     (eval)
__global__ void bn_act_bwd_kernel(const float* __restrict__ y, long long ldy, const float* __restrict__ dout, long long ldd,
                                  long long M, int C, const float* __restrict__ mean, const float* __restrict__ invstd,
                                  const float* __restrict__ gamma, const float* __restrict__ beta, float slope,
                                  const float* __restrict__ sum_dz, const float* __restrict__ sum_dz_xh, int training,
                                  float* __restrict__ dy, long long ld_dy) {
    const long long total = M * (long long)C;
    const float invM = 1.f / (float)M;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        long long r = t / C; int c = (int)(t - r * C);
        float is = __ldg(invstd + c);
        float xh = (__ldg(y + r * ldy + c) - __ldg(mean + c)) * is;
        float g = gamma ? __ldg(gamma + c) : 1.f, b = beta ? __ldg(beta + c) : 0.f;
        float z = g * xh + b;
        float dz = __ldg(dout + r * ldd + c) * (z > 0.f ? 1.f : slope);
        float v = training ? dz - __ldg(sum_dz + c) * invM - xh * (__ldg(sum_dz_xh + c) * invM) : dz;
        dy[r * ld_dy + c] = g * is * v;
    }
}

__global__ void bn_act_bwd_vec4_kernel(const float* __restrict__ y, long long ldy, const float* __restrict__ dout, long long ldd,
                                       long long M, int C4, const float* __restrict__ mean, const float* __restrict__ invstd,
                                       const float* __restrict__ gamma, const float* __restrict__ beta, float slope,
                                       const float* __restrict__ sum_dz, const float* __restrict__ sum_dz_xh, int training,
                                       float* __restrict__ dy, long long ld_dy) {
    const long long total = M * (long long)C4;
    const float invM = 1.f / (float)M;
    for (RowCol rc(C4); rc.r < M; rc.next()) {            // grid-stride over [M, C4] without a 64-bit division per element
        const long long r = rc.r;
        const int c = rc.c * 4;
        const float4 yv = ld_stream4(reinterpret_cast<const float4*>(y + r * ldy + c));
        const float4 dv = ld_stream4(reinterpret_cast<const float4*>(dout + r * ldd + c));
        const float4 mu = __ldg(reinterpret_cast<const float4*>(mean + c)), is = __ldg(reinterpret_cast<const float4*>(invstd + c));
        const float4 g = gamma ? __ldg(reinterpret_cast<const float4*>(gamma + c)) : make_float4(1.f, 1.f, 1.f, 1.f);
        const float4 b = beta ? __ldg(reinterpret_cast<const float4*>(beta + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 s1 = __ldg(reinterpret_cast<const float4*>(sum_dz + c)), s2 = __ldg(reinterpret_cast<const float4*>(sum_dz_xh + c));
        auto one = [&](float yy, float dd, float m, float i_, float gg, float bb, float a1, float a2) {
            float xh = (yy - m) * i_;
            float z = gg * xh + bb;
            float dz = dd * (z > 0.f ? 1.f : slope);
            float v = training ? dz - a1 * invM - xh * (a2 * invM) : dz;
            return gg * i_ * v;
        };
        float4 o;
        o.x = one(yv.x, dv.x, mu.x, is.x, g.x, b.x, s1.x, s2.x); o.y = one(yv.y, dv.y, mu.y, is.y, g.y, b.y, s1.y, s2.y);
        o.z = one(yv.z, dv.z, mu.z, is.z, g.z, b.z, s1.z, s2.z); o.w = one(yv.w, dv.w, mu.w, is.w, g.w, b.w, s1.w, s2.w);
        *reinterpret_cast<float4*>(dy + r * ld_dy + c) = o;
    }
}

// out[r, :] = a[r, :] + b[r, :] on column slices of wider buffers (gradient of a tensor that feeds two consumers)
__global__ void add_rows_kernel(const float* __restrict__ a, long long lda, const float* __restrict__ b, long long ldb,
                                long long M, int C, int vec, float* __restrict__ out, long long ldo) {
    const int cw = vec ? C / 4 : C;
    const long long total = M * (long long)cw;
    for (RowCol rc(cw); rc.r < M; rc.next()) {
        const long long r = rc.r;
        const int c = rc.c;
        if (vec) {
            const float4 x = ld_stream4(reinterpret_cast<const float4*>(a + r * lda) + c), y = ld_stream4(reinterpret_cast<const float4*>(b + r * ldb) + c);
            reinterpret_cast<float4*>(out + r * ldo)[c] = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
        } else {
            out[r * ldo + c] = __ldg(a + r * lda + c) + __ldg(b + r * ldb + c);
        }
    }
}

// ---- SGD with momentum and weight decay on flat buffers (torch.optim.SGD, dampening 0, no Nesterov) ----------------
// g' = gs * g + wd * p;  buf = mu * buf + g';  p -= lr * buf.   A zero-initialised buf makes the first step buf = g' as in torch.
// lr is read from device memory so that a captured step follows learning-rate schedules.  gs (grad_scale) is 1 / world size
// after a SUM all-reduce: the averaging costs no pass of its own.
__global__ void sgd_step_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ buf, long long n,
                                const float* __restrict__ lr_ptr, float momentum, float weight_decay, float gs) {
    const float lr = __ldg(lr_ptr);
    const long long n4 = n >> 2;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n4; t += (long long)gridDim.x * blockDim.x) {
        float4 pv = reinterpret_cast<float4*>(p)[t], bv = reinterpret_cast<float4*>(buf)[t];
        const float4 gv = ld_stream4(reinterpret_cast<const float4*>(g) + t);
        bv.x = fmaf(momentum, bv.x, fmaf(weight_decay, pv.x, gs * gv.x)); pv.x = fmaf(-lr, bv.x, pv.x);
        bv.y = fmaf(momentum, bv.y, fmaf(weight_decay, pv.y, gs * gv.y)); pv.y = fmaf(-lr, bv.y, pv.y);
        bv.z = fmaf(momentum, bv.z, fmaf(weight_decay, pv.z, gs * gv.z)); pv.z = fmaf(-lr, bv.z, pv.z);
        bv.w = fmaf(momentum, bv.w, fmaf(weight_decay, pv.w, gs * gv.w)); pv.w = fmaf(-lr, bv.w, pv.w);
        reinterpret_cast<float4*>(buf)[t] = bv;
        reinterpret_cast<float4*>(p)[t] = pv;
    }
    for (long long t = (n4 << 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const float b = fmaf(momentum, buf[t], fmaf(weight_decay, p[t], gs * g[t]));
        buf[t] = b;
        p[t] = fmaf(-lr, b, p[t]);
    }
}

// ---- Adam on flat buffers (torch.optim.Adam: L2 weight decay added to the gradient, no amsgrad) -----------------
// t = *step + 1;  m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;  p -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps).
// The step counter and the learning rate live in device memory (captured steps); adam_advance_kernel increments the counter.
__global__ void adam_step_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                                 long long n, const float* __restrict__ lr_ptr, const int* __restrict__ step_ptr,
                                 float beta1, float beta2, float eps, float weight_decay, float gs) {
    const float t = (float)(__ldg(step_ptr) + 1);
    const float bc1 = 1.f - powf(beta1, t), bc2_sqrt = sqrtf(1.f - powf(beta2, t));
    const float step_size = __ldg(lr_ptr) / bc1;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float pi = p[i];
        const float gi = fmaf(weight_decay, pi, gs * g[i]);
        const float mi = fmaf(beta1, m[i], (1.f - beta1) * gi);
        const float vi = fmaf(beta2, v[i], (1.f - beta2) * gi * gi);
        m[i] = mi; v[i] = vi;
        p[i] = pi - step_size * (mi / (sqrtf(vi) / bc2_sqrt + eps));
    }
}
__global__ void adam_advance_kernel(int* step_ptr) { *step_ptr += 1; }

// ---- BatchNorm + activation fused with the global pools, backward -------------------------------
// The gradient of [max | mean] pooling w.r.t. the pooled activation at (r, c) is (r == arg[b,c] ? d_max[b,c] : 0) + d_mean[b,c] / n_b
// (same expression as pool_bwd): it is rebuilt on the fly.  Work is split by cloud so that the [B,C] quantities are loaded once per
// thread: a unit = (cloud, split of its rows); a CTA = 32 float4 column lanes x 8 row lanes.
struct PoolGrad {
    const int* cloud_ptr; const int* arg;
    const float* d_max; long long ld_dmax; const float* d_mean; long long ld_dmean; int C;
};

struct PoolGradCol {      // the per-(cloud, 4 columns) part, loaded once
    int4 am; float4 gm, ga;
    __device__ __forceinline__ void load(const PoolGrad& pg, int b, int c, float inv) {
        am = make_int4(-1, -1, -1, -1); gm = make_float4(0.f, 0.f, 0.f, 0.f); ga = gm;
        if (pg.d_max) {
            am = __ldg(reinterpret_cast<const int4*>(pg.arg + (size_t)b * pg.C + c));
            gm = __ldg(reinterpret_cast<const float4*>(pg.d_max + (size_t)b * pg.ld_dmax + c));
        }
        if (pg.d_mean) {
            ga = __ldg(reinterpret_cast<const float4*>(pg.d_mean + (size_t)b * pg.ld_dmean + c));
            ga.x *= inv; ga.y *= inv; ga.z *= inv; ga.w *= inv;
        }
    }
    __device__ __forceinline__ float4 at(int r) const {
        return make_float4((am.x == r ? gm.x : 0.f) + ga.x, (am.y == r ? gm.y : 0.f) + ga.y,
                           (am.z == r ? gm.z : 0.f) + ga.z, (am.w == r ? gm.w : 0.f) + ga.w);
    }
};

struct BnCol {            // BatchNorm constants of 4 columns
    float4 mu, is, g, bt;
    __device__ __forceinline__ void load(const float* mean, const float* invstd, const float* gamma, const float* beta, int c) {
        mu = __ldg(reinterpret_cast<const float4*>(mean + c)); is = __ldg(reinterpret_cast<const float4*>(invstd + c));
        g = gamma ? __ldg(reinterpret_cast<const float4*>(gamma + c)) : make_float4(1.f, 1.f, 1.f, 1.f);
        bt = beta ? __ldg(reinterpret_cast<const float4*>(beta + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
};

// pass 1: per-CTA fp64 partial sums of dz and dz * xhat; partial[(q * 2 + v) * C + c], q = blockIdx.x
__global__ void __launch_bounds__(256) pool_bn_bwd_reduce_kernel(const float* __restrict__ y, long long ldy, PoolGrad pg, int B, int splits, int C,
                                                                 const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                 const float* __restrict__ gamma, const float* __restrict__ beta, float slope,
                                                                 double* __restrict__ partial) {
    __shared__ double sh[8][2][128];
    const int lane = threadIdx.x & 31, rl = threadIdx.x >> 5;
    const int c = (blockIdx.y * 32 + lane) * 4;
    double a0[4] = {0.0, 0.0, 0.0, 0.0}, a1[4] = {0.0, 0.0, 0.0, 0.0};
    if (c < C) {
        BnCol k; k.load(mean, invstd, gamma, beta, c);
        for (int u = blockIdx.x; u < B * splits; u += gridDim.x) {          // units in a fixed order: deterministic
            const int b = u / splits, sp = u - b * splits;
            const int s = __ldg(pg.cloud_ptr + b), e = __ldg(pg.cloud_ptr + b + 1);
            const int per = (e - s + splits - 1) / splits;
            const int r0 = s + sp * per, r1 = min(e, r0 + per);
            PoolGradCol col; col.load(pg, b, c, 1.f / (float)max(e - s, 1));
#pragma unroll 4
            for (int r = r0 + rl; r < r1; r += 8) {
                const float4 yv = ld_stream4(reinterpret_cast<const float4*>(y + (size_t)r * ldy + c));
                const float4 d = col.at(r);
                auto one = [&](float yy, float dd, float m, float i_, float gg, float bb, double& s0, double& s1) {
                    float xh = (yy - m) * i_;
                    float z = gg * xh + bb;
                    float dz = dd * (z > 0.f ? 1.f : slope);
                    s0 += (double)dz; s1 += (double)(dz * xh);
                };
                one(yv.x, d.x, k.mu.x, k.is.x, k.g.x, k.bt.x, a0[0], a1[0]); one(yv.y, d.y, k.mu.y, k.is.y, k.g.y, k.bt.y, a0[1], a1[1]);
                one(yv.z, d.z, k.mu.z, k.is.z, k.g.z, k.bt.z, a0[2], a1[2]); one(yv.w, d.w, k.mu.w, k.is.w, k.g.w, k.bt.w, a0[3], a1[3]);
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) { sh[rl][0][lane * 4 + q] = a0[q]; sh[rl][1][lane * 4 + q] = a1[q]; }
    __syncthreads();
    const int v = threadIdx.x >> 7, cc = threadIdx.x & 127;                 // 256 threads = 2 values x 128 columns
    const int col = blockIdx.y * 128 + cc;
    if (col < C) {
        double t = 0.0;
#pragma unroll
        for (int q = 0; q < 8; ++q) t += sh[q][v][cc];
        partial[((size_t)blockIdx.x * 2 + v) * C + col] = t;
    }
}

// pass 2: dy = gamma * invstd * (dz - sum_dz / M - xhat * sum_dz_xhat / M), grid (splits, column blocks, B)
__global__ void __launch_bounds__(256) pool_bn_bwd_apply_kernel(const float* __restrict__ y, long long ldy, PoolGrad pg, long long M, int C,
                                                                const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                const float* __restrict__ gamma, const float* __restrict__ beta, float slope,
                                                                const float* __restrict__ sum_dz, const float* __restrict__ sum_dz_xh,
                                                                float* __restrict__ dy, long long ld_dy) {
    const int lane = threadIdx.x & 31, rl = threadIdx.x >> 5;
    const int c = (blockIdx.y * 32 + lane) * 4;
    if (c >= C) return;
    const int b = blockIdx.z;
    const int s = __ldg(pg.cloud_ptr + b), e = __ldg(pg.cloud_ptr + b + 1);
    const float invM = 1.f / (float)M;
    BnCol k; k.load(mean, invstd, gamma, beta, c);
    PoolGradCol col; col.load(pg, b, c, 1.f / (float)max(e - s, 1));
    float4 s1 = __ldg(reinterpret_cast<const float4*>(sum_dz + c)), s2 = __ldg(reinterpret_cast<const float4*>(sum_dz_xh + c));
    s1.x *= invM; s1.y *= invM; s1.z *= invM; s1.w *= invM;
    s2.x *= invM; s2.y *= invM; s2.z *= invM; s2.w *= invM;
#pragma unroll 4
    for (int r = s + blockIdx.x * 8 + rl; r < e; r += gridDim.x * 8) {
        const float4 yv = ld_stream4(reinterpret_cast<const float4*>(y + (size_t)r * ldy + c));
        const float4 d = col.at(r);
        auto one = [&](float yy, float dd, float m, float i_, float gg, float bb, float a1, float a2) {
            float xh = (yy - m) * i_;
            float z = gg * xh + bb;
            float dz = dd * (z > 0.f ? 1.f : slope);
            return gg * i_ * (dz - a1 - xh * a2);
        };
        float4 o;
        o.x = one(yv.x, d.x, k.mu.x, k.is.x, k.g.x, k.bt.x, s1.x, s2.x); o.y = one(yv.y, d.y, k.mu.y, k.is.y, k.g.y, k.bt.y, s1.y, s2.y);
        o.z = one(yv.z, d.z, k.mu.z, k.is.z, k.g.z, k.bt.z, s1.z, s2.z); o.w = one(yv.w, d.w, k.mu.w, k.is.w, k.g.w, k.bt.w, s1.w, s2.w);
        *reinterpret_cast<float4*>(dy + (size_t)r * ld_dy + c) = o;
    }
}

// ---- VectorNonLin (nonlin.py:63-82) -----------------------------------------------------------
// n = |v|;  s = relu(gamma (n - mean) invstd + beta)  [batchnorm]   or  relu(n + bias)  [no BN]
// out = v * s / max(n, 1e-8)
__device__ __forceinline__ float vnl_shift(float n, int c, const float* mean, const float* invstd, const float* gamma,
                                           const float* beta, const float* bias) {
    if (mean) {
        float g = gamma ? __ldg(gamma + c) : 1.f, b = beta ? __ldg(beta + c) : 0.f;
        return g * ((n - __ldg(mean + c)) * __ldg(invstd + c)) + b;
    }
    return n + __ldg(bias + c);
}

__global__ void vecnonlin_fwd_kernel(const float* __restrict__ v, long long ldv, long long M, int C,
                                     const float* __restrict__ mean, const float* __restrict__ invstd,
                                     const float* __restrict__ gamma, const float* __restrict__ beta,
                                     const float* __restrict__ bias, float* __restrict__ out, long long ldo) {
    const long long total = M * (long long)C;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        long long r = t / C; int c = (int)(t - r * C);
        float a = __ldg(v + (2 * r) * ldv + c), b = __ldg(v + (2 * r + 1) * ldv + c);
        float n = sqrtf(a * a + b * b);
        float s = fmaxf(vnl_shift(n, c, mean, invstd, gamma, beta, bias), 0.f) / fmaxf(n, 1e-8f);
        out[(2 * r) * ldo + c] = a * s;
        out[(2 * r + 1) * ldo + c] = b * s;
    }
}

// Backward.  With q = relu(shift(n)), h = q / max(n, eps):  out = v h.
//   d out / d v = h I + v (dh/dn) (v^T / n),  dh/dn = q'/max(n,eps) - q [n > eps] / n^2
// dq (the gradient reaching the BN output through q) = (dout . v) / max(n, eps); its BN backward
// needs the column sums of dq*relu' and dq*relu'*xhat over all points (training mode).
struct VnlBwdVal {
    const float* v; long long ldv; const float* dout; long long ldd;
    const float* mean; const float* invstd; const float* gamma; const float* beta;
    __device__ void operator()(long long r, int c, float (&o)[2]) const {
        float a = __ldg(v + (2 * r) * ldv + c), b = __ldg(v + (2 * r + 1) * ldv + c);
        float n = sqrtf(a * a + b * b);
        float xh = (n - __ldg(mean + c)) * __ldg(invstd + c);
        float g = gamma ? __ldg(gamma + c) : 1.f, be = beta ? __ldg(beta + c) : 0.f;
        float z = g * xh + be;
        float dq = (__ldg(dout + (2 * r) * ldd + c) * a + __ldg(dout + (2 * r + 1) * ldd + c) * b) / fmaxf(n, 1e-8f);
        float dz = z > 0.f ? dq : 0.f;
        o[0] = dz; o[1] = dz * xh;
    }
    __device__ BnColumns columns(int c) const { return bn_columns(mean, invstd, gamma, beta, c); }
    __device__ void quad(long long r, int c, const BnColumns& k, float (&o)[4][2]) const {
        const float4 a4 = __ldg((const float4*)(v + (2 * r) * ldv + c)), b4 = __ldg((const float4*)(v + (2 * r + 1) * ldv + c));
        const float4 da4 = __ldg((const float4*)(dout + (2 * r) * ldd + c)), db4 = __ldg((const float4*)(dout + (2 * r + 1) * ldd + c));
        const float aa[4] = {a4.x, a4.y, a4.z, a4.w}, bb[4] = {b4.x, b4.y, b4.z, b4.w}, da[4] = {da4.x, da4.y, da4.z, da4.w},
                    db[4] = {db4.x, db4.y, db4.z, db4.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float n = sqrtf(aa[u] * aa[u] + bb[u] * bb[u]);
            const float xh = (n - k.mean[u]) * k.invstd[u];
            const float z = k.gamma[u] * xh + k.beta[u];
            const float dq = (da[u] * aa[u] + db[u] * bb[u]) / fmaxf(n, 1e-8f);
            const float dz = z > 0.f ? dq : 0.f;
            o[u][0] = dz; o[u][1] = dz * xh;
        }
    }
};

// float4 variants of the two element-wise VectorNonLin kernels (C % 4 == 0, aligned): same per-element math
__global__ void vecnonlin_fwd_vec4_kernel(const float* __restrict__ v, long long ldv, long long M, int C4,
                                          const float* __restrict__ mean, const float* __restrict__ invstd,
                                          const float* __restrict__ gamma, const float* __restrict__ beta,
                                          const float* __restrict__ bias, float* __restrict__ out, long long ldo) {
    const long long total = M * (long long)C4;
    for (RowCol rc(C4); rc.r < M; rc.next()) {            // grid-stride over [M, C4] without a 64-bit division per element
        const long long r = rc.r;
        const int c = rc.c * 4;
        const float4 a = ld_stream4(reinterpret_cast<const float4*>(v + (2 * r) * ldv + c));
        const float4 b = ld_stream4(reinterpret_cast<const float4*>(v + (2 * r + 1) * ldv + c));
        const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
        float oa[4], ob[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            float n = sqrtf(av[u] * av[u] + bv[u] * bv[u]);
            float s_ = fmaxf(vnl_shift(n, c + u, mean, invstd, gamma, beta, bias), 0.f) / fmaxf(n, 1e-8f);
            oa[u] = av[u] * s_; ob[u] = bv[u] * s_;
        }
        *reinterpret_cast<float4*>(out + (2 * r) * ldo + c) = make_float4(oa[0], oa[1], oa[2], oa[3]);
        *reinterpret_cast<float4*>(out + (2 * r + 1) * ldo + c) = make_float4(ob[0], ob[1], ob[2], ob[3]);
    }
}

__device__ __forceinline__ void vnl_bwd_one(float a, float b, float da, float db, int c, long long r, int C, const float* mean, const float* invstd,
                                            const float* gamma, const float* beta, const float* bias, const float* sum_dz,
                                            const float* sum_dz_xh, int training, float invM, float* dbias_rows, float& oa, float& ob) {
    float n = sqrtf(a * a + b * b), nc = fmaxf(n, 1e-8f);
    float z, dn_from_q;        // dn_from_q: gradient w.r.t. n through q = relu(shift(n))
    float dq = (da * a + db * b) / nc;
    if (mean) {
        float is = __ldg(invstd + c);
        float xh = (n - __ldg(mean + c)) * is;
        float g = gamma ? __ldg(gamma + c) : 1.f, be = beta ? __ldg(beta + c) : 0.f;
        z = g * xh + be;
        float dz = z > 0.f ? dq : 0.f;
        float core = training ? dz - __ldg(sum_dz + c) * invM - xh * (__ldg(sum_dz_xh + c) * invM) : dz;
        dn_from_q = g * is * core;
    } else {
        z = n + __ldg(bias + c);
        dn_from_q = z > 0.f ? dq : 0.f;
        if (dbias_rows) dbias_rows[r * C + c] = dn_from_q;
    }
    float q = fmaxf(z, 0.f);
    float h = q / nc;
    // through the 1/max(n,eps) factor: -q/n^2 * (dout . v) when n > eps
    float dn_from_h = n > 1e-8f ? -(da * a + db * b) * q / (nc * nc) : 0.f;
    float dn = dn_from_q + dn_from_h;
    float inv_n = n > 0.f ? 1.f / n : 0.f;     // d n / d v = v / n   (0 at the origin)
    oa = da * h + dn * a * inv_n;
    ob = db * h + dn * b * inv_n;
}

__global__ void vecnonlin_bwd_kernel(const float* __restrict__ v, long long ldv, const float* __restrict__ dout, long long ldd,
                                     long long M, int C, const float* __restrict__ mean, const float* __restrict__ invstd,
                                     const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ bias,
                                     const float* __restrict__ sum_dz, const float* __restrict__ sum_dz_xh, int training,
                                     float* __restrict__ dv, long long ld_dv, float* __restrict__ dbias_rows) {
    const long long total = M * (long long)C;
    const float invM = 1.f / (float)M;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        long long r = t / C; int c = (int)(t - r * C);
        float oa, ob;
        vnl_bwd_one(__ldg(v + (2 * r) * ldv + c), __ldg(v + (2 * r + 1) * ldv + c), __ldg(dout + (2 * r) * ldd + c),
                    __ldg(dout + (2 * r + 1) * ldd + c), c, r, C, mean, invstd, gamma, beta, bias, sum_dz, sum_dz_xh, training, invM,
                    dbias_rows, oa, ob);
        dv[(2 * r) * ld_dv + c] = oa;
        dv[(2 * r + 1) * ld_dv + c] = ob;
    }
}

__global__ void vecnonlin_bwd_vec4_kernel(const float* __restrict__ v, long long ldv, const float* __restrict__ dout, long long ldd,
                                          long long M, int C, const float* __restrict__ mean, const float* __restrict__ invstd,
                                          const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ bias,
                                          const float* __restrict__ sum_dz, const float* __restrict__ sum_dz_xh, int training,
                                          float* __restrict__ dv, long long ld_dv, float* __restrict__ dbias_rows) {
    const int C4 = C >> 2;
    const long long total = M * (long long)C4;
    const float invM = 1.f / (float)M;
    for (RowCol rc(C4); rc.r < M; rc.next()) {            // grid-stride over [M, C4] without a 64-bit division per element
        const long long r = rc.r;
        const int c = rc.c * 4;
        const float4 a = ld_stream4(reinterpret_cast<const float4*>(v + (2 * r) * ldv + c));
        const float4 b = ld_stream4(reinterpret_cast<const float4*>(v + (2 * r + 1) * ldv + c));
        const float4 da = ld_stream4(reinterpret_cast<const float4*>(dout + (2 * r) * ldd + c));
        const float4 db = ld_stream4(reinterpret_cast<const float4*>(dout + (2 * r + 1) * ldd + c));
        const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w}, dav[4] = {da.x, da.y, da.z, da.w}, dbv[4] = {db.x, db.y, db.z, db.w};
        float oa[4], ob[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            vnl_bwd_one(av[u], bv[u], dav[u], dbv[u], c + u, r, C, mean, invstd, gamma, beta, bias, sum_dz, sum_dz_xh, training, invM,
                        dbias_rows, oa[u], ob[u]);
        *reinterpret_cast<float4*>(dv + (2 * r) * ld_dv + c) = make_float4(oa[0], oa[1], oa[2], oa[3]);
        *reinterpret_cast<float4*>(dv + (2 * r + 1) * ld_dv + c) = make_float4(ob[0], ob[1], ob[2], ob[3]);
    }
}

// ---- kNN max aggregation (deltaconv.py:54) ----------------------------------------------------
// out[i,c] = max_e y[idx[i,e], c]; arg[i,c] = lowest slot attaining it (parity rule)
__global__ void gathermax_fwd_kernel(const int* __restrict__ idx, int N, int k, int C, const float* __restrict__ y, long long ldy,
                                     float* __restrict__ out, long long ldo, unsigned char* __restrict__ arg) {
    // one warp per point, lanes over channels (coalesced row gathers)
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= N) return;
    const int i = (int)warp;
    for (int c0 = 0; c0 < C; c0 += 32 * 4) {
        float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        int slot[4] = {0, 0, 0, 0};
        for (int e = 0; e < k; ++e) {
            int j = __ldg(idx + (size_t)i * k + e);
            if ((unsigned)j >= (unsigned)N) continue;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                int c = c0 + u * 32 + lane;
                if (c < C) {
                    float val = __ldg(y + (size_t)j * ldy + c);
                    if (val > best[u]) { best[u] = val; slot[u] = e; }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            int c = c0 + u * 32 + lane;
            if (c < C) {
                out[(size_t)i * ldo + c] = best[u] == -INFINITY ? 0.f : best[u];    // empty group -> 0 (torch_scatter)
                arg[(size_t)i * C + c] = (unsigned char)slot[u];
            }
        }
    }
}

// dY[j,c] = sum over incoming edges (i,e) with arg[i,c] == e of dOut[i,c]
__global__ void gathermax_bwd_kernel(const int* __restrict__ in_ptr, const int* __restrict__ in_edge, int N, int k, int C,
                                     const unsigned char* __restrict__ arg, const float* __restrict__ dout, long long ldd,
                                     float* __restrict__ dy, long long ld_dy) {
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= N) return;
    const int j = (int)warp;
    const int b = __ldg(in_ptr + j), e_ = __ldg(in_ptr + j + 1);
    for (int c0 = 0; c0 < C; c0 += 32 * 4) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int t = b; t < e_; ++t) {
            int edge = __ldg(in_edge + t);
            int i = edge / k, e = edge - i * k;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                int c = c0 + u * 32 + lane;
                if (c < C && arg[(size_t)i * C + c] == (unsigned char)e) acc[u] += __ldg(dout + (size_t)i * ldd + c);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            int c = c0 + u * 32 + lane;
            if (c < C) dy[(size_t)j * ld_dy + c] = acc[u];
        }
    }
}

struct BnAct {
    const float* mean; const float* invstd; const float* gamma; const float* beta; float slope;
};

// float4 variants: a group of LPP lanes owns one point and 4*LPP consecutive channels (like spmm_kernel).
// BN = true: the gathered rows are PRE-activations y and the result is max_e act(bn(y[idx[i,e]])) without the activation ever
// being written: z -> leaky_relu(gamma (z - mean) invstd + beta) is monotone per channel (increasing for gamma >= 0, decreasing
// otherwise; invstd > 0, slope > 0), so the maximum of the activations is the activation of the per-channel maximum — or
// minimum — of y.  The arg slot is that extreme's, and the value is computed from that very element: bitwise what a separate
// BatchNorm + activation pass followed by the plain gather-max gives.
template <bool BN>
__global__ void __launch_bounds__(256) gathermax_fwd_vec4_kernel(const int* __restrict__ idx, int N, int k, int C, const float* __restrict__ y,
                                                                 long long ldy, float* __restrict__ out, long long ldo,
                                                                 unsigned char* __restrict__ arg, int lpp_log2, BnAct bn) {
    const int lpp = 1 << lpp_log2, lane = threadIdx.x & 31, sub = lane & (lpp - 1);
    const long long row_ll = ((long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (32 >> lpp_log2) + (lane >> lpp_log2);
    const bool row_ok = row_ll < N;
    const int i = row_ok ? (int)row_ll : 0;
    const int ch = (blockIdx.y * lpp + sub) * 4;
    const bool ch_ok = ch < C;
    float4 best = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    uchar4 slot = make_uchar4(0, 0, 0, 0);
    // sign of the channel's slope: invstd > 0, so it is the sign bit of gamma; one bit per channel in ONE register (the
    // kernel hides its gather latency by occupancy: the other BatchNorm constants are loaded after the loop)
    unsigned neg = 0u;
    if (BN && ch_ok && bn.gamma) {
        const float4 ga = __ldg(reinterpret_cast<const float4*>(bn.gamma + ch));
        neg = (__float_as_uint(ga.x) >> 31) | ((__float_as_uint(ga.y) >> 31) << 1) | ((__float_as_uint(ga.z) >> 31) << 2) | ((__float_as_uint(ga.w) >> 31) << 3);
    }
    for (int e0 = 0; e0 < k; e0 += lpp) {
        int my_j = -1;
        if (row_ok && e0 + sub < k) { my_j = __ldg(idx + (size_t)i * k + e0 + sub); if ((unsigned)my_j >= (unsigned)N) my_j = -1; }
        const int cnt = min(lpp, k - e0);
#pragma unroll 4
        for (int q = 0; q < cnt; ++q) {
            const int j = __shfl_sync(FULL, my_j, q, lpp);
            if (!row_ok || !ch_ok || j < 0) continue;
            float4 v = __ldg(reinterpret_cast<const float4*>(y + (size_t)j * ldy + ch));
            if (BN) {            // flip the sign bit where gamma < 0: the order of (+-)y is the order of act(bn(y))
                v.x = __uint_as_float(__float_as_uint(v.x) ^ ((neg & 1u) << 31)); v.y = __uint_as_float(__float_as_uint(v.y) ^ ((neg & 2u) << 30));
                v.z = __uint_as_float(__float_as_uint(v.z) ^ ((neg & 4u) << 29)); v.w = __uint_as_float(__float_as_uint(v.w) ^ ((neg & 8u) << 28));
            }
            const unsigned char e = (unsigned char)(e0 + q);
            if (v.x > best.x) { best.x = v.x; slot.x = e; }
            if (v.y > best.y) { best.y = v.y; slot.y = e; }
            if (v.z > best.z) { best.z = v.z; slot.z = e; }
            if (v.w > best.w) { best.w = v.w; slot.w = e; }
        }
    }
    if (!row_ok || !ch_ok) return;
    if (BN) {
        const float4 mu = __ldg(reinterpret_cast<const float4*>(bn.mean + ch)), is = __ldg(reinterpret_cast<const float4*>(bn.invstd + ch));
        const float4 ga = bn.gamma ? __ldg(reinterpret_cast<const float4*>(bn.gamma + ch)) : make_float4(1.f, 1.f, 1.f, 1.f);
        const float4 be = bn.beta ? __ldg(reinterpret_cast<const float4*>(bn.beta + ch)) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float b4[4] = {best.x, best.y, best.z, best.w}, m4[4] = {mu.x, mu.y, mu.z, mu.w};
        const float i4[4] = {is.x, is.y, is.z, is.w}, g4[4] = {ga.x, ga.y, ga.z, ga.w}, e4[4] = {be.x, be.y, be.z, be.w};
        float o4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float ysel = __uint_as_float(__float_as_uint(b4[u]) ^ (((neg >> u) & 1u) << 31));      // the selected y itself
            float z = g4[u] * ((ysel - m4[u]) * i4[u]) + e4[u];                    // the expression of bn_act_fwd_kernel on it
            z = z > 0.f ? z : bn.slope * z;
            o4[u] = b4[u] == -INFINITY ? 0.f : z;                                  // empty group -> 0 (torch_scatter)
        }
        best = make_float4(o4[0], o4[1], o4[2], o4[3]);
    } else {
        best.x = best.x == -INFINITY ? 0.f : best.x; best.y = best.y == -INFINITY ? 0.f : best.y;
        best.z = best.z == -INFINITY ? 0.f : best.z; best.w = best.w == -INFINITY ? 0.f : best.w;
    }
    *reinterpret_cast<float4*>(out + (size_t)i * ldo + ch) = best;
    *reinterpret_cast<uchar4*>(arg + (size_t)i * C + ch) = slot;
}

__global__ void __launch_bounds__(256) gathermax_bwd_vec4_kernel(const int* __restrict__ in_ptr, const int* __restrict__ in_edge, int N, int k,
                                                                 int C, const unsigned char* __restrict__ arg, const float* __restrict__ dout,
                                                                 long long ldd, float* __restrict__ dy, long long ld_dy, int lpp_log2) {
    const int lpp = 1 << lpp_log2, lane = threadIdx.x & 31, sub = lane & (lpp - 1);
    const long long row_ll = ((long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (32 >> lpp_log2) + (lane >> lpp_log2);
    const bool row_ok = row_ll < N;
    const int j = row_ok ? (int)row_ll : 0;
    const int ch = (blockIdx.y * lpp + sub) * 4;
    const bool ch_ok = ch < C;
    const int t_begin = row_ok ? __ldg(in_ptr + j) : 0, len = row_ok ? __ldg(in_ptr + j + 1) - t_begin : 0;
    int max_len = len;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) max_len = max(max_len, __shfl_xor_sync(FULL, max_len, o));
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int t0 = 0; t0 < max_len; t0 += lpp) {
        // the lane that loads an edge id also splits it into (source row, slot): one integer division per edge, not per lane
        int my_src = -1, my_slot = 0;
        if (t0 + sub < len) {
            const int edge = __ldg(in_edge + t_begin + t0 + sub);
            my_src = edge / k;
            my_slot = edge - my_src * k;
        }
        const int cnt = min(lpp, max_len - t0);
        // (loading dout only where one of the lane's four channels matches — per edge, or batched four edges at a time with the
        // arg bytes first — was measured 1.5-1.7x slower than this unconditional form: the kernel is bound by instruction issue
        // and load latency, not by the L2 traffic of the gather)
#pragma unroll 4
        for (int q = 0; q < cnt; ++q) {
            const int i = __shfl_sync(FULL, my_src, q, lpp);
            const unsigned char e = (unsigned char)__shfl_sync(FULL, my_slot, q, lpp);
            if (!ch_ok || i < 0) continue;
            const uchar4 a = *reinterpret_cast<const uchar4*>(arg + (size_t)i * C + ch);
            const float4 d = __ldg(reinterpret_cast<const float4*>(dout + (size_t)i * ldd + ch));
            if (a.x == e) acc.x += d.x;
            if (a.y == e) acc.y += d.y;
            if (a.z == e) acc.z += d.z;
            if (a.w == e) acc.w += d.w;
        }
    }
    if (row_ok && ch_ok) *reinterpret_cast<float4*>(dy + (size_t)j * ld_dy + ch) = acc;
}

// ---- centralized first layer helpers (deltaconv.py:50-52) -----------------------------------
// edge features x_j - x_i for every slot: [N*k, C]
__global__ void edge_diff_kernel(const int* __restrict__ idx, int N, int k, int C, const float* __restrict__ x, long long ldx,
                                 float* __restrict__ out) {
    const long long total = (long long)N * k * C;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        long long edge = t / C; int c = (int)(t - edge * C);
        int i = (int)(edge / k);
        int j = __ldg(idx + edge);
        if ((unsigned)j >= (unsigned)N) j = i;
        out[t] = __ldg(x + (size_t)j * ldx + c) - __ldg(x + (size_t)i * ldx + c);
    }
}

// transpose of the above: dX[p] = sum_{in-edges (i,e)->p} dE[i,e] - sum_e dE[p,e]
__global__ void edge_diff_bwd_kernel(const int* __restrict__ in_ptr, const int* __restrict__ in_edge, int N, int k, int C,
                                     const float* __restrict__ dE, float* __restrict__ dx, long long ld_dx) {
    const long long total = (long long)N * C;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        int p = (int)(t / C), c = (int)(t - (long long)p * C);
        float acc = 0.f;
        for (int q = __ldg(in_ptr + p); q < __ldg(in_ptr + p + 1); ++q) acc += __ldg(dE + (size_t)__ldg(in_edge + q) * C + c);
        for (int e = 0; e < k; ++e) acc -= __ldg(dE + ((size_t)p * k + e) * C + c);
        dx[(size_t)p * ld_dx + c] = acc;
    }
}

// max over the k consecutive rows of each point: in [N*k, C] -> out [N, C], arg slot
__global__ void segmax_fwd_kernel(const float* __restrict__ in, int N, int k, int C, const int* __restrict__ idx,
                                  float* __restrict__ out, long long ldo, unsigned char* __restrict__ arg) {
    const long long total = (long long)N * C;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        int i = (int)(t / C), c = (int)(t - (long long)i * C);
        float best = -INFINITY; int slot = 0;
        for (int e = 0; e < k; ++e) {
            if (idx && (unsigned)__ldg(idx + (size_t)i * k + e) >= (unsigned)N) continue;
            float v = __ldg(in + ((size_t)i * k + e) * C + c);
            if (v > best) { best = v; slot = e; }
        }
        out[(size_t)i * ldo + c] = best == -INFINITY ? 0.f : best;
        arg[t] = (unsigned char)slot;
    }
}

__global__ void segmax_bwd_kernel(const float* __restrict__ dout, long long ldd, int N, int k, int C,
                                  const unsigned char* __restrict__ arg, float* __restrict__ din) {
    const long long total = (long long)N * k * C;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        long long edge = t / C; int c = (int)(t - edge * C);
        int i = (int)(edge / k), e = (int)(edge - (long long)i * k);
        din[t] = arg[(size_t)i * C + c] == (unsigned char)e ? __ldg(dout + (size_t)i * ldd + c) : 0.f;
    }
}

// float4 variants of the two kernels above (C % 4 == 0, 16-byte aligned rows): the [N*k, C] tensors of the generic centralized
// layer are hundreds of MB (ShapeNet configuration: 983,040 edge rows x 64 channels), the scalar kernels spend their time on
// 64-bit index arithmetic and 4-byte accesses.  Forward: a thread owns (point, 4 channels) and keeps 6 edge rows in flight.
__global__ void __launch_bounds__(256) segmax_fwd_vec4_kernel(const float* __restrict__ in, int N, int k, int C4, const int* __restrict__ idx,
                                                              float* __restrict__ out, long long ldo, unsigned char* __restrict__ arg) {
    for (RowCol rc(C4); rc.r < N; rc.next()) {
        const int i = (int)rc.r, c = rc.c * 4;
        const float4* row = reinterpret_cast<const float4*>(in + ((size_t)i * k) * (4 * C4) + c);
        float4 best = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        uchar4 slot = make_uchar4(0, 0, 0, 0);
        for (int e0 = 0; e0 < k; e0 += 6) {
            float4 v[6];
            bool ok[6];
#pragma unroll
            for (int u = 0; u < 6; ++u) {
                ok[u] = e0 + u < k && !(idx && (unsigned)__ldg(idx + (size_t)i * k + e0 + u) >= (unsigned)N);
                if (e0 + u < k) v[u] = ld_stream4(row + (size_t)(e0 + u) * C4);
            }
#pragma unroll
            for (int u = 0; u < 6; ++u)
                if (ok[u]) {
                    const unsigned char e = (unsigned char)(e0 + u);
                    if (v[u].x > best.x) { best.x = v[u].x; slot.x = e; }
                    if (v[u].y > best.y) { best.y = v[u].y; slot.y = e; }
                    if (v[u].z > best.z) { best.z = v[u].z; slot.z = e; }
                    if (v[u].w > best.w) { best.w = v[u].w; slot.w = e; }
                }
        }
        best.x = best.x == -INFINITY ? 0.f : best.x; best.y = best.y == -INFINITY ? 0.f : best.y;
        best.z = best.z == -INFINITY ? 0.f : best.z; best.w = best.w == -INFINITY ? 0.f : best.w;
        *reinterpret_cast<float4*>(out + (size_t)i * ldo + c) = best;
        *reinterpret_cast<uchar4*>(arg + (size_t)i * (4 * C4) + c) = slot;
    }
}

// Backward: a thread owns (point, 4 channels), reads arg and dOut once and streams the k rows of dIn (one element per row
// and channel is non-zero).
__global__ void __launch_bounds__(256) segmax_bwd_vec4_kernel(const float* __restrict__ dout, long long ldd, int N, int k, int C4,
                                                              const unsigned char* __restrict__ arg, float* __restrict__ din) {
    for (RowCol rc(C4); rc.r < N; rc.next()) {
        const int i = (int)rc.r, c = rc.c * 4;
        const uchar4 a = *reinterpret_cast<const uchar4*>(arg + (size_t)i * (4 * C4) + c);
        const float4 d = __ldg(reinterpret_cast<const float4*>(dout + (size_t)i * ldd + c));
        float4* row = reinterpret_cast<float4*>(din + ((size_t)i * k) * (4 * C4) + c);
        for (int e = 0; e < k; ++e) {
            const unsigned char ee = (unsigned char)e;
            st_stream4(row + (size_t)e * C4, make_float4(a.x == ee ? d.x : 0.f, a.y == ee ? d.y : 0.f, a.z == ee ? d.z : 0.f, a.w == ee ? d.w : 0.f));
        }
    }
}

// The last block of a deeper edge MLP fused with the max over the edges (the segment-max analogue of gathermax_fwd_vec4_kernel
// <true>): `in` holds the PRE-activations y [N*k, C]; out[i,c] = max_e act(bn(y[i*k+e, c])) by selecting the per-channel extreme
// of (+-)y — the activation [N*k, C] is neither written nor read back.  The value is computed from the selected element with the
// expression of bn_act_fwd_kernel: bitwise what BatchNorm + activation followed by segmax_fwd_vec4_kernel gives.
__global__ void __launch_bounds__(256) segmax_bn_fwd_vec4_kernel(const float* __restrict__ in, int N, int k, int C4, const int* __restrict__ idx,
                                                                 float* __restrict__ out, long long ldo, unsigned char* __restrict__ arg, BnAct bn) {
    for (RowCol rc(C4); rc.r < N; rc.next()) {
        const int i = (int)rc.r, c = rc.c * 4;
        const float4* row = reinterpret_cast<const float4*>(in + ((size_t)i * k) * (4 * C4) + c);
        unsigned neg = 0u;
        if (bn.gamma) {
            const float4 ga = __ldg(reinterpret_cast<const float4*>(bn.gamma + c));
            neg = (__float_as_uint(ga.x) >> 31) | ((__float_as_uint(ga.y) >> 31) << 1) | ((__float_as_uint(ga.z) >> 31) << 2) | ((__float_as_uint(ga.w) >> 31) << 3);
        }
        float4 best = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        uchar4 slot = make_uchar4(0, 0, 0, 0);
        for (int e0 = 0; e0 < k; e0 += 6) {
            float4 v[6];
            bool ok[6];
#pragma unroll
            for (int u = 0; u < 6; ++u) {
                ok[u] = e0 + u < k && !(idx && (unsigned)__ldg(idx + (size_t)i * k + e0 + u) >= (unsigned)N);
                if (e0 + u < k) v[u] = ld_stream4(row + (size_t)(e0 + u) * C4);
            }
#pragma unroll
            for (int u = 0; u < 6; ++u)
                if (ok[u]) {
                    const unsigned char e = (unsigned char)(e0 + u);
                    const float vx = __uint_as_float(__float_as_uint(v[u].x) ^ ((neg & 1u) << 31)), vy = __uint_as_float(__float_as_uint(v[u].y) ^ ((neg & 2u) << 30));
                    const float vz = __uint_as_float(__float_as_uint(v[u].z) ^ ((neg & 4u) << 29)), vw = __uint_as_float(__float_as_uint(v[u].w) ^ ((neg & 8u) << 28));
                    if (vx > best.x) { best.x = vx; slot.x = e; }
                    if (vy > best.y) { best.y = vy; slot.y = e; }
                    if (vz > best.z) { best.z = vz; slot.z = e; }
                    if (vw > best.w) { best.w = vw; slot.w = e; }
                }
        }
        const float4 mu = __ldg(reinterpret_cast<const float4*>(bn.mean + c)), is = __ldg(reinterpret_cast<const float4*>(bn.invstd + c));
        const float4 ga = bn.gamma ? __ldg(reinterpret_cast<const float4*>(bn.gamma + c)) : make_float4(1.f, 1.f, 1.f, 1.f);
        const float4 be = bn.beta ? __ldg(reinterpret_cast<const float4*>(bn.beta + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float b4[4] = {best.x, best.y, best.z, best.w}, m4[4] = {mu.x, mu.y, mu.z, mu.w};
        const float i4[4] = {is.x, is.y, is.z, is.w}, g4[4] = {ga.x, ga.y, ga.z, ga.w}, e4[4] = {be.x, be.y, be.z, be.w};
        float o4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float ysel = __uint_as_float(__float_as_uint(b4[u]) ^ (((neg >> u) & 1u) << 31));
            float z = g4[u] * ((ysel - m4[u]) * i4[u]) + e4[u];
            z = z > 0.f ? z : bn.slope * z;
            o4[u] = b4[u] == -INFINITY ? 0.f : z;
        }
        *reinterpret_cast<float4*>(out + (size_t)i * ldo + c) = make_float4(o4[0], o4[1], o4[2], o4[3]);
        *reinterpret_cast<uchar4*>(arg + (size_t)i * (4 * C4) + c) = slot;
    }
}

// Backward of that node.  The gradient of the activation is non-zero only at the arg-max edge of each (point, channel), so the
// two BatchNorm sums run over the N points (y is read at the selected rows only) ...
struct SegBnBwdVal {
    const float* y; int k, C; const unsigned char* arg; const float* dout; long long ldd;
    const float* mean; const float* invstd; const float* gamma; const float* beta; float slope;
    __device__ void operator()(long long r, int c, float (&o)[2]) const {
        const int e = arg[(size_t)r * C + c];
        const float xh = (__ldg(y + ((size_t)r * k + e) * C + c) - __ldg(mean + c)) * __ldg(invstd + c);
        const float g = gamma ? __ldg(gamma + c) : 1.f, b = beta ? __ldg(beta + c) : 0.f;
        const float z = g * xh + b;
        const float dz = __ldg(dout + r * ldd + c) * (z > 0.f ? 1.f : slope);
        o[0] = dz; o[1] = dz * xh;
    }
    __device__ BnColumns columns(int c) const { return bn_columns(mean, invstd, gamma, beta, c); }
    __device__ void quad(long long r, int c, const BnColumns& kc, float (&o)[4][2]) const {
        const uchar4 a = *reinterpret_cast<const uchar4*>(arg + (size_t)r * C + c);
        const float4 dv = __ldg(reinterpret_cast<const float4*>(dout + r * ldd + c));
        const float* base = y + (size_t)r * k * C + c;
        const float yy[4] = {__ldg(base + (size_t)a.x * C), __ldg(base + (size_t)a.y * C + 1), __ldg(base + (size_t)a.z * C + 2),
                             __ldg(base + (size_t)a.w * C + 3)};
        const float dd[4] = {dv.x, dv.y, dv.z, dv.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float xh = (yy[u] - kc.mean[u]) * kc.invstd[u];
            const float z = kc.gamma[u] * xh + kc.beta[u];
            const float dz = dd[u] * (z > 0.f ? 1.f : slope);
            o[u][0] = dz; o[u][1] = dz * xh;
        }
    }
};

// ... and the elementwise part streams y [N*k, C] once and writes dy once: a thread owns (point, 4 channels), the sparse
// gradient of the activation (dout at the arg slot) is formed in registers.  M = N*k rows share the statistics.
__global__ void __launch_bounds__(256) segmax_bn_bwd_vec4_kernel(const float* __restrict__ y, const float* __restrict__ dout, long long ldd,
                                                                 int N, int k, int C4, const unsigned char* __restrict__ arg, BnAct bn,
                                                                 const float* __restrict__ sum_dz, const float* __restrict__ sum_dz_xh,
                                                                 int training, float* __restrict__ dy) {
    const float invM = 1.f / ((float)N * (float)k);
    for (RowCol rc(C4); rc.r < N; rc.next()) {
        const int i = (int)rc.r, c = rc.c * 4;
        const uchar4 a = *reinterpret_cast<const uchar4*>(arg + (size_t)i * (4 * C4) + c);
        const float4 d = __ldg(reinterpret_cast<const float4*>(dout + (size_t)i * ldd + c));
        const float4 mu = __ldg(reinterpret_cast<const float4*>(bn.mean + c)), is = __ldg(reinterpret_cast<const float4*>(bn.invstd + c));
        const float4 g = bn.gamma ? __ldg(reinterpret_cast<const float4*>(bn.gamma + c)) : make_float4(1.f, 1.f, 1.f, 1.f);
        const float4 b = bn.beta ? __ldg(reinterpret_cast<const float4*>(bn.beta + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 s1 = make_float4(0.f, 0.f, 0.f, 0.f), s2 = s1;
        if (training) {
            s1 = __ldg(reinterpret_cast<const float4*>(sum_dz + c)); s2 = __ldg(reinterpret_cast<const float4*>(sum_dz_xh + c));
            s1.x *= invM; s1.y *= invM; s1.z *= invM; s1.w *= invM; s2.x *= invM; s2.y *= invM; s2.z *= invM; s2.w *= invM;
        }
        const float slope = bn.slope;
        auto one = [&](float yy, float dd, bool hit, float m, float i_, float gg, float bb, float a1, float a2) {
            const float xh = (yy - m) * i_;
            const float z = gg * xh + bb;
            const float dz = hit ? dd * (z > 0.f ? 1.f : slope) : 0.f;
            return gg * i_ * (dz - a1 - xh * a2);
        };
        const float4* yrow = reinterpret_cast<const float4*>(y + ((size_t)i * k) * (4 * C4) + c);
        float4* drow = reinterpret_cast<float4*>(dy + ((size_t)i * k) * (4 * C4) + c);
        for (int e0 = 0; e0 < k; e0 += 4) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (e0 + u < k) v[u] = ld_stream4(yrow + (size_t)(e0 + u) * C4);
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (e0 + u < k) {
                    const unsigned char ee = (unsigned char)(e0 + u);
                    float4 o;
                    o.x = one(v[u].x, d.x, a.x == ee, mu.x, is.x, g.x, b.x, s1.x, s2.x); o.y = one(v[u].y, d.y, a.y == ee, mu.y, is.y, g.y, b.y, s1.y, s2.y);
                    o.z = one(v[u].z, d.z, a.z == ee, mu.z, is.z, g.z, b.z, s1.z, s2.z); o.w = one(v[u].w, d.w, a.w == ee, mu.w, is.w, g.w, b.w, s1.w, s2.w);
                    st_stream4(drow + (size_t)(e0 + u) * C4, o);
                }
        }
    }
}

// ---- global pools (one CTA per (cloud, 32-channel block); rows of a cloud are contiguous) ------
// out_max[b,c] = max_i x[i,c] (arg kept), out_mean[b,c] = mean_i x[i,c]
// With BN the pooled quantity is leaky_relu(gamma * (x - mean) * invstd + beta): the activation of the embedding layer is
// applied while its pre-activation is read, so the [N,C] activation never exists in memory.

template <bool BN>
__global__ void __launch_bounds__(256) pool_fwd_kernel(const float* __restrict__ x, long long ldx, const int* __restrict__ cloud_ptr,
                                                       int C, float* __restrict__ out_max, long long ld_max,
                                                       float* __restrict__ out_mean, long long ld_mean, int* __restrict__ arg, BnAct bn) {
    __shared__ float smax[8][32];
    __shared__ int sarg[8][32];
    __shared__ float ssum[8][32];
    const int b = blockIdx.x, lane = threadIdx.x & 31, rl = threadIdx.x >> 5;
    const int c = blockIdx.y * 32 + lane;
    const int s = __ldg(cloud_ptr + b), e = __ldg(cloud_ptr + b + 1);
    float best = -INFINITY, sum = 0.f; int bi = s;
    if (c < C) {
        float mu = 0.f, is = 1.f, g = 1.f, bt = 0.f;
        if (BN) {
            mu = __ldg(bn.mean + c); is = __ldg(bn.invstd + c);
            g = bn.gamma ? __ldg(bn.gamma + c) : 1.f; bt = bn.beta ? __ldg(bn.beta + c) : 0.f;
        }
        for (int i = s + rl; i < e; i += 8) {
            float v = __ldg(x + (size_t)i * ldx + c);
            if (BN) { v = g * ((v - mu) * is) + bt; v = v > 0.f ? v : bn.slope * v; }
            sum += v;
            if (v > best) { best = v; bi = i; }
        }
    }
    smax[rl][lane] = best; sarg[rl][lane] = bi; ssum[rl][lane] = sum;
    __syncthreads();
    if (rl == 0 && c < C) {
        for (int q = 1; q < 8; ++q) {
            float v = smax[q][lane]; int a = sarg[q][lane];
            if (v > best || (v == best && a < bi)) { best = v; bi = a; }
            sum += ssum[q][lane];
        }
        int cnt = e - s;
        if (out_max) { out_max[(size_t)b * ld_max + c] = cnt > 0 ? best : 0.f; arg[(size_t)b * C + c] = bi; }
        if (out_mean) out_mean[(size_t)b * ld_mean + c] = sum / (float)(cnt > 0 ? cnt : 1);
    }
}

// one CTA column per cloud: no per-element cloud search, the 1/count is computed once per thread
__global__ void __launch_bounds__(256) pool_bwd_cloud_kernel(const int* __restrict__ cloud_ptr, int C, const int* __restrict__ arg,
                                                             const float* __restrict__ d_max, long long ld_dmax, const float* __restrict__ d_mean,
                                                             long long ld_dmean, float* __restrict__ dx, long long ld_dx) {
    const int b = blockIdx.y, lane = threadIdx.x & 31, rl = threadIdx.x >> 5;
    const int s = __ldg(cloud_ptr + b), e = __ldg(cloud_ptr + b + 1);
    const float inv = 1.f / (float)max(e - s, 1);
    for (int c = lane; c < C; c += 32) {
        const float gm = d_max ? __ldg(d_max + (size_t)b * ld_dmax + c) : 0.f;
        const int am = d_max ? __ldg(arg + (size_t)b * C + c) : -1;
        const float ga = d_mean ? __ldg(d_mean + (size_t)b * ld_dmean + c) * inv : 0.f;
        for (int i = s + blockIdx.x * 8 + rl; i < e; i += gridDim.x * 8) dx[(size_t)i * ld_dx + c] = (i == am ? gm : 0.f) + ga;
    }
}

__global__ void pool_bwd_kernel(const int* __restrict__ cloud_ptr, int B, int N, int C, const int* __restrict__ arg,
                                const float* __restrict__ d_max, long long ld_dmax, const float* __restrict__ d_mean, long long ld_dmean,
                                float* __restrict__ dx, long long ld_dx) {
    const long long total = (long long)N * C;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        int i = (int)(t / C), c = (int)(t - (long long)i * C);
        int b = cloud_of(cloud_ptr, B, i);
        float g = 0.f;
        if (d_max && __ldg(arg + (size_t)b * C + c) == i) g += __ldg(d_max + (size_t)b * ld_dmax + c);
        if (d_mean) g += __ldg(d_mean + (size_t)b * ld_dmean + c) / (float)(__ldg(cloud_ptr + b + 1) - __ldg(cloud_ptr + b));
        dx[(size_t)i * ld_dx + c] = g;
    }
}

// out[i, :] = src[cloud(i), :]  (x_max[batch], deltanet_segmentation.py:61,64) and its transpose
__global__ void bcast_rows_kernel(const float* __restrict__ src, long long lds, const int* __restrict__ cloud_ptr, int B, int N, int C,
                                  float* __restrict__ out, long long ldo) {
    const long long total = (long long)N * C;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        int i = (int)(t / C), c = (int)(t - (long long)i * C);
        out[(size_t)i * ldo + c] = __ldg(src + (size_t)cloud_of(cloud_ptr, B, i) * lds + c);
    }
}

// float4 form: a thread owns (row, 4 columns); no 64-bit division per element, 16-byte accesses (the segmentation head broadcasts
// a [B,1024] embedding over 32768 rows: 134 MB written per step).
__global__ void __launch_bounds__(256) bcast_rows_vec4_kernel(const float* __restrict__ src, long long lds, const int* __restrict__ cloud_ptr,
                                                              int B, int N, int C4, float* __restrict__ out, long long ldo) {
    for (RowCol rc(C4); rc.r < N; rc.next()) {
        const int i = (int)rc.r, c = rc.c * 4;
        const int b = cloud_of(cloud_ptr, B, i);
        st_stream4(reinterpret_cast<float4*>(out + (size_t)i * ldo + c), __ldg(reinterpret_cast<const float4*>(src + (size_t)b * lds + c)));
    }
}

// out[i, :] += src[cloud(i), :] — the per-cloud part of a Linear whose input is [broadcast per-cloud features | per-point
// features]: the product with the broadcast block is formed on the B cloud rows and added here.
__global__ void __launch_bounds__(256) bcast_rows_add_vec4_kernel(const float* __restrict__ src, long long lds, const int* __restrict__ cloud_ptr,
                                                                  int B, int N, int C4, float* __restrict__ out, long long ldo) {
    for (RowCol rc(C4); rc.r < N; rc.next()) {
        const int i = (int)rc.r, c = rc.c * 4;
        const int b = cloud_of(cloud_ptr, B, i);
        float4* o = reinterpret_cast<float4*>(out + (size_t)i * ldo + c);
        const float4 a = *o, s = __ldg(reinterpret_cast<const float4*>(src + (size_t)b * lds + c));
        *o = make_float4(a.x + s.x, a.y + s.y, a.z + s.z, a.w + s.w);
    }
}
__global__ void bcast_rows_add_kernel(const float* __restrict__ src, long long lds, const int* __restrict__ cloud_ptr, int B, int N, int C,
                                      float* __restrict__ out, long long ldo) {
    const long long total = (long long)N * C;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        int i = (int)(t / C), c = (int)(t - (long long)i * C);
        out[(size_t)i * ldo + c] += __ldg(src + (size_t)cloud_of(cloud_ptr, B, i) * lds + c);
    }
}

template <class... P>
bool all_aligned16(P... ptrs) {
    const void* a[] = {(const void*)ptrs...};
    for (const void* p : a)
        if (p != nullptr && (reinterpret_cast<uintptr_t>(p) & 15u)) return false;
    return true;
}

int elementwise_grid(long long total) {
    long long g = (total + 255) / 256;
    long long cap = (long long)dc_sm_count() * 16;
    return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

}  // namespace

// ================================ C ABI ===========================================================

extern "C" size_t dc_colreduce_ws_bytes(int C) { return (size_t)RED_MAX_CHUNKS * 2 * (size_t)C * sizeof(double); }

// mode 0: statistics of x [M,C]; mode 1: statistics of the norms of a vector field x [2M,C]
extern "C" int dc_bn_stats(const float* x, int64_t ldx, int64_t M, int C, int mode, float eps, float momentum,
                           float* mean, float* invstd, float* running_mean_or_null, float* running_var_or_null,
                           int64_t* num_batches_tracked_or_null, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(M >= 1 && C >= 1, "dc_bn_stats: empty input");
    DC_REQUIRE(ws_bytes >= dc_colreduce_ws_bytes(C), "dc_bn_stats: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    double* partial = (double*)ws;
    const bool vec4_ok = ldx % 4 == 0 && all_aligned16(x);
    int chunks;
    if (mode == 0) {
        PlainVal f = {x, ldx};
        chunks = col_reduce_launch<2>(f, M, C, vec4_ok, partial, st);
        bn_finalize_kernel<<<dc_ceil_div(C, FIN_COLS), FIN_T, 0, st>>>(f, partial, chunks, M, C, eps, momentum, mean, invstd,
                                                                        running_mean_or_null, running_var_or_null, (long long*)num_batches_tracked_or_null);
        return dc_check_launch_n("dc_bn_stats", 2);
    }
    NormVal f = {x, ldx};
    chunks = col_reduce_launch<2>(f, M, C, vec4_ok, partial, st);
    bn_finalize_kernel<<<dc_ceil_div(C, FIN_COLS), FIN_T, 0, st>>>(f, partial, chunks, M, C, eps, momentum, mean, invstd,
                                                            running_mean_or_null, running_var_or_null,
                                                            (long long*)num_batches_tracked_or_null);
    return dc_check_launch_n("dc_bn_stats", 2);
}

// out[c] = sum_r x[r,c]  (bias gradients)
extern "C" int dc_col_sum(const float* x, int64_t ldx, int64_t M, int C, float* out, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(M >= 1 && C >= 1, "dc_col_sum: empty input");
    DC_REQUIRE(ws_bytes >= dc_colreduce_ws_bytes(C), "dc_col_sum: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    ColLayout L = col_layout(M, C);
    dim3 grid(L.chunks, L.col_blocks);
    SumVal f = {x, ldx};
    col_reduce_kernel<1><<<grid, RED_T, 0, st>>>(f, M, C, L.col_lanes, L.rows_per_chunk, (double*)ws);
    sum1_finalize_kernel<<<dc_ceil_div(C, FIN_COLS), FIN_T, 0, st>>>((double*)ws, L.chunks, C, out);
    return dc_check_launch_n("dc_col_sum", 2);
}

extern "C" int dc_bn_act_fwd(const float* y, int64_t ldy, int64_t M, int C, const float* mean, const float* invstd,
                             const float* gamma_or_null, const float* beta_or_null, float slope,
                             const float* addend_or_null, int64_t ld_add, float* out, int64_t ldo, void* stream) {
    if (M <= 0 || C <= 0) return DC_OK;
    if (C % 4 == 0 && ldy % 4 == 0 && ldo % 4 == 0 && ld_add % 4 == 0 && all_aligned16(y, out, addend_or_null, mean, invstd, gamma_or_null, beta_or_null))
        bn_act_fwd_vec4_kernel<<<elementwise_grid(M * C / 4), 256, 0, (cudaStream_t)stream>>>(y, ldy, M, C / 4, mean, invstd, gamma_or_null,
                                                                                          beta_or_null, slope, addend_or_null, ld_add, out, ldo);
    else
        bn_act_fwd_kernel<<<elementwise_grid(M * C), 256, 0, (cudaStream_t)stream>>>(y, ldy, M, C, mean, invstd, gamma_or_null,
                                                                                     beta_or_null, slope, addend_or_null, ld_add, out, ldo);
    return dc_check_launch("dc_bn_act_fwd");
}

// d_gamma = sum dz*xhat, d_beta = sum dz (written to the two [C] outputs), dy as documented above
extern "C" int dc_bn_act_bwd(const float* y, int64_t ldy, const float* dout, int64_t ldd, int64_t M, int C,
                             const float* mean, const float* invstd, const float* gamma_or_null, const float* beta_or_null,
                             float slope, int training, float* d_gamma, float* d_beta, float* dy, int64_t ld_dy,
                             void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(M >= 1 && C >= 1, "dc_bn_act_bwd: empty input");
    DC_REQUIRE(ws_bytes >= dc_colreduce_ws_bytes(C), "dc_bn_act_bwd: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    double* partial = (double*)ws;
    BnBwdVal f = {y, ldy, dout, ldd, mean, invstd, gamma_or_null, beta_or_null, slope};
    const int chunks = col_reduce_launch<2>(f, M, C, ldy % 4 == 0 && ldd % 4 == 0 && all_aligned16(y, dout, mean, invstd, gamma_or_null, beta_or_null),
                                            partial, st);
    sum2_finalize_kernel<<<dc_ceil_div(C, FIN_COLS), FIN_T, 0, st>>>(partial, chunks, C, d_beta, d_gamma);
    if (C % 4 == 0 && ldy % 4 == 0 && ldd % 4 == 0 && ld_dy % 4 == 0 && all_aligned16(y, dout, dy, mean, invstd, gamma_or_null, beta_or_null) &&
        all_aligned16(d_beta, d_gamma))
        bn_act_bwd_vec4_kernel<<<elementwise_grid(M * C / 4), 256, 0, st>>>(y, ldy, dout, ldd, M, C / 4, mean, invstd, gamma_or_null, beta_or_null,
                                                                        slope, d_beta, d_gamma, training, dy, ld_dy);
    else
        bn_act_bwd_kernel<<<elementwise_grid(M * C), 256, 0, st>>>(y, ldy, dout, ldd, M, C, mean, invstd, gamma_or_null, beta_or_null,
                                                                   slope, d_beta, d_gamma, training, dy, ld_dy);
    return dc_check_launch_n("dc_bn_act_bwd", 3);
}

extern "C" int dc_vecnonlin_fwd(const float* v, int64_t ldv, int64_t M, int C, const float* mean_or_null,
                                const float* invstd_or_null, const float* gamma_or_null, const float* beta_or_null,
                                const float* bias_or_null, float* out, int64_t ldo, void* stream) {
    DC_REQUIRE(mean_or_null != nullptr || bias_or_null != nullptr, "dc_vecnonlin_fwd: need BN statistics or a bias");
    if (M <= 0 || C <= 0) return DC_OK;
    if (C % 4 == 0 && ldv % 4 == 0 && ldo % 4 == 0 && all_aligned16(v, out))
        vecnonlin_fwd_vec4_kernel<<<elementwise_grid(M * C / 4), 256, 0, (cudaStream_t)stream>>>(v, ldv, M, C / 4, mean_or_null, invstd_or_null,
                                                                                             gamma_or_null, beta_or_null, bias_or_null, out, ldo);
    else
        vecnonlin_fwd_kernel<<<elementwise_grid(M * C), 256, 0, (cudaStream_t)stream>>>(v, ldv, M, C, mean_or_null, invstd_or_null,
                                                                                        gamma_or_null, beta_or_null, bias_or_null, out, ldo);
    return dc_check_launch("dc_vecnonlin_fwd");
}

// With BN: d_gamma/d_beta [C] are produced. Without BN: dbias_rows [M,C] receives the per-row bias
// gradient (the caller sums it over rows), d_gamma/d_beta are ignored.
extern "C" int dc_vecnonlin_bwd(const float* v, int64_t ldv, const float* dout, int64_t ldd, int64_t M, int C,
                                const float* mean_or_null, const float* invstd_or_null, const float* gamma_or_null,
                                const float* beta_or_null, const float* bias_or_null, int training,
                                float* d_gamma, float* d_beta, float* dbias_rows_or_null, float* dv, int64_t ld_dv,
                                void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(M >= 1 && C >= 1, "dc_vecnonlin_bwd: empty input");
    cudaStream_t st = (cudaStream_t)stream;
    if (mean_or_null) {
        DC_REQUIRE(ws_bytes >= dc_colreduce_ws_bytes(C), "dc_vecnonlin_bwd: workspace too small");
        double* partial = (double*)ws;
        VnlBwdVal f = {v, ldv, dout, ldd, mean_or_null, invstd_or_null, gamma_or_null, beta_or_null};
        const int chunks = col_reduce_launch<2>(f, M, C, ldv % 4 == 0 && ldd % 4 == 0 &&
                                                all_aligned16(v, dout, mean_or_null, invstd_or_null, gamma_or_null, beta_or_null), partial, st);
        sum2_finalize_kernel<<<dc_ceil_div(C, FIN_COLS), FIN_T, 0, st>>>(partial, chunks, C, d_beta, d_gamma);
    }
    if (C % 4 == 0 && ldv % 4 == 0 && ldd % 4 == 0 && ld_dv % 4 == 0 && all_aligned16(v, dout, dv))
        vecnonlin_bwd_vec4_kernel<<<elementwise_grid(M * C / 4), 256, 0, st>>>(v, ldv, dout, ldd, M, C, mean_or_null, invstd_or_null, gamma_or_null,
                                                                           beta_or_null, bias_or_null, d_beta, d_gamma, training, dv, ld_dv,
                                                                           dbias_rows_or_null);
    else
        vecnonlin_bwd_kernel<<<elementwise_grid(M * C), 256, 0, st>>>(v, ldv, dout, ldd, M, C, mean_or_null, invstd_or_null, gamma_or_null,
                                                                      beta_or_null, bias_or_null, d_beta, d_gamma, training, dv, ld_dv,
                                                                      dbias_rows_or_null);
    return dc_check_launch_n("dc_vecnonlin_bwd", mean_or_null ? 3 : 1);
}

extern "C" int dc_gathermax_fwd(const int32_t* idx, int N, int k, int C, const float* y, int64_t ldy, float* out, int64_t ldo,
                                uint8_t* arg, void* stream) {
    DC_REQUIRE(k <= 255, "dc_gathermax_fwd: k must be <= 255 (arg slots are stored as u8)");
    if (N <= 0 || C <= 0) return DC_OK;
    if (C % 4 == 0 && ldy % 4 == 0 && ldo % 4 == 0 && all_aligned16(y, out) && (reinterpret_cast<uintptr_t>(arg) & 3u) == 0) {
        int lpp_log2 = 0;
        while ((1 << lpp_log2) < C / 4 && lpp_log2 < 5) ++lpp_log2;
        dim3 grid(dc_ceil_div(N, 8 * (32 >> lpp_log2)), dc_ceil_div(C / 4, 1 << lpp_log2));
        gathermax_fwd_vec4_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(idx, N, k, C, y, ldy, out, ldo, arg, lpp_log2, BnAct());
    } else
        gathermax_fwd_kernel<<<dc_ceil_div((long long)N * 32, 256), 256, 0, (cudaStream_t)stream>>>(idx, N, k, C, y, ldy, out, ldo, arg);
    return dc_check_launch("dc_gathermax_fwd");
}

// max over the neighbours of leaky_relu(BatchNorm(y)) from the pre-activations y (deltaconv.py:54 with the last block of
// s_mlp_max, mlp.py:9, folded in): the [N,C] activation is never written.  Needs the float4 layout (C % 4 == 0, 16-byte rows).
extern "C" int dc_gathermax_bn_fwd(const int32_t* idx, int N, int k, int C, const float* y, int64_t ldy, const float* mean,
                                   const float* invstd, const float* gamma_or_null, const float* beta_or_null, float slope,
                                   float* out, int64_t ldo, uint8_t* arg, void* stream) {
    DC_REQUIRE(k <= 255, "dc_gathermax_bn_fwd: k must be <= 255 (arg slots are stored as u8)");
    DC_REQUIRE(slope > 0.f, "dc_gathermax_bn_fwd: the activation must be strictly increasing (LeakyReLU slope > 0)");
    DC_REQUIRE(C % 4 == 0 && ldy % 4 == 0 && ldo % 4 == 0 && all_aligned16(y, out, mean, invstd, gamma_or_null, beta_or_null) &&
               (reinterpret_cast<uintptr_t>(arg) & 3u) == 0, "dc_gathermax_bn_fwd: needs C %% 4 == 0 and 16-byte aligned rows");
    if (N <= 0 || C <= 0) return DC_OK;
    int lpp_log2 = 0;
    while ((1 << lpp_log2) < C / 4 && lpp_log2 < 5) ++lpp_log2;
    dim3 grid(dc_ceil_div(N, 8 * (32 >> lpp_log2)), dc_ceil_div(C / 4, 1 << lpp_log2));
    BnAct bn = {mean, invstd, gamma_or_null, beta_or_null, slope};
    gathermax_fwd_vec4_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(idx, N, k, C, y, ldy, out, ldo, arg, lpp_log2, bn);
    return dc_check_launch("dc_gathermax_bn_fwd");
}

extern "C" int dc_gathermax_bwd(const int32_t* in_ptr, const int32_t* in_edge, int N, int k, int C, const uint8_t* arg,
                                const float* dout, int64_t ldd, float* dy, int64_t ld_dy, void* stream) {
    if (N <= 0 || C <= 0) return DC_OK;
    if (C % 4 == 0 && ldd % 4 == 0 && ld_dy % 4 == 0 && all_aligned16(dout, dy) && (reinterpret_cast<uintptr_t>(arg) & 3u) == 0) {
        int lpp_log2 = 0;
        while ((1 << lpp_log2) < C / 4 && lpp_log2 < 5) ++lpp_log2;
        dim3 grid(dc_ceil_div(N, 8 * (32 >> lpp_log2)), dc_ceil_div(C / 4, 1 << lpp_log2));
        gathermax_bwd_vec4_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in_ptr, in_edge, N, k, C, arg, dout, ldd, dy, ld_dy, lpp_log2);
    } else
        gathermax_bwd_kernel<<<dc_ceil_div((long long)N * 32, 256), 256, 0, (cudaStream_t)stream>>>(in_ptr, in_edge, N, k, C, arg, dout, ldd, dy, ld_dy);
    return dc_check_launch("dc_gathermax_bwd");
}

extern "C" int dc_edge_diff_fwd(const int32_t* idx, int N, int k, int C, const float* x, int64_t ldx, float* out, void* stream) {
    if (N <= 0 || C <= 0) return DC_OK;
    edge_diff_kernel<<<elementwise_grid((long long)N * k * C), 256, 0, (cudaStream_t)stream>>>(idx, N, k, C, x, ldx, out);
    return dc_check_launch("dc_edge_diff_fwd");
}

extern "C" int dc_edge_diff_bwd(const int32_t* in_ptr, const int32_t* in_edge, int N, int k, int C, const float* d_edge,
                                float* dx, int64_t ld_dx, void* stream) {
    if (N <= 0 || C <= 0) return DC_OK;
    edge_diff_bwd_kernel<<<elementwise_grid((long long)N * C), 256, 0, (cudaStream_t)stream>>>(in_ptr, in_edge, N, k, C, d_edge, dx, ld_dx);
    return dc_check_launch("dc_edge_diff_bwd");
}

extern "C" int dc_segmax_fwd(const float* in, int N, int k, int C, const int32_t* idx_or_null, float* out, int64_t ldo,
                             uint8_t* arg, void* stream) {
    DC_REQUIRE(k <= 255, "dc_segmax_fwd: k must be <= 255");
    if (N <= 0 || C <= 0) return DC_OK;
    if (C % 4 == 0 && ldo % 4 == 0 && all_aligned16(in, out, arg))
        segmax_fwd_vec4_kernel<<<elementwise_grid((long long)N * C / 4), 256, 0, (cudaStream_t)stream>>>(in, N, k, C / 4, idx_or_null, out, ldo, arg);
    else
        segmax_fwd_kernel<<<elementwise_grid((long long)N * C), 256, 0, (cudaStream_t)stream>>>(in, N, k, C, idx_or_null, out, ldo, arg);
    return dc_check_launch("dc_segmax_fwd");
}

extern "C" int dc_segmax_bwd(const float* dout, int64_t ldd, int N, int k, int C, const uint8_t* arg, float* din, void* stream) {
    if (N <= 0 || C <= 0) return DC_OK;
    if (C % 4 == 0 && ldd % 4 == 0 && all_aligned16(dout, din, arg))
        segmax_bwd_vec4_kernel<<<elementwise_grid((long long)N * C / 4), 256, 0, (cudaStream_t)stream>>>(dout, ldd, N, k, C / 4, arg, din);
    else
        segmax_bwd_kernel<<<elementwise_grid((long long)N * k * C), 256, 0, (cudaStream_t)stream>>>(dout, ldd, N, k, C, arg, din);
    return dc_check_launch("dc_segmax_bwd");
}

// max over each point's k edge rows of leaky_relu(BatchNorm(y)) from the pre-activations y [N*k, C] (the last block of a
// centralized s_mlp_max of depth >= 2, deltaconv.py:50-52): the activation is never written.  float4 layout only.
extern "C" int dc_segmax_bn_fwd(const float* y, int N, int k, int C, const int32_t* idx_or_null, const float* mean, const float* invstd,
                                const float* gamma_or_null, const float* beta_or_null, float slope, float* out, int64_t ldo, uint8_t* arg,
                                void* stream) {
    DC_REQUIRE(k >= 1 && k <= 255, "dc_segmax_bn_fwd: k must be in [1,255]");
    DC_REQUIRE(slope > 0.f, "dc_segmax_bn_fwd: the activation must be strictly increasing (LeakyReLU slope > 0)");
    DC_REQUIRE(C % 4 == 0 && ldo % 4 == 0 && all_aligned16(y, out, mean, invstd, gamma_or_null, beta_or_null) &&
               (reinterpret_cast<uintptr_t>(arg) & 3u) == 0, "dc_segmax_bn_fwd: needs C %% 4 == 0 and 16-byte aligned rows");
    if (N <= 0 || C <= 0) return DC_OK;
    BnAct bn = {mean, invstd, gamma_or_null, beta_or_null, slope};
    segmax_bn_fwd_vec4_kernel<<<elementwise_grid((long long)N * C / 4), 256, 0, (cudaStream_t)stream>>>(y, N, k, C / 4, idx_or_null, out, ldo, arg, bn);
    return dc_check_launch("dc_segmax_bn_fwd");
}

// Backward of dc_segmax_bn_fwd: dout [N,C] (gradient of the max) -> dy [N*k, C] (gradient of the pre-activations), d_gamma,
// d_beta.  training != 0: batch statistics over the N*k rows (the mean terms reach every row); 0: running statistics.
// Workspace: dc_colreduce_ws_bytes(C).
extern "C" int dc_segmax_bn_bwd(const float* y, int N, int k, int C, const uint8_t* arg, const float* dout, int64_t ldd, const float* mean,
                                const float* invstd, const float* gamma_or_null, const float* beta_or_null, float slope, int training,
                                float* d_gamma, float* d_beta, float* dy, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(k >= 1 && k <= 255, "dc_segmax_bn_bwd: k must be in [1,255]");
    DC_REQUIRE(C % 4 == 0 && ldd % 4 == 0 && all_aligned16(y, dout, dy, mean, invstd, gamma_or_null, beta_or_null) &&
               all_aligned16(d_gamma, d_beta) && (reinterpret_cast<uintptr_t>(arg) & 3u) == 0,
               "dc_segmax_bn_bwd: needs C %% 4 == 0 and 16-byte aligned rows");
    DC_REQUIRE(ws_bytes >= dc_colreduce_ws_bytes(C), "dc_segmax_bn_bwd: workspace too small");
    if (N <= 0 || C <= 0) return DC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    double* partial = (double*)ws;
    SegBnBwdVal f = {y, k, C, arg, dout, ldd, mean, invstd, gamma_or_null, beta_or_null, slope};
    const int chunks = col_reduce_launch<2>(f, (long long)N, C, true, partial, st);
    sum2_finalize_kernel<<<dc_ceil_div(C, FIN_COLS), FIN_T, 0, st>>>(partial, chunks, C, d_beta, d_gamma);
    BnAct bn = {mean, invstd, gamma_or_null, beta_or_null, slope};
    segmax_bn_bwd_vec4_kernel<<<elementwise_grid((long long)N * C / 4), 256, 0, st>>>(y, dout, ldd, N, k, C / 4, arg, bn, d_beta, d_gamma, training, dy);
    return dc_check_launch_n("dc_segmax_bn_bwd", 3);
}

extern "C" int dc_pool_fwd(const float* x, int64_t ldx, const int32_t* cloud_ptr, int B, int C, float* out_max_or_null,
                           int64_t ld_max, float* out_mean_or_null, int64_t ld_mean, int32_t* arg, void* stream) {
    if (B <= 0 || C <= 0) return DC_OK;
    dim3 grid(B, dc_ceil_div(C, 32));
    pool_fwd_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(x, ldx, cloud_ptr, C, out_max_or_null, ld_max, out_mean_or_null, ld_mean, arg, BnAct{});
    return dc_check_launch("dc_pool_fwd");
}

extern "C" int dc_bn_act_pool_fwd(const float* y, int64_t ldy, const int32_t* cloud_ptr, int B, int C, const float* mean, const float* invstd,
                                  const float* gamma_or_null, const float* beta_or_null, float slope, float* out_max_or_null, int64_t ld_max,
                                  float* out_mean_or_null, int64_t ld_mean, int32_t* arg, void* stream) {
    DC_REQUIRE(mean != nullptr && invstd != nullptr, "dc_bn_act_pool_fwd: statistics missing");
    if (B <= 0 || C <= 0) return DC_OK;
    dim3 grid(B, dc_ceil_div(C, 32));
    BnAct bn = {mean, invstd, gamma_or_null, beta_or_null, slope};
    pool_fwd_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(y, ldy, cloud_ptr, C, out_max_or_null, ld_max, out_mean_or_null, ld_mean, arg, bn);
    return dc_check_launch("dc_bn_act_pool_fwd");
}

// Backward of pool([max|mean]) o leaky_relu o BatchNorm in two passes over y: the pooled gradient of row r is rebuilt from the
// [B,C] gradients (the argmax row receives d_max, every row d_mean / count), so neither it nor the activation is ever stored.
extern "C" int dc_bn_act_pool_bwd(const float* y, int64_t ldy, int64_t M, int C, const int32_t* cloud_ptr, int B,
                                  const int32_t* arg, const float* d_max_or_null, int64_t ld_dmax, const float* d_mean_or_null, int64_t ld_dmean,
                                  const float* mean, const float* invstd, const float* gamma_or_null, const float* beta_or_null, float slope,
                                  float* d_gamma, float* d_beta, float* dy, int64_t ld_dy, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(M >= 1 && C >= 1 && B >= 1, "dc_bn_act_pool_bwd: empty input");
    DC_REQUIRE(M <= 0x7fffffffLL, "dc_bn_act_pool_bwd: too many rows");
    DC_REQUIRE(d_max_or_null == nullptr || arg != nullptr, "dc_bn_act_pool_bwd: d_max without the argmax rows");
    DC_REQUIRE(C % 4 == 0 && ldy % 4 == 0 && ld_dy % 4 == 0 && ld_dmax % 4 == 0 && ld_dmean % 4 == 0 &&
               all_aligned16(y, dy, mean, invstd, gamma_or_null, beta_or_null) && all_aligned16(d_gamma, d_beta, d_max_or_null, d_mean_or_null, arg),
               "dc_bn_act_pool_bwd: needs C %% 4 == 0 and 16-byte aligned rows");
    DC_REQUIRE(ws_bytes >= dc_colreduce_ws_bytes(C), "dc_bn_act_pool_bwd: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int col_blocks = dc_ceil_div(C, 128);
    // about four CTAs per SM in total, at least 32 rows per thread
    int splits = dc_ceil_div((long long)dc_sm_count() * 4, (long long)col_blocks * B);
    const int by_rows = (int)dc_ceil_div(dc_ceil_div(M, B), 8 * 32);
    if (splits > by_rows) splits = by_rows;
    if (splits < 1) splits = 1;
    long long units = (long long)B * splits;
    const int chunks = (int)(units < RED_MAX_CHUNKS ? units : RED_MAX_CHUNKS);
    double* partial = (double*)ws;
    PoolGrad pg = {cloud_ptr, arg, d_max_or_null, ld_dmax, d_mean_or_null, ld_dmean, C};
    pool_bn_bwd_reduce_kernel<<<dim3(chunks, col_blocks), 256, 0, st>>>(y, ldy, pg, B, splits, C, mean, invstd, gamma_or_null, beta_or_null, slope, partial);
    sum2_finalize_kernel<<<dc_ceil_div(C, FIN_COLS), FIN_T, 0, st>>>(partial, chunks, C, d_beta, d_gamma);
    DC_REQUIRE(B <= 65535, "dc_bn_act_pool_bwd: more than 65535 clouds");
    pool_bn_bwd_apply_kernel<<<dim3(splits, col_blocks, B), 256, 0, st>>>(y, ldy, pg, M, C, mean, invstd, gamma_or_null, beta_or_null, slope,
                                                                        d_beta, d_gamma, dy, ld_dy);
    return dc_check_launch_n("dc_bn_act_pool_bwd", 3);
}

extern "C" int dc_pool_bwd(const int32_t* cloud_ptr, int B, int N, int C, const int32_t* arg, const float* d_max_or_null,
                           int64_t ld_dmax, const float* d_mean_or_null, int64_t ld_dmean, float* dx, int64_t ld_dx, void* stream) {
    if (N <= 0 || C <= 0) return DC_OK;
    {
        int per_cloud = dc_ceil_div((long long)dc_sm_count() * 4, B);
        int by_rows = dc_ceil_div(dc_ceil_div(N, B), 8 * 4);
        if (per_cloud > by_rows) per_cloud = by_rows;
        if (per_cloud < 1) per_cloud = 1;
        dim3 grid(per_cloud, B);
        pool_bwd_cloud_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(cloud_ptr, C, arg, d_max_or_null, ld_dmax, d_mean_or_null, ld_dmean, dx, ld_dx);
    }
    return dc_check_launch("dc_pool_bwd");
}

extern "C" int dc_sgd_step(float* params, const float* grads, float* momentum_buf, int64_t n, const float* lr_device,
                           float momentum, float weight_decay, float grad_scale, void* stream) {
    DC_REQUIRE(lr_device != nullptr, "dc_sgd_step: the learning rate is read from device memory");
    DC_REQUIRE(all_aligned16(params, grads, momentum_buf), "dc_sgd_step: buffers must be 16-byte aligned");
    if (n <= 0) return DC_OK;
    sgd_step_kernel<<<elementwise_grid((n + 3) / 4), 256, 0, (cudaStream_t)stream>>>(params, grads, momentum_buf, n, lr_device, momentum, weight_decay, grad_scale);
    return dc_check_launch("dc_sgd_step");
}

extern "C" int dc_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, const float* lr_device,
                            int32_t* step_device, float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                            void* stream) {
    DC_REQUIRE(lr_device != nullptr && step_device != nullptr, "dc_adam_step: learning rate and step counter are read from device memory");
    if (n <= 0) return DC_OK;
    adam_step_kernel<<<elementwise_grid(n), 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, n, lr_device, step_device,
                                                                           beta1, beta2, eps, weight_decay, grad_scale);
    adam_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_device);
    return dc_check_launch_n("dc_adam_step", 2);
}

extern "C" int dc_add_rows(const float* a, int64_t lda, const float* b, int64_t ldb, int64_t M, int C, float* out, int64_t ldo, void* stream) {
    if (M <= 0 || C <= 0) return DC_OK;
    const int vec = (C % 4 == 0 && lda % 4 == 0 && ldb % 4 == 0 && ldo % 4 == 0 && all_aligned16(a, b, out)) ? 1 : 0;
    add_rows_kernel<<<elementwise_grid(M * (vec ? C / 4 : C)), 256, 0, (cudaStream_t)stream>>>(a, lda, b, ldb, M, C, vec, out, ldo);
    return dc_check_launch("dc_add_rows");
}

extern "C" int dc_bcast_rows(const float* src, int64_t lds, const int32_t* cloud_ptr, int B, int N, int C, float* out, int64_t ldo,
                             void* stream) {
    if (N <= 0 || C <= 0) return DC_OK;
    if (C % 4 == 0 && lds % 4 == 0 && ldo % 4 == 0 && all_aligned16(src, out))
        bcast_rows_vec4_kernel<<<elementwise_grid((long long)N * C / 4), 256, 0, (cudaStream_t)stream>>>(src, lds, cloud_ptr, B, N, C / 4, out, ldo);
    else
        bcast_rows_kernel<<<elementwise_grid((long long)N * C), 256, 0, (cudaStream_t)stream>>>(src, lds, cloud_ptr, B, N, C, out, ldo);
    return dc_check_launch("dc_bcast_rows");
}

// out[i, :] += src[cloud of i, :]   (out [N,C] with row pitch ldo, src [B,C] with row pitch lds)
extern "C" int dc_bcast_rows_add(const float* src, int64_t lds, const int32_t* cloud_ptr, int B, int N, int C, float* out, int64_t ldo,
                                 void* stream) {
    if (N <= 0 || C <= 0) return DC_OK;
    if (C % 4 == 0 && lds % 4 == 0 && ldo % 4 == 0 && all_aligned16(src, out))
        bcast_rows_add_vec4_kernel<<<elementwise_grid((long long)N * C / 4), 256, 0, (cudaStream_t)stream>>>(src, lds, cloud_ptr, B, N, C / 4, out, ldo);
    else
        bcast_rows_add_kernel<<<elementwise_grid((long long)N * C), 256, 0, (cudaStream_t)stream>>>(src, lds, cloud_ptr, B, N, C, out, ldo);
    return dc_check_launch("dc_bcast_rows_add");
}
