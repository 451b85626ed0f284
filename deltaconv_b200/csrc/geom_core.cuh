// Per-point geometry math shared by the CUDA kernels and by the host check library
// (csrc/hostcheck.cu, CPU-only test infrastructure).  Everything here is __host__ __device__.
//
// Reference arithmetic being reproduced (paths relative to /root/reference):
//   deltaconv/geometry/grad_div_mls.py:10-47   estimate_basis
//   deltaconv/geometry/grad_div_mls.py:50-69   build_tangent_basis
//   deltaconv/geometry/grad_div_mls.py:72-277  coords_projected, gaussian_weights,
//                                              weighted_least_squares, fit_vector_mapping, build_grad_div
#pragma once
#include "common.cuh"
#include <math.h>

#define DC_GEOM_EPS 1e-5f   // grad_div_mls.py:7

struct V3 { float x, y, z; };
DC_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
DC_HD V3 v3_load(const float* p) { return v3(p[0], p[1], p[2]); }
DC_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
DC_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
DC_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
DC_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
DC_HD V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
DC_HD V3 unit(V3 a, float floor_) { float n = sqrtf(dot(a, a)); n = n < floor_ ? floor_ : n; return v3(a.x / n, a.y / n, a.z / n); }

// ---- kNN distance: the formula the parity suite pins: every product and sum
// rounded separately, no FMA contraction, (dx*dx + dy*dy) + dz*dz.
DC_HD float knn_dist2(float qx, float qy, float qz, float px, float py, float pz) {
#if defined(__CUDA_ARCH__)
    float dx = __fsub_rn(px, qx), dy = __fsub_rn(py, qy), dz = __fsub_rn(pz, qz);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
#else
    volatile float dx = px - qx, dy = py - qy, dz = pz - qz;
    volatile float a = dx * dx, b = dy * dy, c = dz * dz;
    volatile float ab = a + b;
    return ab + c;
#endif
}

// ---- build_tangent_basis (grad_div_mls.py:58-68)
DC_HD void tangent_basis(V3 n, V3& xb, V3& yb) {
    V3 probe = fabsf(n.x) > 0.9f ? v3(0.f, 1.f, 0.f) : v3(1.f, 0.f, 0.f);
    xb = unit(cross(probe, n), DC_GEOM_EPS);
    yb = unit(cross(n, xb), DC_GEOM_EPS);
}

// ---- estimate_basis (grad_div_mls.py:26-45).  The reference takes U of the SVD of the 3 x k
// offset matrix; the columns of U are the eigenvectors of the 3x3 scatter matrix S = sum d d^T,
// found here by cyclic Jacobi in fp64 (fp64 on S resolves the same singular-value ratios fp32
// SVD on the offsets does).  cov = {xx, xy, xz, yy, yz, zz}.  Signs of the axes are a gauge.
DC_HD void sym3_eigen(const double cov[6], double evec[3][3], double eval[3]) {
    double a[3][3] = {{cov[0], cov[1], cov[2]}, {cov[1], cov[3], cov[4]}, {cov[2], cov[4], cov[5]}};
    double v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int sweep = 0; sweep < 12; ++sweep) {
        double off = a[0][1] * a[0][1] + a[0][2] * a[0][2] + a[1][2] * a[1][2];
        double diag = a[0][0] * a[0][0] + a[1][1] * a[1][1] + a[2][2] * a[2][2];
        if (off <= 1e-32 * diag || off == 0.0) break;
#pragma unroll
        for (int pq = 0; pq < 3; ++pq) {
            const int p = pq == 2 ? 1 : 0, q = pq == 0 ? 1 : 2;
            double apq = a[p][q];
            if (apq == 0.0) continue;
            double theta = (a[q][q] - a[p][p]) / (2.0 * apq);
            double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
#pragma unroll
            for (int r = 0; r < 3; ++r) {       // A <- A J
                double arp = a[r][p], arq = a[r][q];
                a[r][p] = c * arp - s * arq;
                a[r][q] = s * arp + c * arq;
            }
#pragma unroll
            for (int r = 0; r < 3; ++r) {       // A <- J^T A
                double apr = a[p][r], aqr = a[q][r];
                a[p][r] = c * apr - s * aqr;
                a[q][r] = s * apr + c * aqr;
            }
#pragma unroll
            for (int r = 0; r < 3; ++r) {       // V <- V J
                double vrp = v[r][p], vrq = v[r][q];
                v[r][p] = c * vrp - s * vrq;
                v[r][q] = s * vrp + c * vrq;
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        eval[i] = a[i][i];
#pragma unroll
        for (int r = 0; r < 3; ++r) evec[i][r] = v[r][i];
    }
}

DC_HD void frame_from_scatter(const double cov[6], const float* orient /*nullable*/, V3& n, V3& xb, V3& yb) {
    double evec[3][3], eval[3];
    sym3_eigen(cov, evec, eval);
    int lo = 0, hi = 0;
#pragma unroll
    for (int i = 1; i < 3; ++i) {
        if (eval[i] < eval[lo]) lo = i;
        if (eval[i] > eval[hi]) hi = i;
    }
    if (lo == hi) { lo = 2; hi = 0; }          // fully degenerate neighbourhood
    n = unit(v3((float)evec[lo][0], (float)evec[lo][1], (float)evec[lo][2]), DC_GEOM_EPS);
    if (orient != nullptr && dot(n, v3_load(orient)) < 0.f) n = n * -1.f;
    xb = unit(v3((float)evec[hi][0], (float)evec[hi][1], (float)evec[hi][2]), DC_GEOM_EPS);
    yb = unit(cross(n, xb), DC_GEOM_EPS);
}

// ---- WLS fit ------------------------------------------------------------------------------
// 6x6 SPD system in packed lower-triangular storage: index(r,c) = r(r+1)/2 + c, c <= r.
DC_HD int tri(int r, int c) { return r * (r + 1) / 2 + c; }

// In-place Cholesky A = L L^T.  Pivots are floored so that lambda -> 0 never yields NaN.
DC_HD void chol6(float L[21]) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
        for (int j = 0; j <= i; ++j) {
            float s = L[tri(i, j)];
#pragma unroll
            for (int t = 0; t < j; ++t) s -= L[tri(i, t)] * L[tri(j, t)];
            if (i == j) L[tri(i, i)] = sqrtf(fmaxf(s, 1e-30f));
            else L[tri(i, j)] = s / L[tri(j, j)];
        }
    }
}

// x <- A^{-1} x given the Cholesky factor
DC_HD void chol6_solve(const float L[21], float x[6]) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        float s = x[i];
#pragma unroll
        for (int t = 0; t < i; ++t) s -= L[tri(i, t)] * x[t];
        x[i] = s / L[tri(i, i)];
    }
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        float s = x[i];
#pragma unroll
        for (int t = i + 1; t < 6; ++t) s -= L[tri(t, i)] * x[t];
        x[i] = s / L[tri(i, i)];
    }
}

DC_HD void quadric_basis(float u, float v, float b[6]) {   // grad_div_mls.py:133-137
    b[0] = 1.f; b[1] = u; b[2] = v; b[3] = u * u; b[4] = u * v; b[5] = v * v;
}

struct FitParams {
    float lam;          // regulariser of the signal fit (grad_div_mls.py:140)
    float lam_shape;    // regulariser of the surface fit, used when use_shape != 0 (:146-150, :266-267)
    int use_shape;
};

// Slot storage of one point: element e lives at [e * stride] (slot-major shared memory on the
// device, stride 1 on the host).  dat holds (u, v, w, height) after pass 1 and
// (g_x, g_y, D_0, D_1) after pass 3.
struct SlotView {
    int* idx;
    float4* dat;
    int stride;
    DC_HD int& j(int e) const { return idx[(size_t)e * stride]; }
    DC_HD float4& d(int e) const { return dat[(size_t)e * stride]; }
};

// Pass 1 (grad_div_mls.py:91-95, :233): tangent-plane coordinates, height over the plane, raw
// gaussian weight of ONE slot.  `denom` = (kernel_width * mean edge length of the cloud)^2.  Neighbour
// index N is the padding sentinel (:226-227): treated as neighbour 0 with final weight 0.
DC_HD float4 fit_slot_coords_loaded(V3 pj, V3 p, V3 n, V3 ex, V3 ey, float denom) {
    V3 d = pj - p;
    float h = dot(n, d);
    V3 t = d - n * h;
    float r = sqrtf(dot(d, d));
    float w = expf(-(r * r) / denom);
    return make_float4(dot(t, ex), dot(t, ey), w, h);
}
DC_HD float4 fit_slot_coords(const float* __restrict__ pos, int N, int j, V3 p, V3 n, V3 ex, V3 ey, float denom) {
    int jj = (j < 0 || j >= N) ? 0 : j;
    return fit_slot_coords_loaded(v3_load(pos + 3 * (size_t)jj), p, n, ex, ey, denom);
}

DC_HD float fit_pass1(const float* __restrict__ pos, int N, int i, int k, V3 p, V3 n, V3 ex, V3 ey,
                      float denom, const SlotView& s) {
    float wsum = 0.f;
    for (int e = 0; e < k; ++e) {
        float4 c = fit_slot_coords(pos, N, s.j(e), p, n, ex, ey, denom);
        wsum += c.z;
        s.d(e) = c;
    }
    return wsum;
}

// Pass 2 (grad_div_mls.py:113-114, :139-144, :164-165): normalise weights, assemble the normal
// equations and the surface right-hand side, factor, and solve for rows 1 and 2 of the inverse
// (the only rows the gradient needs, :255) and for the quadric coefficients of the surface.
DC_HD void fit_pass2(int N, int k, float wsum, const FitParams& fp, const SlotView& s,
                     float z1[6], float z2[6], float coef[6]) {
    float A[21], rhs[6];
#pragma unroll
    for (int q = 0; q < 21; ++q) A[q] = 0.f;
#pragma unroll
    for (int q = 0; q < 6; ++q) rhs[q] = 0.f;
    const float inv_wnorm = 1.f / fmaxf(wsum, DC_GEOM_EPS);
    for (int e = 0; e < k; ++e) {
        float4 c = s.d(e);
        int j = s.j(e);
        float w = (j < 0 || j >= N) ? 0.f : c.z * inv_wnorm;
        c.z = w;
        s.d(e) = c;
        float b[6];
        quadric_basis(c.x, c.y, b);
        float wh = w * c.w;
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            float wb = w * b[r];
#pragma unroll
            for (int q = 0; q <= r; ++q) A[tri(r, q)] += wb * b[q];
            rhs[r] += wh * b[r];
        }
    }
    float L[21];
#pragma unroll
    for (int q = 0; q < 21; ++q) L[q] = A[q];
#pragma unroll
    for (int r = 0; r < 6; ++r) L[tri(r, r)] += fp.lam;
    chol6(L);
#pragma unroll
    for (int q = 0; q < 6; ++q) { z1[q] = (q == 1) ? 1.f : 0.f; z2[q] = (q == 2) ? 1.f : 0.f; coef[q] = rhs[q]; }
    chol6_solve(L, z1);
    chol6_solve(L, z2);
    if (fp.use_shape) {
#pragma unroll
        for (int q = 0; q < 21; ++q) L[q] = A[q];
#pragma unroll
        for (int r = 0; r < 6; ++r) L[tri(r, r)] += fp.lam_shape;
        chol6(L);
    }
    chol6_solve(L, coef);
}

// gradient entries of slot e before normalisation (grad_div_mls.py:144, :255)
DC_HD void fit_grad_entry(float4 c, const float z1[6], const float z2[6], float& gx, float& gy) {
    float b[6];
    quadric_basis(c.x, c.y, b);
    float a1 = 0.f, a2 = 0.f;
#pragma unroll
    for (int q = 0; q < 6; ++q) { a1 += z1[q] * b[q]; a2 += z2[q] * b[q]; }
    gx = c.z * a1;
    gy = c.z * a2;
}

// Pass 3a: s_i = || (sum_e |g_x|, sum_e |g_y|) ||_2   (grad_div_mls.py:259)
DC_HD float fit_row_norm(int k, const SlotView& s, const float z1[6], const float z2[6]) {
    float ax = 0.f, ay = 0.f;
    for (int e = 0; e < k; ++e) {
        float gx, gy;
        fit_grad_entry(s.d(e), z1, z2, gx, gy);
        ax += fabsf(gx);
        ay += fabsf(gy);
    }
    return sqrtf(ax * ax + ay * ay);
}

// Pass 3b: final entries of ONE slot.  `S` is the cloud's infinity norm (<= 0 means "do not normalise");
// the frame mapping follows fit_vector_mapping (:168-194) and the divergence row :271-272.
// Optional debug outputs mirror the reference's public helpers (coords [2], weights [1], mapping [4]).
// Returns (g_x, g_y, D_0, D_1).
DC_HD float4 fit_slot_entry_loaded(V3 xj, V3 yj, bool valid, float4 c, V3 n, V3 ex, V3 ey, float S,
                                   const float z1[6], const float z2[6], const float coef[6],
                                   float* dbg_coords, float* dbg_weights, float* dbg_mapping) {
    float gx, gy;
    fit_grad_entry(c, z1, z2, gx, gy);
    if (S > 1e-5f) { gx = gx / S; gy = gy / S; }
    float u = c.x, v = c.y;
    float hx = coef[1] + 2.f * coef[3] * u + coef[4] * v;
    float hy = coef[2] + coef[4] * u + 2.f * coef[5] * v;
    V3 gam_x = ex + n * hx, gam_y = ey + n * hy;
    float inv_det = 1.f / (1.f + hx * hx + hy * hy);
    float Eb = 1.f + hx * hx, Fb = hx * hy, Gb = 1.f + hy * hy;
    float t00 = dot(gam_x, xj), t01 = dot(gam_x, yj), t10 = dot(gam_y, xj), t11 = dot(gam_y, yj);
    float m00 = (Gb * t00 - Fb * t10) * inv_det, m01 = (Gb * t01 - Fb * t11) * inv_det;
    float m10 = (-Fb * t00 + Eb * t10) * inv_det, m11 = (-Fb * t01 + Eb * t11) * inv_det;
    if (dbg_coords) { dbg_coords[0] = u; dbg_coords[1] = v; }
    if (dbg_weights) dbg_weights[0] = c.z;
    if (dbg_mapping) { dbg_mapping[0] = m00; dbg_mapping[1] = m01; dbg_mapping[2] = m10; dbg_mapping[3] = m11; }
    float4 out;
    out.x = valid ? gx : 0.f;
    out.y = valid ? gy : 0.f;
    out.z = valid ? gx * m00 + gy * m10 : 0.f;
    out.w = valid ? gx * m01 + gy * m11 : 0.f;
    return out;
}
DC_HD float4 fit_slot_entry(const float* __restrict__ xb_all, const float* __restrict__ yb_all, int N, int j, float4 c,
                            V3 n, V3 ex, V3 ey, float S, const float z1[6], const float z2[6], const float coef[6],
                            float* dbg_coords, float* dbg_weights, float* dbg_mapping) {
    bool valid = !(j < 0 || j >= N);
    int jj = valid ? j : 0;
    return fit_slot_entry_loaded(v3_load(xb_all + 3 * (size_t)jj), v3_load(yb_all + 3 * (size_t)jj), valid, c, n, ex, ey, S,
                                 z1, z2, coef, dbg_coords, dbg_weights, dbg_mapping);
}

DC_HD void fit_pass3(const float* __restrict__ xb_all, const float* __restrict__ yb_all, int N, int k,
                     V3 n, V3 ex, V3 ey, float S, const SlotView& s,
                     const float z1[6], const float z2[6], const float coef[6],
                     float* dbg_coords, float* dbg_weights, float* dbg_mapping) {
    for (int e = 0; e < k; ++e)
        s.d(e) = fit_slot_entry(xb_all, yb_all, N, s.j(e), s.d(e), n, ex, ey, S, z1, z2, coef,
                                dbg_coords ? dbg_coords + 2 * e : nullptr, dbg_weights ? dbg_weights + e : nullptr,
                                dbg_mapping ? dbg_mapping + 4 * e : nullptr);
}
