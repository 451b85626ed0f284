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
// One pass per point (grad_div_mls.py:72-194, :230-272), k slots e with neighbour j = idx[i,e]:
//   A  (per slot)  d = p_j - p_i -> tangent coordinates (u, v), height h, raw gaussian weight w; accumulate the 15 distinct
//                  moments sum w u^a v^b (a + b <= 4) the 6x6 normal matrix consists of, and the surface right-hand side
//                  sum w h b_q; keep (u, v, w) per slot
//   solve          scale the sums by 1 / max(sum w, 1e-5) (the weight normalisation, :113-114, is linear in the sums), add
//                  lambda, Cholesky with one reciprocal square root per pivot, rows 1 and 2 of the inverse (the only rows the
//                  gradient needs, :255) and the quadric coefficients of the surface
//   B  (per slot)  gradient entry g = w (z . b), frame mapping M (:168-194), divergence entry g M (:271-272) — NOT yet
//                  divided by the cloud's infinity norm S_b (:258-260): D = g M is linear in g, so the one scalar per cloud
//                  is applied afterwards (fit_scale), which is what lets the fit run once instead of twice.
// 6x6 SPD system in packed lower-triangular storage: index(r,c) = r(r+1)/2 + c, c <= r.
DC_HD int tri(int r, int c) { return r * (r + 1) / 2 + c; }

// 1/sqrt(x): hardware approximation + one Newton step on the device (about 1 ulp), exact division on the host
DC_HD float dc_rsqrt(float x) {
#if defined(__CUDA_ARCH__)
    float y = rsqrtf(x);
    float h = 0.5f * x * y;
    return fmaf(y, fmaf(-h, y, 0.5f), y);
#else
    return 1.0f / sqrtf(x);
#endif
}
DC_HD float dc_rcp(float x) {          // x in [1, 1e30]: approximation + one Newton step, no special-case branches
#if defined(__CUDA_ARCH__)
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return fmaf(r, fmaf(-x, r, 1.f), r);
#else
    return 1.0f / x;
#endif
}
DC_HD float dc_exp2(float x) {
#if defined(__CUDA_ARCH__)
    return exp2f(x);
#else
    return exp2f(x);
#endif
}

// In-place Cholesky A = L L^T with the reciprocals of the pivots (6 reciprocal square roots instead of 15 divisions and 6
// square roots).  Pivots are floored so that lambda -> 0 never yields NaN.
DC_HD void chol6(float L[21], float inv[6]) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
        for (int j = 0; j <= i; ++j) {
            float s = L[tri(i, j)];
#pragma unroll
            for (int t = 0; t < j; ++t) s -= L[tri(i, t)] * L[tri(j, t)];
            if (i == j) { float d = fmaxf(s, 1e-30f); inv[i] = dc_rsqrt(d); L[tri(i, i)] = d * inv[i]; }
            else L[tri(i, j)] = s * inv[j];
        }
    }
}

// x <- A^{-1} x given the Cholesky factor; entries of x below FIRST are known to be zero (unit right-hand sides)
template <int FIRST>
DC_HD void chol6_solve(const float L[21], const float inv[6], float x[6]) {
#pragma unroll
    for (int i = FIRST; i < 6; ++i) {
        float s = x[i];
#pragma unroll
        for (int t = FIRST; t < i; ++t) s -= L[tri(i, t)] * x[t];
        x[i] = s * inv[i];
    }
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        float s = x[i];
#pragma unroll
        for (int t = i + 1; t < 6; ++t) s -= L[tri(t, i)] * x[t];
        x[i] = s * inv[i];
    }
}

struct FitParams {
    float lam;          // regulariser of the signal fit (grad_div_mls.py:140)
    float lam_shape;    // regulariser of the surface fit, used when use_shape != 0 (:146-150, :266-267)
    int use_shape;
};

// sums over the slots of one point
struct FitSums {
    float m[15];        // sum w u^a v^b in the order 1, u, v, uu, uv, vv, uuu, uuv, uvv, vvv, u4, u3v, u2v2, uv3, v4
    float r[6];         // sum w h b_q, b = [1, u, v, uu, uv, vv]  (grad_div_mls.py:133-137, :164-165)
    float wsum;         // sum of the raw weights of ALL slots (gaussian_weights normalises before padded slots are masked, :113-114, :235)
};
DC_HD void fit_sums_zero(FitSums& s) {
#pragma unroll
    for (int q = 0; q < 15; ++q) s.m[q] = 0.f;
#pragma unroll
    for (int q = 0; q < 6; ++q) s.r[q] = 0.f;
    s.wsum = 0.f;
}

// Pass A of ONE slot (grad_div_mls.py:91-95, :233): returns (u, v, w, h) with w = 0 for a padded slot and adds the slot to
// the sums.  `neg_inv_denom_log2e` = -log2(e) / (kernel_width * mean edge length of the cloud)^2.  A padded slot (neighbour
// index N, :226-227) is evaluated on neighbour 0 like the reference does: its raw weight counts in the normalisation only.
DC_HD float4 fit_slot_accumulate(V3 pj, bool valid, V3 p, V3 n, V3 ex, V3 ey, float neg_inv_denom_log2e, FitSums& s) {
    V3 d = pj - p;
    float h = dot(n, d);
    V3 t = d - n * h;
    float u = dot(t, ex), v = dot(t, ey);
    float wraw = dc_exp2(dot(d, d) * neg_inv_denom_log2e);
    s.wsum += wraw;
    float w = valid ? wraw : 0.f;
    float uu = u * u, uv = u * v, vv = v * v;
    s.m[0] += w;
    s.m[1] = fmaf(w, u, s.m[1]); s.m[2] = fmaf(w, v, s.m[2]);
    s.m[3] = fmaf(w, uu, s.m[3]); s.m[4] = fmaf(w, uv, s.m[4]); s.m[5] = fmaf(w, vv, s.m[5]);
    s.m[6] = fmaf(w, uu * u, s.m[6]); s.m[7] = fmaf(w, uu * v, s.m[7]); s.m[8] = fmaf(w, uv * v, s.m[8]); s.m[9] = fmaf(w, vv * v, s.m[9]);
    s.m[10] = fmaf(w, uu * uu, s.m[10]); s.m[11] = fmaf(w, uu * uv, s.m[11]); s.m[12] = fmaf(w, uu * vv, s.m[12]);
    s.m[13] = fmaf(w, uv * vv, s.m[13]); s.m[14] = fmaf(w, vv * vv, s.m[14]);
    float wh = w * h;
    s.r[0] += wh;
    s.r[1] = fmaf(wh, u, s.r[1]); s.r[2] = fmaf(wh, v, s.r[2]);
    s.r[3] = fmaf(wh, uu, s.r[3]); s.r[4] = fmaf(wh, uv, s.r[4]); s.r[5] = fmaf(wh, vv, s.r[5]);
    return make_float4(u, v, w, h);
}

// Solve (grad_div_mls.py:113-114, :139-150, :164-165).  z1, z2: rows 1 and 2 of (B^T W B + lambda I)^-1 ALREADY multiplied by
// the weight normalisation 1 / max(sum w, 1e-5), so that g_x[e] = w_raw[e] * (z1 . b[e]); coef: the quadric coefficients of
// the surface.  Returns the weight normalisation.
DC_HD float fit_solve(const FitSums& s, const FitParams& fp, float z1[6], float z2[6], float coef[6]) {
    const float inv_wnorm = 1.f / fmaxf(s.wsum, DC_GEOM_EPS);
    float m[15];
#pragma unroll
    for (int q = 0; q < 15; ++q) m[q] = s.m[q] * inv_wnorm;
    // A[r][q] = sum w b_r b_q with b = [1, u, v, uu, uv, vv]: monomial of (r, q) -> index into m
    float A[21];
    A[tri(0, 0)] = m[0];
    A[tri(1, 0)] = m[1]; A[tri(1, 1)] = m[3];
    A[tri(2, 0)] = m[2]; A[tri(2, 1)] = m[4]; A[tri(2, 2)] = m[5];
    A[tri(3, 0)] = m[3]; A[tri(3, 1)] = m[6]; A[tri(3, 2)] = m[7]; A[tri(3, 3)] = m[10];
    A[tri(4, 0)] = m[4]; A[tri(4, 1)] = m[7]; A[tri(4, 2)] = m[8]; A[tri(4, 3)] = m[11]; A[tri(4, 4)] = m[12];
    A[tri(5, 0)] = m[5]; A[tri(5, 1)] = m[8]; A[tri(5, 2)] = m[9]; A[tri(5, 3)] = m[12]; A[tri(5, 4)] = m[13]; A[tri(5, 5)] = m[14];
    float L[21], inv[6];
#pragma unroll
    for (int q = 0; q < 21; ++q) L[q] = A[q];
#pragma unroll
    for (int r = 0; r < 6; ++r) L[tri(r, r)] += fp.lam;
    chol6(L, inv);
#pragma unroll
    for (int q = 0; q < 6; ++q) { z1[q] = (q == 1) ? 1.f : 0.f; z2[q] = (q == 2) ? 1.f : 0.f; coef[q] = s.r[q] * inv_wnorm; }
    chol6_solve<1>(L, inv, z1);
    chol6_solve<2>(L, inv, z2);
    if (fp.use_shape) {
#pragma unroll
        for (int q = 0; q < 21; ++q) L[q] = A[q];
#pragma unroll
        for (int r = 0; r < 6; ++r) L[tri(r, r)] += fp.lam_shape;
        chol6(L, inv);
    }
    chol6_solve<0>(L, inv, coef);
#pragma unroll
    for (int q = 0; q < 6; ++q) { z1[q] *= inv_wnorm; z2[q] *= inv_wnorm; }
    return inv_wnorm;
}

// Pass B of ONE slot: (g_x, g_y, D_0, D_1) before the division by the cloud's infinity norm; c = (u, v, w, .) from pass A.
// The frame mapping follows fit_vector_mapping (:168-194), the divergence row :271-272.  Optional debug outputs mirror the
// reference's public helpers (coords [2], normalised weights [1], mapping [4]).
DC_HD float4 fit_slot_entry(V3 xj, V3 yj, float4 c, V3 n, V3 ex, V3 ey, const float z1[6], const float z2[6], const float coef[6],
                            float inv_wnorm, float* dbg_coords, float* dbg_weights, float* dbg_mapping) {
    const float u = c.x, v = c.y, w = c.z;
    const float uu = u * u, uv = u * v, vv = v * v;
    float a1 = z1[0], a2 = z2[0];
    a1 = fmaf(z1[1], u, a1); a1 = fmaf(z1[2], v, a1); a1 = fmaf(z1[3], uu, a1); a1 = fmaf(z1[4], uv, a1); a1 = fmaf(z1[5], vv, a1);
    a2 = fmaf(z2[1], u, a2); a2 = fmaf(z2[2], v, a2); a2 = fmaf(z2[3], uu, a2); a2 = fmaf(z2[4], uv, a2); a2 = fmaf(z2[5], vv, a2);
    const float gx = w * a1, gy = w * a2;                 // w = 0 for a padded slot: all four entries vanish
    const float hx = coef[1] + 2.f * coef[3] * u + coef[4] * v;
    const float hy = coef[2] + coef[4] * u + 2.f * coef[5] * v;
    const V3 gam_x = ex + n * hx, gam_y = ey + n * hy;
    const float Eb = 1.f + hx * hx, Fb = hx * hy, Gb = 1.f + hy * hy;
    const float inv_det = dc_rcp(Eb + hy * hy);
    const float t00 = dot(gam_x, xj), t01 = dot(gam_x, yj), t10 = dot(gam_y, xj), t11 = dot(gam_y, yj);
    const float m00 = (Gb * t00 - Fb * t10) * inv_det, m01 = (Gb * t01 - Fb * t11) * inv_det;
    const float m10 = (Eb * t10 - Fb * t00) * inv_det, m11 = (Eb * t11 - Fb * t01) * inv_det;
    if (dbg_coords) { dbg_coords[0] = u; dbg_coords[1] = v; }
    if (dbg_weights) dbg_weights[0] = w * inv_wnorm;
    if (dbg_mapping) { dbg_mapping[0] = m00; dbg_mapping[1] = m01; dbg_mapping[2] = m10; dbg_mapping[3] = m11; }
    return make_float4(gx, gy, gx * m00 + gy * m10, gx * m01 + gy * m11);
}

// The per-cloud normalisation (grad_div_mls.py:258-260): s_i = || (sum_e |g_x|, sum_e |g_y|) ||_2, S_b = max over the cloud,
// entries divided by S_b when S_b > 1e-5.  fit_scale returns the factor to multiply a point's entries with.
DC_HD float fit_row_norm(float abs_gx_sum, float abs_gy_sum) { return sqrtf(abs_gx_sum * abs_gx_sum + abs_gy_sum * abs_gy_sum); }
DC_HD float fit_scale(float S) { return S > 1e-5f ? 1.f / S : 1.f; }
