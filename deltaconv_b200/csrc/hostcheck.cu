// TEST INFRASTRUCTURE (CPU only): runs the per-point math of geom_core.cuh on the host so the
// numerics of the fit and of the frame estimation can be checked against the oracle in the
// build container, which has no GPU.  Never loaded by the product.
#include "geom_core.cuh"
#include <vector>

void dc_set_error(const char*, ...) {}
int dc_check_launch(const char*) { return 0; }
int dc_check_launch_n(const char*, int) { return 0; }
int dc_sm_count() { return 0; }

extern "C" void hc_knn_dist2(const float* q, const float* p, int n, float* out) {
    for (int i = 0; i < n; ++i) out[i] = knn_dist2(q[0], q[1], q[2], p[3 * i], p[3 * i + 1], p[3 * i + 2]);
}

extern "C" void hc_tangent_basis(const float* normal, int N, float* xb, float* yb) {
    for (int i = 0; i < N; ++i) {
        V3 ex, ey;
        tangent_basis(v3_load(normal + 3 * i), ex, ey);
        xb[3 * i] = ex.x; xb[3 * i + 1] = ex.y; xb[3 * i + 2] = ex.z;
        yb[3 * i] = ey.x; yb[3 * i + 1] = ey.y; yb[3 * i + 2] = ey.z;
    }
}

extern "C" void hc_estimate_basis(const float* pos, const int* idx, int N, int k, const float* orient,
                                  float* normal, float* xb, float* yb) {
    for (int i = 0; i < N; ++i) {
        V3 p = v3_load(pos + 3 * i);
        double cov[6] = {0, 0, 0, 0, 0, 0};
        for (int e = 0; e < k; ++e) {
            int j = idx[(size_t)i * k + e];
            if (j < 0 || j >= N) j = 0;
            V3 d = v3_load(pos + 3 * j) - p;
            cov[0] += (double)d.x * d.x; cov[1] += (double)d.x * d.y; cov[2] += (double)d.x * d.z;
            cov[3] += (double)d.y * d.y; cov[4] += (double)d.y * d.z; cov[5] += (double)d.z * d.z;
        }
        V3 n, ex, ey;
        frame_from_scatter(cov, orient ? orient + 3 * i : nullptr, n, ex, ey);
        normal[3 * i] = n.x; normal[3 * i + 1] = n.y; normal[3 * i + 2] = n.z;
        xb[3 * i] = ex.x; xb[3 * i + 1] = ex.y; xb[3 * i + 2] = ex.z;
        yb[3 * i] = ey.x; yb[3 * i + 1] = ey.y; yb[3 * i + 2] = ey.z;
    }
}

// Mirrors dc_build_grad_div's kernel sequence on the host (same passes, same per-point code): means, one fit pass that
// writes un-normalised entries and the per-cloud maxima, then the per-cloud scale.
extern "C" void hc_build_grad_div(const float* pos, const float* normal, const float* xb, const float* yb,
                                  const int* idx, const int* cloud_ptr, int B, int N, int k, float kernel_width,
                                  float lam, float lam_shape, int use_shape, int normalized, float* G, float* D,
                                  float* dbg_coords, float* dbg_weights, float* dbg_mapping) {
    std::vector<float> m(N), abar(B), S(B, 0.f);
    for (int i = 0; i < N; ++i) {
        V3 p = v3_load(pos + 3 * i);
        float acc = 0.f;
        for (int e = 0; e < k; ++e) {
            int j = idx[(size_t)i * k + e];
            if (j < 0 || j >= N) j = 0;
            V3 d = v3_load(pos + 3 * j) - p;
            float d2 = dot(d, d);
            acc += d2 > 0.f ? d2 * dc_rsqrt(d2) : 0.f;
        }
        m[i] = acc / (float)k;
    }
    for (int c = 0; c < B; ++c) {
        double acc = 0;
        for (int i = cloud_ptr[c]; i < cloud_ptr[c + 1]; ++i) acc += m[i];
        int cnt = cloud_ptr[c + 1] - cloud_ptr[c];
        abar[c] = (float)(acc / (cnt > 0 ? cnt : 1));
    }
    FitParams fp; fp.lam = lam; fp.lam_shape = lam_shape; fp.use_shape = use_shape;
    std::vector<float4> sdat(k);
    int c = 0;
    for (int i = 0; i < N; ++i) {
        while (cloud_ptr[c + 1] <= i) ++c;
        V3 p = v3_load(pos + 3 * i), n = v3_load(normal + 3 * i), ex = v3_load(xb + 3 * i), ey = v3_load(yb + 3 * i);
        float sc = kernel_width * abar[c];
        float nl2 = -1.4426950408889634f / (sc * sc);
        FitSums sums;
        fit_sums_zero(sums);
        for (int e = 0; e < k; ++e) {
            int j = idx[(size_t)i * k + e];
            bool valid = !(j < 0 || j >= N);
            sdat[e] = fit_slot_accumulate(v3_load(pos + 3 * (valid ? j : 0)), valid, p, n, ex, ey, nl2, sums);
        }
        float z1[6], z2[6], coef[6];
        float inv_wnorm = fit_solve(sums, fp, z1, z2, coef);
        float ax = 0.f, ay = 0.f;
        for (int e = 0; e < k; ++e) {
            int j = idx[(size_t)i * k + e];
            int jj = (j < 0 || j >= N) ? 0 : j;
            float4 v = fit_slot_entry(v3_load(xb + 3 * jj), v3_load(yb + 3 * jj), sdat[e], n, ex, ey, z1, z2, coef, inv_wnorm,
                                      dbg_coords ? dbg_coords + ((size_t)i * k + e) * 2 : nullptr,
                                      dbg_weights ? dbg_weights + (size_t)i * k + e : nullptr,
                                      dbg_mapping ? dbg_mapping + ((size_t)i * k + e) * 4 : nullptr);
            ax += fabsf(v.x); ay += fabsf(v.y);
            G[((size_t)i * k + e) * 2] = v.x; G[((size_t)i * k + e) * 2 + 1] = v.y;
            D[((size_t)i * k + e) * 2] = v.z; D[((size_t)i * k + e) * 2 + 1] = v.w;
        }
        float s = fit_row_norm(ax, ay);
        if (s > S[c]) S[c] = s;
    }
    if (!normalized) return;
    c = 0;
    for (int i = 0; i < N; ++i) {
        while (cloud_ptr[c + 1] <= i) ++c;
        float f = fit_scale(S[c]);
        for (int e = 0; e < 2 * k; ++e) { G[(size_t)i * k * 2 + e] *= f; D[(size_t)i * k * 2 + e] *= f; }
    }
}
