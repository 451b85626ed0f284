/* TEST INFRASTRUCTURE — CPU restatement of the reference's geodesic farthest-point sampler, used only as the checker of
 * the CUDA path (tests/, __graft_entry__.smoke(), bench.py's CPU legs).  The product never loads this file.
 *
 * Follows /root/reference/deltaconv/cpp/sampling.cpp:
 *   generateKNN          (:5-17)   10 nearest other points per point; the reference asks geometry-central's kd-tree
 *                                  (nanoflann, L2_Simple_Adaptor<double>: d2 = ((0 + dx*dx) + dy*dy) + dz*dz), here a
 *                                  plain scan with the same metric, ties to the lower index
 *   computeDijkstra      (:56-81)  min-heap on distance, no "visited" set, strict `<` relaxation, edge length
 *                                  sqrt(x*x + y*y + z*z) (geometry-central vector3.ipp:53)
 *   constructGeodesicFPS (:21-53)  D starts at +inf and is shared by all passes; next sample = first index of max(D)
 * The reference draws its first sample from std::random_device (:33-40); here the caller passes it.
 * Pinned against the compiled reference (oracle/_ref/libref_fps.so) by tests/test_oracle_fps.py and against the
 * committed outputs of that library in tests/golden/geodesic_fps.npz.
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC (oracle/Makefile).
 */
#include <math.h>
#include <stdlib.h>

#define FPS_K 10

typedef struct { double d; int v; } item_t;
typedef struct { item_t* a; int n, cap; } heap_t;

static void heap_push(heap_t* h, double d, int v) {
    if (h->n == h->cap) { h->cap = h->cap ? 2 * h->cap : 1024; h->a = (item_t*)realloc(h->a, sizeof(item_t) * (size_t)h->cap); }
    int i = h->n++;
    while (i > 0) {
        int p = (i - 1) / 2;
        if (h->a[p].d <= d) break;
        h->a[i] = h->a[p]; i = p;
    }
    h->a[i].d = d; h->a[i].v = v;
}

static item_t heap_pop(heap_t* h) {
    item_t top = h->a[0], last = h->a[--h->n];
    int i = 0;
    for (;;) {
        int c = 2 * i + 1;
        if (c >= h->n) break;
        if (c + 1 < h->n && h->a[c + 1].d < h->a[c].d) ++c;
        if (last.d <= h->a[c].d) break;
        h->a[i] = h->a[c]; i = c;
    }
    h->a[i] = last;
    return top;
}

/* neighbours [n, FPS_K] (index -1 where the cloud has fewer than FPS_K other points) and edge lengths */
void oracle_fps_knn(const double* xyz, int n, int* nbr, double* len) {
    for (int i = 0; i < n; ++i) {
        double bd[FPS_K]; int bi[FPS_K]; int m = 0;
        for (int j = 0; j < n; ++j) {
            if (j == i) continue;
            double dx = xyz[3 * i] - xyz[3 * j], dy = xyz[3 * i + 1] - xyz[3 * j + 1], dz = xyz[3 * i + 2] - xyz[3 * j + 2];
            double d = dx * dx + dy * dy; d = d + dz * dz;
            if (m == FPS_K && !(d < bd[FPS_K - 1])) continue;
            int t = m < FPS_K ? m++ : FPS_K - 1;
            while (t > 0 && d < bd[t - 1]) { bd[t] = bd[t - 1]; bi[t] = bi[t - 1]; --t; }
            bd[t] = d; bi[t] = j;
        }
        for (int t = 0; t < FPS_K; ++t) {
            nbr[i * FPS_K + t] = t < m ? bi[t] : -1;
            len[i * FPS_K + t] = t < m ? sqrt(bd[t]) : 0.0;
        }
    }
}

/* one pass from `source` into the shared distance vector */
void oracle_fps_dijkstra(int n, const int* nbr, const double* len, int source, double* D) {
    heap_t h = {0, 0, 0};
    (void)n;
    D[source] = 0.0;
    heap_push(&h, 0.0, source);
    while (h.n > 0) {
        item_t top = heap_pop(&h);
        for (int t = 0; t < FPS_K; ++t) {
            int v = nbr[top.v * FPS_K + t];
            if (v < 0) continue;
            double cand = top.d + len[top.v * FPS_K + t];
            if (cand < D[v]) { D[v] = cand; heap_push(&h, cand, v); }
        }
    }
    free(h.a);
}

/* samples[n_samples] from the given first sample; D_out (optional, [n]) = distances after the pass from the LAST sample */
int oracle_geodesic_fps(const double* xyz, int n, int n_samples, int start, int* samples, double* D_out) {
    if (n <= 0 || n_samples <= 0 || start < 0 || start >= n) return -1;
    int* nbr = (int*)malloc(sizeof(int) * (size_t)n * FPS_K);
    double* len = (double*)malloc(sizeof(double) * (size_t)n * FPS_K);
    double* D = (double*)malloc(sizeof(double) * (size_t)n);
    oracle_fps_knn(xyz, n, nbr, len);
    for (int i = 0; i < n; ++i) D[i] = INFINITY;
    samples[0] = start;
    for (int s = 1; s < n_samples; ++s) {
        oracle_fps_dijkstra(n, nbr, len, samples[s - 1], D);
        int arg = 0;
        for (int i = 1; i < n; ++i) if (D[i] > D[arg]) arg = i;
        samples[s] = arg;
    }
    if (D_out) {
        oracle_fps_dijkstra(n, nbr, len, samples[n_samples - 1], D);
        for (int i = 0; i < n; ++i) D_out[i] = D[i];
    }
    free(nbr); free(len); free(D);
    return 0;
}
