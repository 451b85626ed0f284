// Test infrastructure, not product code: a C-ABI door into the REFERENCE's own geodesic farthest-point sampler so that the
// parity tests can call it.  The sampler itself (deltaconv/cpp/sampling.cpp) and geometry-central's kd-tree
// (deps/geometry-central/src/utilities/knn.cpp) are compiled from where they lie under /root/reference by oracle/Makefile;
// nothing of them is copied here.  Only tests/, smoke() and the CPU legs of bench.py may load the resulting library.
#include "sampling.h"

#include <cstring>

extern "C" {

// constructGeodesicFPS (sampling.cpp:21-53) draws its first sample from std::random_device; the start it chose is
// sample[0], which the tests hand to the CUDA path so that both continue from the same point.
int ref_geodesic_fps(const double* xyz, int n, int n_samples, int* samples) {
    std::vector<Vector3> points(n);
    for (int i = 0; i < n; ++i) points[i] = Vector3{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    Eigen::VectorXi idx = constructGeodesicFPS(points, (size_t)n_samples);
    std::memcpy(samples, idx.data(), sizeof(int) * (size_t)n_samples);
    return 0;
}

// generateKNN (sampling.cpp:5-17): k nearest other points per point, the point itself in front -> [n, k+1]
int ref_generate_knn(const double* xyz, int n, int k, long long* neigh) {
    std::vector<Vector3> points(n);
    for (int i = 0; i < n; ++i) points[i] = Vector3{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    Neighbors_t r = generateKNN(points, (size_t)k);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j <= k; ++j) neigh[(size_t)i * (k + 1) + j] = (long long)r[i][j];
    return 0;
}

// computeDijkstra (sampling.cpp:56-81) from one source into a caller-held distance vector (min-combined, as the sampler uses it)
int ref_dijkstra(const double* xyz, int n, int k, int source, double* dist) {
    std::vector<Vector3> points(n);
    for (int i = 0; i < n; ++i) points[i] = Vector3{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    Neighbors_t neigh = generateKNN(points, (size_t)k);
    Eigen::VectorXd D(n);
    for (int i = 0; i < n; ++i) D(i) = dist[i];
    computeDijkstra(points, source, neigh, D);
    for (int i = 0; i < n; ++i) dist[i] = D(i);
    return 0;
}

}
