// K12 — geodesic farthest-point sampling of a batch of clouds (reference: deltaconv/cpp/sampling.cpp:21-81,
// the pre-transform transforms/geodesic_fps.py:14-43 runs on every shape of a dataset).
//
// The reference handles one cloud on one CPU thread: kd-tree 10-NN graph, then per sample one Dijkstra pass into a distance
// vector D shared by all passes (D = geodesic distance to the nearest sample so far), then argmax D.  Clouds are
// independent, so here a batch of clouds is sampled at once, one CTA per cloud, with the cloud's state (D, two work
// queues, round stamps: 20 bytes per point) in shared memory up to 11,264 points per cloud and in the caller's workspace
// beyond that.
//
// Same numbers as the reference, not just the same algorithm: all arithmetic is fp64 in the reference's order
// (d2 = (dx*dx + dy*dy) + dz*dz, w = sqrt(d2), no contraction).  A Dijkstra pass is replaced by label-correcting
// relaxation rounds over a work queue — D[v] = min(D[v], D[u] + w(u,v)) with an atomic min on the bit pattern of the
// non-negative double.  Rounded addition is monotone and edge weights are non-negative, so the fixed point these rounds
// reach is exactly the vector Dijkstra returns, bit for bit, whatever the order of the relaxations; the argmax takes the
// lowest index among equal maxima like Eigen's maxCoeff.  What is NOT pinned: which of several exactly equidistant
// neighbours the kd-tree of the reference keeps (here: lowest index).
#include "common.cuh"

namespace {

constexpr int FPS_K = 10;            // neighbours per point, fixed in the reference (sampling.cpp:23)
constexpr int KNN_T = 256;           // queries per CTA of the graph kernel
constexpr int KNN_TILE = 1024;       // candidates staged per round (24 KB)
constexpr int FPS_T = 512;           // threads of a sampling CTA
constexpr int FPS_SMEM_POINTS = 11264;   // 20 B per point -> 220 KB
constexpr unsigned long long INF_BITS = 0x7ff0000000000000ULL;

// ---- 10 nearest OTHER points of every point, fp64, ascending (d2, index) ------------------------------------------
// generateKNN (sampling.cpp:5-17) = nanoflann L2_Simple_Adaptor<double>: d2 accumulated x, y, z from zero.  The point
// itself, which the reference puts in front as a zero-length self loop, is left out: it can never shorten a path.
__global__ void __launch_bounds__(KNN_T) fps_knn_kernel(const double* __restrict__ pos, const int* __restrict__ cloud_ptr, int B, int N,
                                                        int* __restrict__ nbr, double* __restrict__ wgt) {
    __shared__ double tile[KNN_TILE * 3];
    const int i = blockIdx.x * KNN_T + threadIdx.x;
    const int first = blockIdx.x * KNN_T, last = min(N, first + KNN_T) - 1;
    // candidates any query of this CTA can need: from the start of the first query's cloud to the end of the last one's
    const int c_first = cloud_of(cloud_ptr, B, first), c_last = cloud_of(cloud_ptr, B, last);
    const int span_s = cloud_ptr[c_first], span_e = cloud_ptr[c_last + 1];
    int s = 0, e = 0;
    double qx = 0., qy = 0., qz = 0.;
    if (i < N) {
        const int c = cloud_of(cloud_ptr, B, i);
        s = cloud_ptr[c]; e = cloud_ptr[c + 1];
        qx = pos[3 * (size_t)i]; qy = pos[3 * (size_t)i + 1]; qz = pos[3 * (size_t)i + 2];
    }
    double bd[FPS_K];
    int bi[FPS_K];
#pragma unroll
    for (int t = 0; t < FPS_K; ++t) { bd[t] = __longlong_as_double(INF_BITS); bi[t] = -1; }

    for (int base = span_s; base < span_e; base += KNN_TILE) {
        const int n_tile = min(KNN_TILE, span_e - base);
        __syncthreads();
        for (int t = threadIdx.x; t < n_tile * 3; t += KNN_T) tile[t] = pos[3 * (size_t)base + t];
        __syncthreads();
        const int lo = max(base, s) - base, hi = min(base + n_tile, e) - base;
        for (int j = lo; j < hi; ++j) {
            const double dx = __dsub_rn(qx, tile[3 * j]), dy = __dsub_rn(qy, tile[3 * j + 1]), dz = __dsub_rn(qz, tile[3 * j + 2]);
            const double d = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
            // candidates arrive in ascending index, so a strict comparison keeps the lower index among equals
            if (d < bd[FPS_K - 1] && base + j != i) {
                bd[FPS_K - 1] = d; bi[FPS_K - 1] = base + j;
#pragma unroll
                for (int t = FPS_K - 1; t > 0; --t) {
                    if (bd[t] < bd[t - 1]) {
                        const double td = bd[t]; bd[t] = bd[t - 1]; bd[t - 1] = td;
                        const int ti = bi[t]; bi[t] = bi[t - 1]; bi[t - 1] = ti;
                    }
                }
            }
        }
    }
    if (i >= N) return;
#pragma unroll
    for (int t = 0; t < FPS_K; ++t) {
        // clouds of fewer than 11 points: the free slots become self loops of length zero
        const bool have = bi[t] >= 0;
        nbr[(size_t)i * FPS_K + t] = have ? bi[t] - s : i - s;             // index inside the cloud
        wgt[(size_t)i * FPS_K + t] = have ? __dsqrt_rn(bd[t]) : 0.;        // Vector3::norm of the edge (sampling.cpp:71)
    }
}

// ---- the sampler: one CTA per cloud ----------------------------------------------------------------------------------
struct FpsState {
    unsigned long long* D;   // bit patterns of non-negative doubles: unsigned order = numeric order
    int* q[2];
    int* stamp;              // round in which the point was last queued
};

__global__ void __launch_bounds__(FPS_T) fps_sample_kernel(const int* __restrict__ cloud_ptr, const int* __restrict__ nbr,
                                                           const double* __restrict__ wgt, const int* __restrict__ start,
                                                           int n_samples, int* __restrict__ samples, double* __restrict__ dist_or_null,
                                                           int smem_points, unsigned long long* gD, int* gq0, int* gq1, int* gstamp) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int cnt[3];
    __shared__ unsigned long long red_v[FPS_T / 32];
    __shared__ int red_i[FPS_T / 32];
    __shared__ int next_src;

    const int b = blockIdx.x;
    const int s = cloud_ptr[b], P = cloud_ptr[b + 1] - s;
    int* out = samples + (size_t)b * n_samples;
    if (P <= 0) {
        for (int t = threadIdx.x; t < n_samples; t += FPS_T) out[t] = -1;
        return;
    }
    FpsState st;
    if (P <= smem_points) {
        st.D = (unsigned long long*)smem_raw;
        st.q[0] = (int*)(st.D + P);
        st.q[1] = st.q[0] + P;
        st.stamp = st.q[1] + P;
    } else {
        st.D = gD + s; st.q[0] = gq0 + s; st.q[1] = gq1 + s; st.stamp = gstamp + s;
    }
    const int* my_nbr = nbr + (size_t)s * FPS_K;
    const double* my_w = wgt + (size_t)s * FPS_K;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    for (int t = threadIdx.x; t < P; t += FPS_T) { st.D[t] = INF_BITS; st.stamp[t] = -1; }
    int src = min(max(start[b], 0), P - 1);
    int round = 0;                                   // grows over the whole kernel: stamps never repeat
    if (threadIdx.x == 0) out[0] = src;
    __syncthreads();

    for (int smp = 1; smp <= n_samples; ++smp) {
        if (smp == n_samples && dist_or_null == nullptr) break;      // the pass after the last sample only serves `dist`
        // one Dijkstra pass of the reference (sampling.cpp:56-81) from `src`, into the shared D
        if (threadIdx.x == 0) {
            st.D[src] = 0ULL;
            st.q[round & 1][0] = src;
            cnt[round % 3] = 1; cnt[(round + 1) % 3] = 0;
        }
        __syncthreads();
        while (true) {
            const int F = cnt[round % 3];
            if (F == 0) break;
            if (threadIdx.x == 0) cnt[(round + 2) % 3] = 0;
            const int* qc = st.q[round & 1];
            int* qn = st.q[(round + 1) & 1];
            int* cn = &cnt[(round + 1) % 3];
            for (int it = threadIdx.x; it < F * FPS_K; it += FPS_T) {
                const int u = qc[it / FPS_K], j = it % FPS_K;
                const int v = my_nbr[(size_t)u * FPS_K + j];
                const double du = __longlong_as_double(*(volatile unsigned long long*)(st.D + u));
                const unsigned long long nd = (unsigned long long)__double_as_longlong(__dadd_rn(du, my_w[(size_t)u * FPS_K + j]));
                if (nd < *(volatile unsigned long long*)(st.D + v)) {
                    const unsigned long long old = atomicMin(st.D + v, nd);
                    if (nd < old && atomicExch(st.stamp + v, round + 1) != round + 1) {
                        qn[atomicAdd(cn, 1)] = v;
                        // next round starts with v's edges: pull them from L2 while this round finishes
                        asm volatile("prefetch.global.L1 [%0];" :: "l"(my_nbr + (size_t)v * FPS_K));
                        asm volatile("prefetch.global.L1 [%0];" :: "l"(my_w + (size_t)v * FPS_K));
                        asm volatile("prefetch.global.L1 [%0];" :: "l"(my_w + (size_t)v * FPS_K + FPS_K - 1));
                    }
                }
            }
            ++round;
            __syncthreads();
        }
        if (smp == n_samples) break;
        // farthest point: largest D, lowest index among equals (Eigen maxCoeff, sampling.cpp:47)
        unsigned long long bv = 0ULL;
        int bidx = 0x7fffffff;
        for (int t = threadIdx.x; t < P; t += FPS_T) {
            const unsigned long long v = st.D[t];
            if (v > bv || (v == bv && t < bidx)) { bv = v; bidx = t; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long ov = __shfl_xor_sync(FULL, bv, o);
            const int oi = __shfl_xor_sync(FULL, bidx, o);
            if (ov > bv || (ov == bv && oi < bidx)) { bv = ov; bidx = oi; }
        }
        if (lane == 0) { red_v[warp] = bv; red_i[warp] = bidx; }
        __syncthreads();
        if (warp == 0) {
            bv = lane < FPS_T / 32 ? red_v[lane] : 0ULL;
            bidx = lane < FPS_T / 32 ? red_i[lane] : 0x7fffffff;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const unsigned long long ov = __shfl_xor_sync(FULL, bv, o);
                const int oi = __shfl_xor_sync(FULL, bidx, o);
                if (ov > bv || (ov == bv && oi < bidx)) { bv = ov; bidx = oi; }
            }
            if (lane == 0) { next_src = bidx; out[smp] = bidx; }
        }
        __syncthreads();
        src = next_src;
    }
    if (dist_or_null != nullptr)
        for (int t = threadIdx.x; t < P; t += FPS_T) dist_or_null[s + t] = __longlong_as_double((long long)st.D[t]);
}

size_t align256(size_t n) { return (n + 255) / 256 * 256; }

}  // namespace

// nbr + weights, then the out-of-shared-memory state (D, two queues, stamps) that only clouds above 11,264 points use
extern "C" size_t dc_geodesic_fps_ws_bytes(int N) {
    const size_t n = (size_t)(N > 0 ? N : 0);
    return align256(n * FPS_K * 4) + align256(n * FPS_K * 8) + align256(n * 8) + 3 * align256(n * 4);
}

extern "C" int dc_geodesic_fps(const double* pos, const int32_t* cloud_ptr, int B, int N, int max_cloud_points, int n_samples,
                               const int32_t* start, int32_t* samples, double* dist_or_null, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(B >= 1 && N >= 0 && n_samples >= 1, "dc_geodesic_fps: need B >= 1, N >= 0, n_samples >= 1");
    DC_REQUIRE(max_cloud_points >= 0 && max_cloud_points <= N, "dc_geodesic_fps: max_cloud_points must be in [0, N]");
    DC_REQUIRE(ws_bytes >= dc_geodesic_fps_ws_bytes(N), "dc_geodesic_fps: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    char* w = (char*)ws;
    int* nbr = (int*)w; w += align256((size_t)N * FPS_K * 4);
    double* wgt = (double*)w; w += align256((size_t)N * FPS_K * 8);
    unsigned long long* gD = (unsigned long long*)w; w += align256((size_t)N * 8);
    int* gq0 = (int*)w; w += align256((size_t)N * 4);
    int* gq1 = (int*)w; w += align256((size_t)N * 4);
    int* gstamp = (int*)w;
    int launches = 1;
    if (N > 0) {
        fps_knn_kernel<<<dc_ceil_div(N, KNN_T), KNN_T, 0, st>>>(pos, cloud_ptr, B, N, nbr, wgt);
        ++launches;
    }
    // clouds up to `smem_points` keep their state in shared memory; larger ones (only if the caller's bound says they exist) in ws
    const int smem_points = max_cloud_points < FPS_SMEM_POINTS ? max_cloud_points : FPS_SMEM_POINTS;
    const size_t smem = (size_t)smem_points * 20;
    if (smem > 48 * 1024)
        DC_CUDA(cudaFuncSetAttribute(fps_sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FPS_SMEM_POINTS * 20));
    fps_sample_kernel<<<B, FPS_T, smem, st>>>(cloud_ptr, nbr, wgt, start, n_samples, samples, dist_or_null, smem_points, gD, gq0, gq1, gstamp);
    return dc_check_launch_n("dc_geodesic_fps", launches);
}
