// K1 — batched brute-force kNN graph build (replaces torch_cluster.knn_graph as called at
// /root/reference/deltaconv/models/deltanet_base.py:52,63 with loop=True, flow='target_to_source').
//
// One warp per query.  A CTA serves a tile of 32 consecutive queries; the points of the cloud the
// tile lies in are staged in shared memory as SoA (conflict-free 32-wide reads).  Each lane scores
// one candidate per step; the warp keeps its current k best as a sorted list distributed over
// lanes (SLOTS registers per lane, list position p = slot*32 + lane) and inserts the few
// candidates that beat the current k-th best with shuffles.  Ordering key = (d2 bits, index), a
// 64-bit integer, so ties resolve to the lower index and results are bit-reproducible.
#include "common.cuh"
#include "geom_core.cuh"
#include "knn_list.cuh"

namespace {

constexpr int KNN_WARPS = 8;
constexpr int KNN_TILE = 32;        // queries per CTA
constexpr int KNN_CHUNK = 4096;     // points staged per pass: 3 * 4096 * 4 B = 48 KB static smem
// Offer `count` candidates (local indices base..base+count) to the warp's list.
template <int SLOTS, bool STAGED>
__device__ __forceinline__ void scan(u64 (&keys)[SLOTS], u64& tau, int k, int lane, float qx, float qy, float qz,
                                     const float* __restrict__ sx, const float* __restrict__ sy,
                                     const float* __restrict__ sz, const float* __restrict__ gpos,
                                     int base, int count, MergeMin mm) {
    for (int j0 = 0; j0 < count; j0 += 32) {
        int j = j0 + lane;
        u64 key = KEY_INF;
        if (j < count) {
            float px, py, pz;
            if (STAGED) { px = sx[j]; py = sy[j]; pz = sz[j]; }
            else { const float* g = gpos + 3 * (size_t)(base + j); px = __ldg(g); py = __ldg(g + 1); pz = __ldg(g + 2); }
            float d2 = knn_dist2(qx, qy, qz, px, py, pz);
            key = ((u64)__float_as_uint(d2) << 32) | (unsigned)(base + j);
        }
        list_offer<SLOTS>(keys, tau, k, lane, key, mm);
    }
}

template <int SLOTS>
__global__ void __launch_bounds__(KNN_WARPS * 32)
knn_kernel(const float* __restrict__ pos, const int* __restrict__ cloud_ptr, int B, int N, int k,
           int* __restrict__ idx_out, long long* __restrict__ edge_out, float* __restrict__ d2_out, MergeMin mm) {
    __shared__ float sx[KNN_CHUNK], sy[KNN_CHUNK], sz[KNN_CHUNK];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int q_begin = blockIdx.x * KNN_TILE;
    const int q_end = min(N, q_begin + KNN_TILE);
    if (q_begin >= N) return;
    const long long E = (long long)N * k;

    for (int c = cloud_of(cloud_ptr, B, q_begin); c < B; ++c) {
        const int s = __ldg(cloud_ptr + c), e = __ldg(cloud_ptr + c + 1);
        if (s >= q_end) break;
        const int P = e - s;
        if (P <= 0) continue;
        const int qa = max(q_begin, s), qb = min(q_end, e);
        const bool staged = P <= KNN_CHUNK;
        __syncthreads();                     // previous cloud's readers are done
        if (staged) {
            for (int t = threadIdx.x; t < P; t += blockDim.x) {
                const float* g = pos + 3 * (size_t)(s + t);
                sx[t] = __ldg(g); sy[t] = __ldg(g + 1); sz[t] = __ldg(g + 2);
            }
        }
        __syncthreads();
        for (int q = qa + warp; q < qb; q += KNN_WARPS) {
            const float qx = __ldg(pos + 3 * (size_t)q), qy = __ldg(pos + 3 * (size_t)q + 1), qz = __ldg(pos + 3 * (size_t)q + 2);
            u64 keys[SLOTS];
#pragma unroll
            for (int t = 0; t < SLOTS; ++t) keys[t] = KEY_INF;
            u64 tau = KEY_INF;
            if (staged) scan<SLOTS, true>(keys, tau, k, lane, qx, qy, qz, sx, sy, sz, nullptr, 0, P, mm);
            else scan<SLOTS, false>(keys, tau, k, lane, qx, qy, qz, nullptr, nullptr, nullptr, pos + 3 * (size_t)s, 0, P, mm);
#pragma unroll
            for (int t = 0; t < SLOTS; ++t) {
                int p = t * 32 + lane;
                if (p < k) {
                    u64 key = keys[t];
                    // fewer than k points in the cloud: pad with the reference's sentinel N
                    int nb = key == KEY_INF ? N : s + (int)(unsigned)(key & 0xffffffffu);
                    size_t o = (size_t)q * k + p;
                    idx_out[o] = nb;
                    if (edge_out) { edge_out[o] = q; edge_out[E + o] = nb; }
                    if (d2_out) d2_out[o] = key == KEY_INF ? INFINITY : __uint_as_float((unsigned)(key >> 32));
                }
            }
        }
    }
}

__global__ void cloud_ptr_kernel(const long long* __restrict__ batch, int N, int B, int* __restrict__ ptr) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > N) return;
    // batch is sorted: ptr[b] = first i with batch[i] >= b
    long long prev = i == 0 ? -1 : batch[i - 1];
    long long cur = i == N ? B : batch[i];
    if (cur > B) cur = B;
    for (long long b = prev + 1; b <= cur; ++b) ptr[b] = i;
}

__global__ void ell_from_edges_kernel(const long long* __restrict__ col, long long E, int* __restrict__ idx) {
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (t < E) idx[t] = (int)col[t];
}

__global__ void edges_from_ell_kernel(const int* __restrict__ idx, long long E, int k, long long* __restrict__ edge) {
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (t < E) { edge[t] = t / k; edge[E + t] = idx[t]; }
}

}  // namespace

// tuning: survivors per round from which the bitonic merge replaces single insertions (one / two list slots per lane)
static MergeMin g_merge_min = {12, 8};      // scripts/exp_knn_merge.py: two slots 2..33 -> 8 is best (k=40: 7.38 ms; 12: 7.50; 4: 8.51)
MergeMin dc_knn_merge_min() { return g_merge_min; }
extern "C" int dc_set_knn_merge_min(int one_slot, int two_slots) {
    if (one_slot > 0) g_merge_min.one_slot = one_slot;
    if (two_slots > 0) g_merge_min.two_slots = two_slots;
    return DC_OK;
}

extern "C" int dc_cloud_ptr(const int64_t* batch, int N, int B, int32_t* cloud_ptr, void* stream) {
    DC_REQUIRE(N >= 0 && B >= 1, "dc_cloud_ptr: need N >= 0 and B >= 1 (got N=%d B=%d)", N, B);
    cloud_ptr_kernel<<<dc_ceil_div(N + 1, 256), 256, 0, (cudaStream_t)stream>>>((const long long*)batch, N, B, cloud_ptr);
    return dc_check_launch("dc_cloud_ptr");
}

extern "C" int dc_knn(const float* pos, const int32_t* cloud_ptr, int B, int N, int k, int32_t* idx,
                      int64_t* edge_index_or_null, float* d2_or_null, void* stream) {
    DC_REQUIRE(k >= 1 && k <= 128, "dc_knn: k must be in [1,128] (got %d)", k);
    DC_REQUIRE(B >= 1 && N >= 0, "dc_knn: need B >= 1, N >= 0");
    if (N == 0) return DC_OK;
    const int grid = dc_ceil_div(N, KNN_TILE);
    cudaStream_t st = (cudaStream_t)stream;
    long long* eo = (long long*)edge_index_or_null;
    if (k <= 32) knn_kernel<1><<<grid, KNN_WARPS * 32, 0, st>>>(pos, cloud_ptr, B, N, k, idx, eo, d2_or_null, dc_knn_merge_min());
    else if (k <= 64) knn_kernel<2><<<grid, KNN_WARPS * 32, 0, st>>>(pos, cloud_ptr, B, N, k, idx, eo, d2_or_null, dc_knn_merge_min());
    else knn_kernel<4><<<grid, KNN_WARPS * 32, 0, st>>>(pos, cloud_ptr, B, N, k, idx, eo, d2_or_null, dc_knn_merge_min());
    return dc_check_launch("dc_knn");
}

extern "C" int dc_ell_from_edge_index(const int64_t* edge_index, int N, int k, int32_t* idx, void* stream) {
    long long E = (long long)N * k;
    if (E == 0) return DC_OK;
    ell_from_edges_kernel<<<dc_ceil_div(E, 256), 256, 0, (cudaStream_t)stream>>>((const long long*)edge_index + E, E, idx);
    return dc_check_launch("dc_ell_from_edge_index");
}

extern "C" int dc_edge_index_from_ell(const int32_t* idx, int N, int k, int64_t* edge_index, void* stream) {
    long long E = (long long)N * k;
    if (E == 0) return DC_OK;
    edges_from_ell_kernel<<<dc_ceil_div(E, 256), 256, 0, (cudaStream_t)stream>>>(idx, E, k, (long long*)edge_index);
    return dc_check_launch("dc_edge_index_from_ell");
}
