// K1b — kNN graph for LARGE clouds (BASELINE configs[3] up to 1M points per cloud): the brute-force
// kernel of knn.cu is O(P^2) per cloud (1.85 s for one cloud of 2^20 points).  Here every cloud gets a uniform grid
// (counting sort of its points by cell), a warp per query scans the 3x3x3 block of cells around it, then further shells,
// until the k-th best distance is provably smaller than the distance to anything unexamined.  The result is the same
// exact k-nearest list as knn.cu — same d2 arithmetic (knn_dist2), same (d2, index) ordering — so it is bit-identical.
// Everything runs on the device without a host synchronisation.
#include "common.cuh"
#include "geom_core.cuh"
#include "knn_list.cuh"

MergeMin dc_knn_merge_min();      // knn.cu

namespace {

constexpr int GRID_MAX_DIM = 1024;     // cells per axis
constexpr int GRID_WARPS = 8;

struct GridCloud {                     // per cloud, written by grid_params_kernel
    float ox, oy, oz, inv_h, h;
    int gx, gy, gz;
    int cell_base;                     // first global cell id of this cloud
    int s, e;                          // point range
};

// order-preserving float <-> int for atomicMin / atomicMax
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void bbox_init_kernel(int* __restrict__ bbox, int B) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < B * 6) bbox[t] = (t % 6) < 3 ? 0x7fffffff : (int)0x80000000;
}

// A warp whose points all belong to one cloud (the usual case) reduces its 32 points first and sends six atomics instead
// of 192: with hundreds of thousands of points per cloud the atomics on one cloud's six words are the whole cost.
__global__ void bbox_kernel(const float* __restrict__ pos, const int* __restrict__ cloud_ptr, int B, int N, int* __restrict__ bbox) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < N;
    if (!valid) i = N - 1;                                         // keeps the warp whole for the reductions; same values, harmless
    const int b = cloud_of(cloud_ptr, B, i);
    const int x = f2ord(pos[3 * (size_t)i]), y = f2ord(pos[3 * (size_t)i + 1]), z = f2ord(pos[3 * (size_t)i + 2]);
    const int b0 = __shfl_sync(FULL, b, 0);
    if (__all_sync(FULL, b == b0)) {
        const int x0 = __reduce_min_sync(FULL, x), y0 = __reduce_min_sync(FULL, y), z0 = __reduce_min_sync(FULL, z);
        const int x1 = __reduce_max_sync(FULL, x), y1 = __reduce_max_sync(FULL, y), z1 = __reduce_max_sync(FULL, z);
        if ((threadIdx.x & 31) == 0) {
            atomicMin(bbox + b * 6 + 0, x0); atomicMin(bbox + b * 6 + 1, y0); atomicMin(bbox + b * 6 + 2, z0);
            atomicMax(bbox + b * 6 + 3, x1); atomicMax(bbox + b * 6 + 4, y1); atomicMax(bbox + b * 6 + 5, z1);
        }
        return;
    }
    if (!valid) return;
    atomicMin(bbox + b * 6 + 0, x); atomicMin(bbox + b * 6 + 1, y); atomicMin(bbox + b * 6 + 2, z);
    atomicMax(bbox + b * 6 + 3, x); atomicMax(bbox + b * 6 + 4, y); atomicMax(bbox + b * 6 + 5, z);
}

// Lays out all grids: cell size from the bounding volume and the point count, at most 4 P + 64 cells per cloud so that the
// cell arrays have a fixed size.  One CTA: a thread per cloud computes that cloud's grid, then the cell counts are scanned over
// the clouds (first cell id of every cloud).  (The first version was one thread walking all clouds: 21 us for 32 clouds.)
constexpr int GPAR_T = 1024;
__device__ __forceinline__ GridCloud grid_of_cloud(const int* __restrict__ cloud_ptr, const int* __restrict__ bbox, int b, float points_per_cell) {
    GridCloud g;
    g.s = cloud_ptr[b]; g.e = cloud_ptr[b + 1];
    g.cell_base = 0;
    const int P = g.e - g.s;
    if (P <= 0) { g.ox = g.oy = g.oz = 0.f; g.h = 1.f; g.inv_h = 1.f; g.gx = g.gy = g.gz = 1; return g; }
    const float x0 = ord2f(bbox[b * 6]), y0 = ord2f(bbox[b * 6 + 1]), z0 = ord2f(bbox[b * 6 + 2]);
    const float ex = ord2f(bbox[b * 6 + 3]) - x0, ey = ord2f(bbox[b * 6 + 4]) - y0, ez = ord2f(bbox[b * 6 + 5]) - z0;
    const float emax = fmaxf(ex, fmaxf(ey, ez));
    // degenerate extents (planar or collinear clouds) count as one cell layer
    const float floor_ext = fmaxf(emax * 1e-3f, 1e-30f);
    const float vol = fmaxf(ex, floor_ext) * fmaxf(ey, floor_ext) * fmaxf(ez, floor_ext);
    // points per bounding-box cell: given, or (points_per_cell <= 0) chosen from the cloud's size — measured optimum
    // (scripts/tune_knn_small.py, k = 10 ... 40): ~6 for clouds of 1-4 K points (17-28 % faster than 2 at 2048 points),
    // 3 around 16 K, 1.5 at 2^20 (a surface cloud fills only a shell of its box: the occupied cells hold several times this)
    const float ppc = points_per_cell > 0.f ? points_per_cell : (P <= 6144 ? 6.f : (P <= 49152 ? 3.f : 1.5f));
    float h = cbrtf(vol / fmaxf((float)P / ppc, 1.f));
    h = fmaxf(h, emax / (float)(GRID_MAX_DIM - 1));
    if (!(h > 0.f)) h = 1.f;                                    // all points coincide
    const long long cap = 4LL * P + 64;
    for (int it = 0; it < 64; ++it) {
        g.gx = min(GRID_MAX_DIM, (int)(ex / h) + 1); g.gy = min(GRID_MAX_DIM, (int)(ey / h) + 1); g.gz = min(GRID_MAX_DIM, (int)(ez / h) + 1);
        if ((long long)g.gx * g.gy * g.gz <= cap) break;
        h *= 1.26f;
    }
    g.ox = x0; g.oy = y0; g.oz = z0; g.h = h; g.inv_h = 1.f / h;
    return g;
}

__global__ void __launch_bounds__(GPAR_T) grid_params_kernel(const int* __restrict__ cloud_ptr, int B, const int* __restrict__ bbox,
                                                             GridCloud* __restrict__ clouds, int* __restrict__ total_cells, float points_per_cell) {
    __shared__ int warp_tot[32];
    __shared__ int carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < B; base += GPAR_T) {
        const int b = base + threadIdx.x;
        GridCloud g;
        int cells = 0;
        if (b < B) { g = grid_of_cloud(cloud_ptr, bbox, b, points_per_cell); cells = g.gx * g.gy * g.gz; }
        int x = cells;
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
        if (b < B) { g.cell_base = carry + incl - cells; clouds[b] = g; }
        __syncthreads();
        if (threadIdx.x == GPAR_T - 1) carry_s = carry + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_cells = carry_s;
}

__device__ __forceinline__ int cell_coord(float p, float o, float inv_h, int G) {
    const int c = (int)floorf((p - o) * inv_h);
    return min(G - 1, max(0, c));
}

// cell id per point + histogram
__global__ void grid_count_kernel(const float* __restrict__ pos, const int* __restrict__ cloud_ptr, int B, int N,
                                  const GridCloud* __restrict__ clouds, int* __restrict__ cell_of, int* __restrict__ cnt) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const GridCloud g = clouds[cloud_of(cloud_ptr, B, i)];
    const int cx = cell_coord(pos[3 * (size_t)i], g.ox, g.inv_h, g.gx);
    const int cy = cell_coord(pos[3 * (size_t)i + 1], g.oy, g.inv_h, g.gy);
    const int cz = cell_coord(pos[3 * (size_t)i + 2], g.oz, g.inv_h, g.gz);
    const int cell = g.cell_base + (cz * g.gy + cy) * g.gx + cx;
    cell_of[i] = cell;
    atomicAdd(cnt + cell, 1);
}

constexpr int GSCAN_T = 1024, GSCAN_V = 4;
// exclusive scan of the cell counts, one CTA per cloud (4096 cells per round): the cells of cloud b are consecutive and its
// points occupy sorted[s_b, e_b), so every cloud's scan starts at s_b and the clouds are independent.  out[total] = N.
__global__ void __launch_bounds__(GSCAN_T) grid_scan_kernel(const int* __restrict__ cnt, const GridCloud* __restrict__ clouds, int B, int N,
                                                            int* __restrict__ out) {
    const GridCloud g = clouds[blockIdx.x];
    const int first = g.cell_base, n = g.gx * g.gy * g.gz;
    __shared__ int warp_tot[32];
    __shared__ int carry_s;
    if (threadIdx.x == 0) carry_s = g.s;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < n; base += GSCAN_T * GSCAN_V) {
        const int i0 = base + threadIdx.x * GSCAN_V;
        int v[GSCAN_V], sum = 0;
#pragma unroll
        for (int u = 0; u < GSCAN_V; ++u) { v[u] = (i0 + u < n) ? cnt[first + i0 + u] : 0; sum += v[u]; }
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
        for (int u = 0; u < GSCAN_V; ++u) { if (i0 + u < n) out[first + i0 + u] = run; run += v[u]; }
        __syncthreads();
        if (threadIdx.x == GSCAN_T - 1) carry_s = carry + incl;
        __syncthreads();
    }
    if (blockIdx.x == B - 1 && threadIdx.x == 0) out[first + n] = N;
}

// points sorted by cell: (x, y, z, local index bits).  The order inside a cell is arbitrary (atomic cursor); the selection
// below does not depend on it because keys are unique.
__global__ void grid_fill_kernel(const float* __restrict__ pos, const int* __restrict__ cloud_ptr, int B, int N,
                                 const int* __restrict__ cell_of, const int* __restrict__ cell_start, int* __restrict__ cursor,
                                 float4* __restrict__ sorted) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const int cell = cell_of[i];
    const int slot = atomicSub(cursor + cell, 1) - 1;           // `cursor` is the cell's count, counted down: no second zeroed array
    const int s = cloud_ptr[cloud_of(cloud_ptr, B, i)];
    sorted[cell_start[cell] + slot] = make_float4(pos[3 * (size_t)i], pos[3 * (size_t)i + 1], pos[3 * (size_t)i + 2], __int_as_float(i - s));
}

// Offer the sorted points [begin, end) to the list (all lanes take part in every round).
template <int SLOTS>
__device__ __forceinline__ void grid_scan_range(u64 (&keys)[SLOTS], u64& tau, int k, int lane, float qx, float qy, float qz,
                                                const float4* __restrict__ sorted, int begin, int end, MergeMin mm) {
    for (int j0 = begin; j0 < end; j0 += 32) {
        const int j = j0 + lane;
        u64 key = KEY_INF;
        if (j < end) {
            const float4 p = __ldg(sorted + j);
            const float d2 = knn_dist2(qx, qy, qz, p.x, p.y, p.z);
            key = ((u64)__float_as_uint(d2) << 32) | (unsigned)__float_as_int(p.w);
        }
        list_offer<SLOTS>(keys, tau, k, lane, key, mm);
    }
}

// One warp per query, queries taken in cell order (neighbouring warps read the same cells).
template <int SLOTS>
__global__ void __launch_bounds__(GRID_WARPS * 32)
grid_query_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start, const int* __restrict__ cloud_ptr, int B, int N, int k,
                  const GridCloud* __restrict__ clouds, int* __restrict__ idx_out, long long* __restrict__ edge_out, float* __restrict__ d2_out, MergeMin mm) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = blockIdx.x * GRID_WARPS + warp;                 // position of the query in the sorted array
    if (j >= N) return;
    const GridCloud g = clouds[cloud_of(cloud_ptr, B, j)];       // cloud b owns sorted[s_b, e_b): its cells are consecutive
    const float4 q4 = __ldg(sorted + j);
    const int q = g.s + __float_as_int(q4.w);
    const float qx = q4.x, qy = q4.y, qz = q4.z;
    const float tx = (qx - g.ox) * g.inv_h, ty = (qy - g.oy) * g.inv_h, tz = (qz - g.oz) * g.inv_h;   // same expressions as cell_coord
    const int cx = min(g.gx - 1, max(0, (int)floorf(tx))), cy = min(g.gy - 1, max(0, (int)floorf(ty))), cz = min(g.gz - 1, max(0, (int)floorf(tz)));
    u64 keys[SLOTS];
#pragma unroll
    for (int t = 0; t < SLOTS; ++t) keys[t] = KEY_INF;
    u64 tau = KEY_INF;
    for (int R = 1;; ++R) {
        for (int dz = -R; dz <= R; ++dz) {
            const int z = cz + dz;
            if (z < 0 || z >= g.gz) continue;
            for (int dy = -R; dy <= R; ++dy) {
                const int y = cy + dy;
                if (y < 0 || y >= g.gy) continue;
                const int row = g.cell_base + (z * g.gy + y) * g.gx;
                if (R == 1 || abs(dz) == R || abs(dy) == R) {     // a whole row of the block: its cells are contiguous in memory
                    const int x0 = max(cx - R, 0), x1 = min(cx + R, g.gx - 1);
                    grid_scan_range<SLOTS>(keys, tau, k, lane, qx, qy, qz, sorted, __ldg(cell_start + row + x0), __ldg(cell_start + row + x1 + 1), mm);
                } else {                                          // interior row of a later shell: only its two end cells are new
                    if (cx - R >= 0)
                        grid_scan_range<SLOTS>(keys, tau, k, lane, qx, qy, qz, sorted, __ldg(cell_start + row + cx - R), __ldg(cell_start + row + cx - R + 1), mm);
                    if (cx + R <= g.gx - 1)
                        grid_scan_range<SLOTS>(keys, tau, k, lane, qx, qy, qz, sorted, __ldg(cell_start + row + cx + R), __ldg(cell_start + row + cx + R + 1), mm);
                }
            }
        }
        // Everything unexamined lies in a cell at index distance > R along some axis.  A point's cell is floor(t) with the same
        // float expression t as above (monotone in the coordinate), so along that axis |t_p - t_q| >= gap; t is at most 1024,
        // whose ulp is 1.2e-4: 2e-3 cells of slack cover the rounding of both t's, and one part in 1e6 the product with h.
        float gap = INFINITY;
        if (cx - R > 0) gap = fminf(gap, tx - (float)(cx - R));
        if (cx + R < g.gx - 1) gap = fminf(gap, (float)(cx + R + 1) - tx);
        if (cy - R > 0) gap = fminf(gap, ty - (float)(cy - R));
        if (cy + R < g.gy - 1) gap = fminf(gap, (float)(cy + R + 1) - ty);
        if (cz - R > 0) gap = fminf(gap, tz - (float)(cz - R));
        if (cz + R < g.gz - 1) gap = fminf(gap, (float)(cz + R + 1) - tz);
        if (gap == INFINITY) break;                               // the block covers the whole grid
        if (tau != KEY_INF) {
            const float dmin = (gap - 2e-3f) * g.h * (1.f - 1e-6f);
            if (dmin > 0.f && __uint_as_float((unsigned)(tau >> 32)) < dmin * dmin) break;
        }
    }
    const long long E = (long long)N * k;
#pragma unroll
    for (int t = 0; t < SLOTS; ++t) {
        const int p = t * 32 + lane;
        if (p < k) {
            const u64 key = keys[t];
            const int nb = key == KEY_INF ? N : g.s + (int)(unsigned)(key & 0xffffffffu);     // fewer than k points: the reference's sentinel N
            const size_t o = (size_t)q * k + p;
            idx_out[o] = nb;
            if (edge_out) { edge_out[o] = q; edge_out[E + o] = nb; }
            if (d2_out) d2_out[o] = key == KEY_INF ? INFINITY : __uint_as_float((unsigned)(key >> 32));
        }
    }
}

float g_points_per_cell = 0.f;       // target occupancy of a bounding-box cell (dc_set_knn_grid_density); 0 = by cloud size
// 30-bit Morton code of a point inside its cloud's bounding box (10 bits per axis), cloud index above it: sorting by this key
// orders every cloud along a space-filling curve without mixing clouds.
__device__ __forceinline__ unsigned spread10(unsigned v) {
    v = (v | (v << 16)) & 0x030000FFu; v = (v | (v << 8)) & 0x0300F00Fu; v = (v | (v << 4)) & 0x030C30C3u; v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__global__ void morton_key_kernel(const float* __restrict__ pos, const int* __restrict__ cloud_ptr, int B, int N, const int* __restrict__ bbox,
                                  long long* __restrict__ keys) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const int b = cloud_of(cloud_ptr, B, i);
    const float x0 = ord2f(bbox[b * 6]), y0 = ord2f(bbox[b * 6 + 1]), z0 = ord2f(bbox[b * 6 + 2]);
    const float ext = fmaxf(fmaxf(ord2f(bbox[b * 6 + 3]) - x0, ord2f(bbox[b * 6 + 4]) - y0), fmaxf(ord2f(bbox[b * 6 + 5]) - z0, 1e-30f));
    const float sc = 1023.999f / ext;
    const unsigned cx = min(1023u, (unsigned)fmaxf((pos[3 * (size_t)i] - x0) * sc, 0.f));
    const unsigned cy = min(1023u, (unsigned)fmaxf((pos[3 * (size_t)i + 1] - y0) * sc, 0.f));
    const unsigned cz = min(1023u, (unsigned)fmaxf((pos[3 * (size_t)i + 2] - z0) * sc, 0.f));
    keys[i] = ((long long)b << 30) | (long long)(spread10(cx) | (spread10(cy) << 1) | (spread10(cz) << 2));
}

size_t align256(size_t n) { return (n + 255) / 256 * 256; }
long long grid_cell_cap(int N, int B) { return 4LL * N + 65LL * B; }

}  // namespace

// keys[i] = (cloud of i) << 30 | Morton code of point i in its cloud's bounding box; ws: B * 24 bytes.  Sorting the points of a
// large cloud by this key (stable) keeps the kNN neighbourhoods of consecutive points in the same cache lines: measured 2-2.3x on
// the SpMM gathers of a 2^20-point cloud (scripts/exp_point_order.py).
extern "C" int dc_morton_keys(const float* pos, const int32_t* cloud_ptr, int B, int N, int64_t* keys, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(B >= 1 && N >= 0, "dc_morton_keys: need B >= 1, N >= 0");
    DC_REQUIRE(ws_bytes >= (size_t)B * 24, "dc_morton_keys: workspace too small");
    if (N == 0) return DC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int* bbox = (int*)ws;
    bbox_init_kernel<<<dc_ceil_div(B * 6, 256), 256, 0, st>>>(bbox, B);
    bbox_kernel<<<dc_ceil_div(N, 256), 256, 0, st>>>(pos, cloud_ptr, B, N, bbox);
    morton_key_kernel<<<dc_ceil_div(N, 256), 256, 0, st>>>(pos, cloud_ptr, B, N, bbox, (long long*)keys);
    return dc_check_launch_n("dc_morton_keys", 3);
}

// tuning: points per bounding-box cell the grid aims at (0 = default: chosen per cloud from its size, grid_of_cloud)
extern "C" int dc_set_knn_grid_density(float points_per_cell) {
    DC_REQUIRE(points_per_cell == 0.f || (points_per_cell >= 0.25f && points_per_cell <= 64.f), "dc_set_knn_grid_density: out of range (0 = automatic)");
    g_points_per_cell = points_per_cell;
    return DC_OK;
}

extern "C" size_t dc_knn_grid_ws_bytes(int N, int B) {
    const size_t cap = (size_t)grid_cell_cap(N, B);
    return align256((size_t)B * 6 * 4) + align256((size_t)B * sizeof(GridCloud)) + 256 + align256((size_t)N * 4) +
           3 * align256((cap + 2) * 4) + align256((size_t)N * sizeof(float4));
}

extern "C" int dc_knn_grid(const float* pos, const int32_t* cloud_ptr, int B, int N, int k, int32_t* idx,
                           int64_t* edge_index_or_null, float* d2_or_null, void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(k >= 1 && k <= 128, "dc_knn_grid: k must be in [1,128] (got %d)", k);
    DC_REQUIRE(B >= 1 && N >= 0, "dc_knn_grid: need B >= 1, N >= 0");
    DC_REQUIRE(ws_bytes >= dc_knn_grid_ws_bytes(N, B), "dc_knn_grid: workspace too small");
    DC_REQUIRE(grid_cell_cap(N, B) < 2147483000LL, "dc_knn_grid: too many points");
    if (N == 0) return DC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int cap = (int)grid_cell_cap(N, B);
    char* w = (char*)ws;
    int* bbox = (int*)w; w += align256((size_t)B * 6 * 4);
    GridCloud* clouds = (GridCloud*)w; w += align256((size_t)B * sizeof(GridCloud));
    int* total = (int*)w; w += 256;
    int* cell_of = (int*)w; w += align256((size_t)N * 4);
    int* cnt = (int*)w; w += align256(((size_t)cap + 2) * 4);
    int* cell_start = (int*)w; w += align256(((size_t)cap + 2) * 4);
    int* cursor = (int*)w; w += align256(((size_t)cap + 2) * 4);
    float4* sorted = (float4*)w;
    DC_CUDA(cudaMemsetAsync(cnt, 0, ((size_t)cap + 2) * 4, st));
    (void)cursor;
    bbox_init_kernel<<<dc_ceil_div(B * 6, 256), 256, 0, st>>>(bbox, B);
    bbox_kernel<<<dc_ceil_div(N, 256), 256, 0, st>>>(pos, cloud_ptr, B, N, bbox);
    grid_params_kernel<<<1, GPAR_T, 0, st>>>(cloud_ptr, B, bbox, clouds, total, g_points_per_cell);
    grid_count_kernel<<<dc_ceil_div(N, 256), 256, 0, st>>>(pos, cloud_ptr, B, N, clouds, cell_of, cnt);
    grid_scan_kernel<<<B, GSCAN_T, 0, st>>>(cnt, clouds, B, N, cell_start);
    grid_fill_kernel<<<dc_ceil_div(N, 256), 256, 0, st>>>(pos, cloud_ptr, B, N, cell_of, cell_start, cnt, sorted);
    const int grid = dc_ceil_div(N, GRID_WARPS);
    long long* eo = (long long*)edge_index_or_null;
    if (k <= 32) grid_query_kernel<1><<<grid, GRID_WARPS * 32, 0, st>>>(sorted, cell_start, cloud_ptr, B, N, k, clouds, idx, eo, d2_or_null, dc_knn_merge_min());
    else if (k <= 64) grid_query_kernel<2><<<grid, GRID_WARPS * 32, 0, st>>>(sorted, cell_start, cloud_ptr, B, N, k, clouds, idx, eo, d2_or_null, dc_knn_merge_min());
    else grid_query_kernel<4><<<grid, GRID_WARPS * 32, 0, st>>>(sorted, cell_start, cloud_ptr, B, N, k, clouds, idx, eo, d2_or_null, dc_knn_merge_min());
    return dc_check_launch_n("dc_knn_grid", 7);
}
