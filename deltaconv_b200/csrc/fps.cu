// K12 — geodesic farthest-point sampling of a batch of clouds (reference: deltaconv/cpp/sampling.cpp:21-81,
// the pre-transform transforms/geodesic_fps.py:14-43 runs on every shape of a dataset).
//
// The reference handles one cloud on one CPU thread: kd-tree 10-NN graph, then per sample one Dijkstra pass into a distance
// vector D shared by all passes (D = geodesic distance to the nearest sample so far), then argmax D.  Clouds are
// independent, so here a batch of clouds is sampled at once:
//   1. fps_sort_box_kernel   points gathered into the caller's visiting order (a space-filling-curve order makes step 2 fast;
//                            any order is correct) + one bounding box per 64 consecutive points, one per 8 boxes;
//   2. fps_knn_kernel        the 10 nearest other points of every point, a thread per query: boxes that lie farther than the
//                            current 10th distance are skipped, the others are screened in fp32 with a proven error margin
//                            and only the few survivors are evaluated in fp64;
//   3. fps_sample_kernel     one CTA per cloud, the cloud's state (D, two work queues, two bit masks: 12.25 bytes per point)
//                            in shared memory up to 18,432 points per cloud — two clouds of 8,192 per SM — and in the
//                            caller's workspace beyond that.
//
// Same numbers as the reference, not just the same algorithm: all arithmetic is fp64 in the reference's order
// (d2 = (dx*dx + dy*dy) + dz*dz, w = sqrt(d2), no contraction).  A Dijkstra pass is replaced by label-correcting
// relaxation rounds over a work queue — D[v] = min(D[v], D[u] + w(u,v)) with an atomic min on the bit pattern of the
// non-negative double.  Rounded addition is monotone and edge weights are non-negative, so the fixed point these rounds
// reach is exactly the vector Dijkstra returns, bit for bit, whatever the order of the relaxations; the argmax takes the
// lowest ORIGINAL index among equal maxima like Eigen's maxCoeff.  The pruning of the neighbour search can not change the
// result either: boxes are stored rounded outwards and their distance is computed with round-down fp32 operations (a lower
// bound of the fp64 distance to every point inside), the fp32 screening of single points carries a proven margin
// (see fps_knn_kernel), and whatever survives is evaluated in fp64.  What is NOT pinned: which of several exactly
// equidistant neighbours the kd-tree of the reference keeps (here: lowest original index).
#include "common.cuh"

namespace {

constexpr int FPS_K = 10;            // neighbours per point, fixed in the reference (sampling.cpp:23)
constexpr int BOX_TILE = 64;         // consecutive points (in visiting order) per bounding box
constexpr int SUPER = 8;              // boxes per super box
constexpr int KNN_WARPS = 8;         // warps of 32 queries per CTA of the graph kernel
constexpr int FPS_T = 512;           // threads of a sampling CTA
constexpr int FPS_SMEM_POINTS = 18432;  // < 65,536: queue entries of shared-memory clouds are 16-bit
constexpr unsigned long long INF_BITS = 0x7ff0000000000000ULL;

// the graph: one 128-byte record (one cache line) per point, so that a relaxation round fetches one line per frontier point
struct __align__(128) FpsEdges {
    double len[FPS_K];       // edge lengths
    int to[FPS_K];           // neighbours, as positions inside the cloud (visiting order)
    int pad[2];
};
static_assert(sizeof(FpsEdges) == 128, "one line per point");

__host__ __device__ constexpr size_t fps_smem_bytes(int points) { return (size_t)points * 12 + 8 * (size_t)((points + 31) / 32); }

__device__ __forceinline__ int orig_of(const int* __restrict__ order, int j) { return order != nullptr ? __ldg(order + j) : j; }

// ---- 1. gather into visiting order + boxes ----------------------------------------------------------------------------
__global__ void __launch_bounds__(128) fps_sort_box_kernel(const double* __restrict__ pos, const int* __restrict__ order, int N,
                                                           double* __restrict__ spos, float4* __restrict__ spos32,
                                                           float4* __restrict__ boxes) {
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * 4 + (threadIdx.x >> 5);
    if ((long long)g * BOX_TILE >= N) return;
    const double inf = __longlong_as_double(INF_BITS);
    double lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
#pragma unroll
    for (int h = 0; h < BOX_TILE / 32; ++h) {
        const int i = g * BOX_TILE + h * 32 + lane;
        float4 f = make_float4(0.f, 0.f, 0.f, 0.f);               // the fp32 copy is padded to whole tiles
        if (i < N) {
            const int o = orig_of(order, i);
            double p[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                p[a] = pos[3 * (size_t)o + a];
                spos[3 * (size_t)i + a] = p[a];
                lo[a] = fmin(lo[a], p[a]); hi[a] = fmax(hi[a], p[a]);
            }
            f = make_float4(__double2float_rn(p[0]), __double2float_rn(p[1]), __double2float_rn(p[2]), 0.f);
        }
        spos32[i] = f;
    }
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fmin(lo[a], __shfl_xor_sync(FULL, lo[a], o));
            hi[a] = fmax(hi[a], __shfl_xor_sync(FULL, hi[a], o));
        }
    // stored in fp32, rounded outwards: the box can only grow, so a distance to it stays a lower bound
    if (lane == 0) {
        boxes[2 * (size_t)g] = make_float4(__double2float_rd(lo[0]), __double2float_rd(lo[1]), __double2float_rd(lo[2]), 0.f);
        boxes[2 * (size_t)g + 1] = make_float4(__double2float_ru(hi[0]), __double2float_ru(hi[1]), __double2float_ru(hi[2]), 0.f);
    }
}

// one box around every SUPER consecutive boxes: the search tests these first
__global__ void fps_super_box_kernel(const float4* __restrict__ boxes, int tiles, float4* __restrict__ super_boxes) {
    const int G = blockIdx.x * blockDim.x + threadIdx.x;
    if ((long long)G * SUPER >= tiles) return;
    const float inf = __int_as_float(0x7f800000);
    float4 lo = make_float4(inf, inf, inf, 0.f), hi = make_float4(-inf, -inf, -inf, 0.f);
    for (int g = G * SUPER; g < min(tiles, G * SUPER + SUPER); ++g) {
        const float4 a = boxes[2 * (size_t)g], b = boxes[2 * (size_t)g + 1];
        lo.x = fminf(lo.x, a.x); lo.y = fminf(lo.y, a.y); lo.z = fminf(lo.z, a.z);
        hi.x = fmaxf(hi.x, b.x); hi.y = fmaxf(hi.y, b.y); hi.z = fmaxf(hi.z, b.z);
    }
    super_boxes[2 * (size_t)G] = lo; super_boxes[2 * (size_t)G + 1] = hi;
}

// ---- 2. 10 nearest OTHER points of every point, fp64, ascending (d2, original index) ---------------------------------
// generateKNN (sampling.cpp:5-17) = nanoflann L2_Simple_Adaptor<double>: d2 accumulated x, y, z from zero.  The point
// itself, which the reference puts in front as a zero-length self loop, is left out: it can never shorten a path.
// Indices here are positions in visiting order; `order` maps them back where ties have to be broken.
__device__ __forceinline__ double dist2_rn(double dx, double dy, double dz) {
    return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// Lower bound, in fp32 with every operation rounded down, of the squared distance from a query (given as the fp32 values just
// above and just below its coordinates) to an outward-rounded box.  It never exceeds the exact squared distance to any
// point inside; the fp64-computed d2 of such a point is below its exact value by at most a few 2^-53, which the caller's
// factor (1 - 1e-15) absorbs.  Rounding down also turns fp32 overflow into FLT_MAX, still a lower bound.
__device__ __forceinline__ float box_lower_bound(const float4 lo, const float4 hi, float xu, float xd, float yu, float yd, float zu, float zd) {
    const float ax = fmaxf(fmaxf(__fsub_rd(lo.x, xu), __fsub_rd(xd, hi.x)), 0.f);
    const float ay = fmaxf(fmaxf(__fsub_rd(lo.y, yu), __fsub_rd(yd, hi.y)), 0.f);
    const float az = fmaxf(fmaxf(__fsub_rd(lo.z, zu), __fsub_rd(zd, hi.z)), 0.f);
    return __fadd_rd(__fadd_rd(__fmul_rd(ax, ax), __fmul_rd(ay, ay)), __fmul_rd(az, az));
}

// A box that survives the test is scanned in two phases.  Phase A runs over its 64 points in fp32 and only marks the ones
// that may be at least as near as the current 10th neighbour; phase B evaluates the marked ones (a handful) in fp64 and
// inserts them.  Phase A can not lose a neighbour: with u = 2^-24, S = largest |coordinate| of the cloud, T = current 10th d2,
//   fp32 inputs differ from the fp64 ones by at most u S, each fp32 difference by another u |a|, the fp32 sum of squares
//   d32 <= |a|^2 (1+u)^3, hence the true distance d >= sqrt(d32) (1 - 3u) - 2 sqrt(3) u S;
//   a point is dropped only if d32 > R^2, R = (sqrt(T) (1 + 1e-15) + 2 sqrt(3) u S) / (1 - 3u), which implies that its
//   fp64-computed d2 exceeds T.  (Threshold rounded up, 1e-12 of slack for the fp64 roundings in R and fp32 underflow.)
// Outside 1e-9 <= S <= 1e18 (fp32 squares would underflow or overflow) phase A marks everything.
__global__ void __launch_bounds__(KNN_WARPS * 32, 3) fps_knn_kernel(const double* __restrict__ spos, const float4* __restrict__ spos32,
                                                                 const float4* __restrict__ boxes, const float4* __restrict__ super_boxes,
                                                                 const int* __restrict__ order, const int* __restrict__ cloud_ptr,
                                                                 int B, int N, FpsEdges* __restrict__ edges) {
    const int lane = threadIdx.x & 31;
    const int qbase = (blockIdx.x * KNN_WARPS + (threadIdx.x >> 5)) * 32;
    if (qbase >= N) return;                                    // whole warp
    const int i = qbase + lane;
    const bool valid = i < N;
    const int iq = valid ? i : N - 1;
    const int c = cloud_of(cloud_ptr, B, iq);
    const int s = cloud_ptr[c], e = cloud_ptr[c + 1];
    const double qx = spos[3 * (size_t)iq], qy = spos[3 * (size_t)iq + 1], qz = spos[3 * (size_t)iq + 2];
    const float fx = __double2float_rn(qx), fy = __double2float_rn(qy), fz = __double2float_rn(qz);
    const float xu = __double2float_ru(qx), xd = __double2float_rd(qx), yu = __double2float_ru(qy), yd = __double2float_rd(qy),
                zu = __double2float_ru(qz), zd = __double2float_rd(qz);
    const int g_lo = __reduce_min_sync(FULL, s) / BOX_TILE, g_hi = (__reduce_max_sync(FULL, e) - 1) / BOX_TILE;
    const int g0 = qbase / BOX_TILE;
    const int G_lo = g_lo / SUPER, G_hi = g_hi / SUPER, G0 = g0 / SUPER;

    // S over every box this warp can visit (boxes shared with a neighbouring cloud, and the outward rounding, only make it
    // larger: still a bound)
    float S32 = 0.f;
    for (int G = G_lo + lane; G <= G_hi; G += 32) {
        const float4 a = __ldg(super_boxes + 2 * (size_t)G), b = __ldg(super_boxes + 2 * (size_t)G + 1);
        S32 = fmaxf(S32, fmaxf(fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fmaxf(fabsf(a.z), fabsf(b.x))), fmaxf(fabsf(b.y), fabsf(b.z))));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) S32 = fmaxf(S32, __shfl_xor_sync(FULL, S32, o));
    const double S = (double)S32;
    const bool use32 = S >= 1e-9 && S <= 1e18;
    const double u = 5.9604644775390625e-08;                   // 2^-24
    const double slack = 2. * 1.7320508075688774 * u * S * (1. + 1e-6);

    double bd[FPS_K];
    int bi[FPS_K];
#pragma unroll
    for (int t = 0; t < FPS_K; ++t) { bd[t] = __longlong_as_double(INF_BITS); bi[t] = -1; }

    auto scan_box = [&](int g) {
        const int j_first = g * BOX_TILE, count = min(BOX_TILE, N - j_first);
        const float lb = box_lower_bound(__ldg(boxes + 2 * (size_t)g), __ldg(boxes + 2 * (size_t)g + 1), xu, xd, yu, yd, zu, zd);
        const bool near = valid && max(j_first, s) < min(j_first + count, e) && (double)lb * (1. - 1e-15) <= bd[FPS_K - 1];
        if (!__any_sync(FULL, near)) return;

        // 32 points at a time: phase A with the threshold of the moment, then phase B, so that the second half of a box is
        // already screened against what the first half found (in the own box nothing is known before)
#pragma unroll
        for (int h = 0; h < BOX_TILE / 32; ++h) {
            float thr = __int_as_float(0x7f800000);
            if (use32) {
                const double R = (sqrt(bd[FPS_K - 1]) * (1. + 1e-15) + slack) / (1. - 3. * u);
                thr = __double2float_ru(R * R * (1. + 1e-12));
            }
            if (!near) thr = -1.f;
            unsigned m = 0u;
#pragma unroll 8
            for (int jj = 0; jj < 32; ++jj) {
                const float4 p = __ldg(spos32 + j_first + h * 32 + jj);
                const float dx = fx - p.x, dy = fy - p.y, dz = fz - p.z;
                const float d32 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                m |= (d32 > thr ? 0u : 1u) << jj;
            }
            const int left = count - h * 32;
            m = left >= 32 ? m : (left <= 0 ? 0u : m & ((1u << left) - 1u));
            if (!near) m = 0u;

            while (m != 0u) {
                const int j = j_first + h * 32 + __ffs(m) - 1;
                m &= m - 1u;
                if ((unsigned)(j - s) >= (unsigned)(e - s) || j == i) continue;
                const double d = dist2_rn(__dsub_rn(qx, __ldg(spos + 3 * (size_t)j)), __dsub_rn(qy, __ldg(spos + 3 * (size_t)j + 1)),
                                          __dsub_rn(qz, __ldg(spos + 3 * (size_t)j + 2)));
                if (!(d <= bd[FPS_K - 1])) continue;
                bool below[FPS_K], tie = false;
#pragma unroll
                for (int t = 0; t < FPS_K; ++t) { below[t] = d < bd[t]; tie |= d == bd[t]; }
                if (!tie) {
                    // d differs from every entry (so d < bd[9]): shift the tail and drop it in, without branches
#pragma unroll
                    for (int t = FPS_K - 1; t > 0; --t) {
                        bd[t] = below[t - 1] ? bd[t - 1] : (below[t] ? d : bd[t]);
                        bi[t] = below[t - 1] ? bi[t - 1] : (below[t] ? j : bi[t]);
                    }
                    bd[0] = below[0] ? d : bd[0];
                    bi[0] = below[0] ? j : bi[0];
                    continue;
                }
                // equal distances: order by original index
                if (d == bd[FPS_K - 1] && (bi[FPS_K - 1] < 0 || orig_of(order, j) > orig_of(order, bi[FPS_K - 1]))) continue;
                bd[FPS_K - 1] = d; bi[FPS_K - 1] = j;
#pragma unroll
                for (int t = FPS_K - 1; t > 0; --t) {
                    const bool up = bd[t] < bd[t - 1] ||
                                    (bd[t] == bd[t - 1] && bi[t - 1] >= 0 && orig_of(order, bi[t]) < orig_of(order, bi[t - 1]));
                    if (up) {
                        const double td = bd[t]; bd[t] = bd[t - 1]; bd[t - 1] = td;
                        const int ti = bi[t]; bi[t] = bi[t - 1]; bi[t - 1] = ti;
                    }
                }
            }
        }
    };

    // own box first: neighbours in a space-filling order are mostly neighbours in space, so the 10th distance is tight
    // right away; then the super boxes from the own one outwards, and inside a near one its boxes
    scan_box(g0);
    const int reach = max(G0 - G_lo, G_hi - G0);
    for (int step = 0; step <= reach; ++step) {
#pragma unroll 1
        for (int side = 0; side < 2; ++side) {
            const int G = side == 0 ? G0 + step : G0 - step;
            if (G > G_hi || G < G_lo || (side == 1 && step == 0)) continue;
            const float lb = box_lower_bound(__ldg(super_boxes + 2 * (size_t)G), __ldg(super_boxes + 2 * (size_t)G + 1), xu, xd, yu, yd, zu, zd);
            if (!__any_sync(FULL, valid && (double)lb * (1. - 1e-15) <= bd[FPS_K - 1])) continue;
#pragma unroll 1
            for (int g = max(G * SUPER, g_lo); g <= min(G * SUPER + SUPER - 1, g_hi); ++g)
                if (g != g0) scan_box(g);
        }
    }
    if (!valid) return;
#pragma unroll
    for (int t = 0; t < FPS_K; ++t) {
        // clouds of fewer than 11 points: the free slots become self loops of length zero
        const bool have = bi[t] >= 0;
        edges[i].to[t] = have ? bi[t] - s : i - s;                          // position inside the cloud, visiting order
        edges[i].len[t] = have ? __dsqrt_rn(bd[t]) : 0.;                    // Vector3::norm of the edge (sampling.cpp:71)
    }
}

// ---- 3. the sampler: one CTA per cloud ---------------------------------------------------------------------------------
struct FpsShared {
    int cnt[3];
    unsigned long long red_v[FPS_T / 32];
    int red_i[FPS_T / 32], red_o[FPS_T / 32];
    int start_pos;
};

// (value, original index) order of the argmax: larger value first, then the lower original index (Eigen maxCoeff keeps the first)
__device__ __forceinline__ bool better(unsigned long long v, int o, unsigned long long bv, int bo) { return v > bv || (v == bv && o < bo); }

// Arg-max over the lanes of a warp with three warp-wide integer reductions instead of a shuffle tree: the high words, the
// low words among the lanes that hold the largest high word, then the lowest original index among the lanes that hold the
// largest value (`o` < 0 = not looked up yet: only those lanes fetch it).  Every lane returns the winner; no lane taking
// part gives (0, 0, INT_MAX).
__device__ __forceinline__ void warp_argmax(bool takes_part, unsigned long long v, int idx, int& o, const int* __restrict__ order, int s,
                                            unsigned long long& best_v, int& best_idx, int& best_o) {
    const unsigned hi = takes_part ? (unsigned)(v >> 32) : 0u, lo = takes_part ? (unsigned)v : 0u;
    const unsigned mh = __reduce_max_sync(FULL, hi);
    const unsigned ml = __reduce_max_sync(FULL, hi == mh ? lo : 0u);
    const bool top = takes_part && hi == mh && lo == ml;
    if (top && o < 0) o = orig_of(order, s + idx);
    const int mo = __reduce_min_sync(FULL, top ? o : 0x7fffffff);
    const unsigned who = __ballot_sync(FULL, top && o == mo);
    const int winner = who != 0u ? __ffs(who) - 1 : 0;
    best_idx = who != 0u ? __shfl_sync(FULL, idx, winner) : 0;
    best_v = who != 0u ? (((unsigned long long)mh << 32) | ml) : 0ULL;
    best_o = who != 0u ? mo : 0x7fffffff;
}

// D: bit patterns of non-negative doubles (unsigned order = numeric order).  Queue `round & 1` holds the points whose
// distance dropped in the round before; mask `round & 1` has a bit per point in that queue, so a point enters it once.
template <typename QT>
__device__ void fps_run(FpsShared& sh, unsigned long long* D, QT* q0, QT* q1, unsigned* m0, unsigned* m1, int mbase, bool own_masks, int s, int P,
                        const FpsEdges* __restrict__ my_edges, const int* __restrict__ order, int want_start,
                        int n_samples, int* __restrict__ out, double* __restrict__ dist_or_null) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int t = threadIdx.x; t < P; t += FPS_T) D[t] = INF_BITS;
    if (own_masks)                                                   // workspace masks are cleared by the host (clouds share words)
        for (int t = threadIdx.x; t < (P + 31) / 32 + 1; t += FPS_T) { m0[t] = 0u; m1[t] = 0u; }
    if (threadIdx.x == 0) sh.start_pos = order != nullptr ? 0 : want_start;
    __syncthreads();
    if (order != nullptr)
        for (int t = threadIdx.x; t < P; t += FPS_T)
            if (order[s + t] - s == want_start) sh.start_pos = t;
    __syncthreads();
    int src = sh.start_pos;
    int round = 0;
    unsigned long long cv = 0ULL;                     // this thread's own maximum (value, point, original index), kept between samples
    int cidx = 0, co = 0x7fffffff;
    bool have = false;
    // a pass starts from `src` with distance zero, alone in the work queue (thread 0 sets it up, one barrier publishes it)
    auto begin_pass = [&](int from) {
        D[from] = 0ULL;
        ((round & 1) ? q1 : q0)[0] = (QT)from;
        atomicOr(((round & 1) ? m1 : m0) + ((mbase + from) >> 5), 1u << ((mbase + from) & 31));
        sh.cnt[round % 3] = 1; sh.cnt[(round + 1) % 3] = 0;
    };
    if (threadIdx.x == 0) { out[0] = want_start; begin_pass(src); }
    __syncthreads();

    for (int smp = 1; smp <= n_samples; ++smp) {
        if (smp == n_samples && dist_or_null == nullptr) break;      // the pass after the last sample only serves `dist`
        // one Dijkstra pass of the reference (sampling.cpp:56-81) into the shared D
        while (true) {
            const int F = sh.cnt[round % 3];
            if (F == 0) break;
            if (threadIdx.x == 0) sh.cnt[(round + 2) % 3] = 0;
            const QT* qc = (round & 1) ? q1 : q0;
            QT* qn = (round & 1) ? q0 : q1;
            unsigned* mc = (round & 1) ? m1 : m0;
            unsigned* mn = (round & 1) ? m0 : m1;
            int* cn = &sh.cnt[(round + 1) % 3];
            for (int it = threadIdx.x; it < F * FPS_K; it += FPS_T) {
                const int u = (int)qc[it / FPS_K], j = it % FPS_K;
                if (j == 0) atomicAnd(mc + ((mbase + u) >> 5), ~(1u << ((mbase + u) & 31)));     // free the bit for round + 2
                const int v = my_edges[u].to[j];
                const double du = __longlong_as_double(*(volatile unsigned long long*)(D + u));
                const unsigned long long nd = (unsigned long long)__double_as_longlong(__dadd_rn(du, my_edges[u].len[j]));
                if (nd < *(volatile unsigned long long*)(D + v)) {
                    const unsigned long long old = atomicMin(D + v, nd);
                    const unsigned bit = 1u << ((mbase + v) & 31);
                    if (nd < old && !(atomicOr(mn + ((mbase + v) >> 5), bit) & bit)) {
                        qn[atomicAdd(cn, 1)] = (QT)v;
                        // next round starts with v's edges: pull them from L2 while this round finishes
                        asm volatile("prefetch.global.L1 [%0];" :: "l"(my_edges + v));
                    }
                }
            }
            ++round;
            __syncthreads();
        }
        if (smp == n_samples) break;
        // farthest point: largest D, lowest original index among equals (Eigen maxCoeff, sampling.cpp:47).
        // Distances only ever decrease, so the maximum a thread found among its own points last time still stands unless
        // that very point was lowered since: one load instead of a scan for most threads.
        if (!(have && D[cidx] == cv)) {
            cv = 0ULL; cidx = threadIdx.x < P ? threadIdx.x : 0; co = 0x7fffffff;
            for (int t = threadIdx.x; t < P; t += FPS_T) {
                const unsigned long long v = D[t];
                if (v > cv) { cv = v; cidx = t; co = -1; }                              // original index fetched only if a tie needs it
                else if (v == cv) {
                    if (co == -1) co = orig_of(order, s + cidx);
                    const int o = orig_of(order, s + t);
                    if (o < co) { cidx = t; co = o; }
                }
            }
            have = threadIdx.x < P;
        }
        unsigned long long bv;
        int bidx, bo;
        warp_argmax(have, cv, cidx, co, order, s, bv, bidx, bo);
        if (lane == 0) { sh.red_v[warp] = bv; sh.red_i[warp] = bidx; sh.red_o[warp] = bo; }
        __syncthreads();
        // every warp finishes the reduction for itself: no second barrier, no broadcast
        {
            const bool mine = lane < FPS_T / 32 && sh.red_o[lane < FPS_T / 32 ? lane : 0] != 0x7fffffff;
            const unsigned long long pv = lane < FPS_T / 32 ? sh.red_v[lane] : 0ULL;
            const int pi = lane < FPS_T / 32 ? sh.red_i[lane] : 0;
            int po = lane < FPS_T / 32 ? sh.red_o[lane] : 0x7fffffff;
            warp_argmax(mine, pv, pi, po, nullptr, 0, bv, bidx, bo);
        }
        src = __shfl_sync(FULL, bidx, 0);
        if (threadIdx.x == 0) { out[smp] = bo - s; begin_pass(src); }
        __syncthreads();
    }
    if (dist_or_null != nullptr)
        for (int t = threadIdx.x; t < P; t += FPS_T) dist_or_null[orig_of(order, s + t)] = __longlong_as_double((long long)D[t]);
}

__global__ void __launch_bounds__(FPS_T, 2) fps_sample_kernel(const int* __restrict__ cloud_ptr, const FpsEdges* __restrict__ edges,
                                                              const int* __restrict__ order,
                                                              const int* __restrict__ start, int n_samples, int* __restrict__ samples,
                                                              double* __restrict__ dist_or_null, int smem_points,
                                                              unsigned long long* gD, int* gq0, int* gq1, unsigned* gm0, unsigned* gm1) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ FpsShared sh;
    const int b = blockIdx.x;
    const int s = cloud_ptr[b], P = cloud_ptr[b + 1] - s;
    int* out = samples + (size_t)b * n_samples;
    if (P <= 0) {
        for (int t = threadIdx.x; t < n_samples; t += FPS_T) out[t] = -1;
        return;
    }
    const int want = min(max(start[b], 0), P - 1);
    const FpsEdges* my_edges = edges + s;
    if (P <= smem_points) {
        unsigned long long* D = (unsigned long long*)smem_raw;
        unsigned short* q0 = (unsigned short*)(D + P);
        unsigned short* q1 = q0 + P;
        unsigned* m0 = (unsigned*)(q1 + P);
        unsigned* m1 = m0 + (P + 31) / 32 + 1;
        fps_run<unsigned short>(sh, D, q0, q1, m0, m1, 0, true, s, P, my_edges, order, want, n_samples, out, dist_or_null);
        return;
    }
    fps_run<int>(sh, gD + s, gq0 + s, gq1 + s, gm0, gm1, s, false, s, P, my_edges, order, want, n_samples, out, dist_or_null);
}

size_t align256(size_t n) { return (n + 255) / 256 * 256; }

}  // namespace

// ordered positions (fp64 and fp32) + boxes + graph (one 128-byte record per point), then the out-of-shared-memory state (D, two queues, two
// masks) that only clouds above 18,432 points use
extern "C" size_t dc_geodesic_fps_ws_bytes(int N) {
    const size_t n = (size_t)(N > 0 ? N : 0);
    const size_t tiles = (n + BOX_TILE - 1) / BOX_TILE;
    return align256(n * 24) + align256(tiles * BOX_TILE * 16) + align256(tiles * 32) + align256((tiles / SUPER + 1) * 32) + align256(n * sizeof(FpsEdges)) + align256(n * 8) +
           2 * align256(n * 4) + 2 * align256((n / 32 + 2) * 4);
}

extern "C" int dc_geodesic_fps(const double* pos, const int32_t* cloud_ptr, int B, int N, int max_cloud_points, int n_samples,
                               const int32_t* order_or_null, const int32_t* start, int32_t* samples, double* dist_or_null,
                               void* ws, size_t ws_bytes, void* stream) {
    DC_REQUIRE(B >= 1 && N >= 0 && n_samples >= 1, "dc_geodesic_fps: need B >= 1, N >= 0, n_samples >= 1");
    DC_REQUIRE(max_cloud_points >= 0 && max_cloud_points <= N, "dc_geodesic_fps: max_cloud_points must be in [0, N]");
    DC_REQUIRE(ws_bytes >= dc_geodesic_fps_ws_bytes(N), "dc_geodesic_fps: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t tiles = ((size_t)N + BOX_TILE - 1) / BOX_TILE;
    char* w = (char*)ws;
    double* spos = (double*)w; w += align256((size_t)N * 24);
    float4* spos32 = (float4*)w; w += align256(tiles * BOX_TILE * 16);
    float4* boxes = (float4*)w; w += align256(tiles * 32);
    float4* super_boxes = (float4*)w; w += align256((tiles / SUPER + 1) * 32);
    FpsEdges* edges = (FpsEdges*)w; w += align256((size_t)N * sizeof(FpsEdges));
    unsigned long long* gD = (unsigned long long*)w; w += align256((size_t)N * 8);
    int* gq0 = (int*)w; w += align256((size_t)N * 4);
    int* gq1 = (int*)w; w += align256((size_t)N * 4);
    unsigned* gm0 = (unsigned*)w; w += align256(((size_t)N / 32 + 2) * 4);
    unsigned* gm1 = (unsigned*)w;
    int launches = 1;
    if (N > 0) {
        fps_sort_box_kernel<<<dc_ceil_div((long long)tiles, 4), 128, 0, st>>>(pos, order_or_null, N, spos, spos32, boxes);
        fps_super_box_kernel<<<dc_ceil_div((long long)dc_ceil_div((long long)tiles, SUPER), 128), 128, 0, st>>>(boxes, (int)tiles, super_boxes);
        fps_knn_kernel<<<dc_ceil_div(N, KNN_WARPS * 32), KNN_WARPS * 32, 0, st>>>(spos, spos32, boxes, super_boxes, order_or_null, cloud_ptr,
                                                                                 B, N, edges);
        launches += 3;
    }
    // clouds up to `smem_points` keep their state in shared memory; larger ones (only if the caller's bound says they exist) in ws
    const int smem_points = max_cloud_points < FPS_SMEM_POINTS ? max_cloud_points : FPS_SMEM_POINTS;
    const size_t smem = fps_smem_bytes(smem_points) + 8;
    if (N > smem_points)      // a cloud that does not fit may exist (also when the caller's bound was too low): it needs clean masks
        DC_CUDA(cudaMemsetAsync(gm0, 0, 2 * align256(((size_t)N / 32 + 2) * 4), st));
    if (smem > 48 * 1024)
        DC_CUDA(cudaFuncSetAttribute(fps_sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(fps_smem_bytes(FPS_SMEM_POINTS) + 8)));
    fps_sample_kernel<<<B, FPS_T, smem, st>>>(cloud_ptr, edges, order_or_null, start, n_samples, samples, dist_or_null, smem_points,
                                              gD, gq0, gq1, gm0, gm1);
    return dc_check_launch_n("dc_geodesic_fps", launches);
}
