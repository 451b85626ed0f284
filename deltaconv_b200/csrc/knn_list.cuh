// Warp-distributed sorted candidate list shared by the brute-force (knn.cu) and the grid (knn_grid.cu) kNN kernels.
// The list holds SLOTS * 32 keys, position p = slot * 32 + lane, ascending; a key is (d2 bits << 32) | local index, so
// ties resolve to the lower index and results are bit-reproducible.
#pragma once
#include "common.cuh"

struct MergeMin { int one_slot, two_slots; };      // survivors per round from which a merge replaces insertions, per slot count

namespace {

typedef unsigned long long u64;
constexpr u64 KEY_INF = ~0ull;

template <int SLOTS>
__device__ __forceinline__ u64 list_get(const u64 (&keys)[SLOTS], int p) {
    u64 r = __shfl_sync(FULL, keys[0], p & 31);
#pragma unroll
    for (int s = 1; s < SLOTS; ++s) {
        u64 t = __shfl_sync(FULL, keys[s], p & 31);
        if ((p >> 5) == s) r = t;
    }
    return r;
}

template <int SLOTS>
__device__ __forceinline__ void list_insert(u64 (&keys)[SLOTS], u64 c, int lane) {
    int pos = 0;
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) pos += __popc(__ballot_sync(FULL, keys[s] < c));
#pragma unroll
    for (int s = SLOTS - 1; s >= 0; --s) {
        u64 up = __shfl_up_sync(FULL, keys[s], 1);
        if (s > 0) {
            u64 carry = __shfl_sync(FULL, keys[s - 1], 31);
            if (lane == 0) up = carry;
        }
        int p = s * 32 + lane;
        if (p == pos) keys[s] = c;
        else if (p > pos) keys[s] = up;
    }
}

__device__ __forceinline__ u64 shfl_xor64(u64 v, int m) { return __shfl_xor_sync(FULL, v, m); }

// One compare-exchange step of a bitonic network over the 32 lanes: the lower lane of each pair keeps the minimum when the
// pair's block is ascending.
__device__ __forceinline__ u64 bitonic_step(u64 key, int lane, int stride, bool ascending) {
    const u64 other = shfl_xor64(key, stride);
    const bool lower = (lane & stride) == 0;
    const u64 lo = key < other ? key : other, hi = key < other ? other : key;
    return (lower == ascending) ? lo : hi;
}

// Merge 32 new candidates into the sorted 32-entry list (k <= 32): sort them across the warp (15 steps), pair the list with
// their reverse (the element-wise minimum is the bitonic lower half of the union), sort that (5 steps).  About 170
// instructions whatever the number of survivors, against ~22 per single insertion: used when a round has many survivors,
// i.e. in the first few rounds of every query.  Keys are unique (they carry the index), so the result is the same list.
__device__ __forceinline__ u64 list_merge32(u64 list, u64 key, int lane) {
#pragma unroll
    for (int size = 2; size <= 32; size <<= 1)
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1)
            key = bitonic_step(key, lane, stride, size == 32 || (lane & size) == 0);
    const u64 rev = __shfl_sync(FULL, key, 31 - lane);
    u64 m = list < rev ? list : rev;
#pragma unroll
    for (int stride = 16; stride > 0; stride >>= 1) m = bitonic_step(m, lane, stride, true);
    return m;
}

// The same for the 64-entry list of 32 < k <= 64 (two keys per lane, position = slot * 32 + lane): sort the 32 candidates
// (15 steps); the 64 smallest of list ∪ candidates are (slot 0, min(slot 1, reversed candidates)) — the candidates padded
// with +inf to 64 and reversed only meet the upper half — which is a bitonic sequence of 64: one register-level
// compare-exchange between the two slots (stride 32), then 5 cross-lane steps on each slot.
__device__ __forceinline__ void list_merge64(u64 (&keys)[2], u64 key, int lane) {
#pragma unroll
    for (int size = 2; size <= 32; size <<= 1)
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1)
            key = bitonic_step(key, lane, stride, size == 32 || (lane & size) == 0);
    const u64 rev = __shfl_sync(FULL, key, 31 - lane);
    const u64 upper = keys[1] < rev ? keys[1] : rev;
    u64 lo = keys[0] < upper ? keys[0] : upper, hi = keys[0] < upper ? upper : keys[0];
#pragma unroll
    for (int stride = 16; stride > 0; stride >>= 1) {
        lo = bitonic_step(lo, lane, stride, true);
        hi = bitonic_step(hi, lane, stride, true);
    }
    keys[0] = lo;
    keys[1] = hi;
}

// survivors in a round from which the merge is cheaper than single insertions.  One slot per lane: 5 / 8 / 12 / 16 / 24
// measured, 12 is best.  Two slots per lane: an insertion costs about twice as much (two 64-bit registers shifted per
// lane), the merge only 26 steps instead of 20 — the threshold is lower (dc_set_knn_merge_min tunes both at run time).


// Offer one key per lane to the list (KEY_INF = no candidate); tau is the current k-th best.
template <int SLOTS>
__device__ __forceinline__ void list_offer(u64 (&keys)[SLOTS], u64& tau, int k, int lane, u64 key, MergeMin mm) {
    unsigned mask = __ballot_sync(FULL, key < tau);
    if (SLOTS <= 2 && __popc(mask) >= (SLOTS == 1 ? mm.one_slot : mm.two_slots)) {
        if constexpr (SLOTS == 1) keys[0] = list_merge32(keys[0], key, lane);
        else if constexpr (SLOTS == 2) list_merge64(keys, key, lane);
        tau = list_get<SLOTS>(keys, k - 1);
        return;
    }
    while (mask) {
        int b = __ffs(mask) - 1;
        u64 c = __shfl_sync(FULL, key, b);
        list_insert<SLOTS>(keys, c, lane);
        tau = list_get<SLOTS>(keys, k - 1);
        mask = __ballot_sync(FULL, key < tau) & ~((2u << b) - 1u);
    }
}

}  // namespace
