// K9 — tensor-core GEMM for the per-point channel mixing on sm_100a: tcgen05.mma (kind::tf32) issued
// by one thread, operands staged in shared memory by TMA (128-byte swizzle), accumulators in TMEM.
//
//   C[m, n] = sum_k A(m, k) * B(n, k)        fp32 in HBM on both sides, fp32 accumulation
//
// Accuracy.  The reference computes its Linears in fp32 and the parity bound on logits is 1e-4, so
// the default mode is a 3xTF32 split: x = hi + lo with hi = x & 0xffffe000 (exactly representable in
// TF32) and lo = x - hi, and  A B ~= A_hi B_hi + A_lo B_hi + A_hi B_lo  (the dropped lo*lo term is
// 2^-22 relative).  Four "converter" warps split every landed stage in shared memory (hi in place, lo
// into a second buffer) before the MMA warp consumes it.  SPLIT3 = false gives plain TF32 (1 MMA).
//
// Operand layouts.  "K-major": the reduction index is contiguous in HBM (X[M,K], W[N,K]); the smem
// tile is rows x 128 B.  "MN-major": the M/N index is contiguous (dY[M,Nout] and X[M,K] reduced over
// rows for the weight gradient; W[Nout,K] for the input gradient); the tile is fetched as 32-wide
// column chunks, each chunk 32 reduction rows x 128 B.  Both are canonical UMMA SWIZZLE_128B layouts,
// so no transposed copy of any activation is ever made.
//
// Roles (320 threads, persistent CTA per SM): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer,
// warps 2..5 = converters, warps 6..9 = epilogue (TMEM -> registers -> HBM; warp w owns TMEM lanes 32*(w%4)..);
// two TMEM accumulators so the epilogue of one tile overlaps the main loop of the next.
#include "common.cuh"
#include <cuda.h>
#include <cstring>

namespace {

constexpr int TC_BM = 128;          // UMMA_M (cta_group::1)
constexpr int TC_BK = 32;           // fp32 elements per k-block = one 128-byte swizzle span
constexpr int TC_THREADS = 320;       // TMA, MMA, 4 converter warps, 4 epilogue warps
constexpr int TC_MAX_STAGES = 4;

struct TcArgs {
    float* C; long long ldc;
    const float* bias;
    int M, N, K;
    int kb_per_split;       // k-blocks per grid.z slice
    float* partial;         // [splits][M][N] when split-K
    int keep_raw_hi;        // 1: leave the raw fp32 tile as the 'hi' B operand (valid iff the MMA truncates fp32 -> tf32)
    int extra_mma;          // measurement only (dc_set_gemm_debug): extra copies of every k-block's MMAs; results are then garbage
    int wide_mma;           // measurement only: the extra copies use N = 256 per instruction
    int tma_store;          // 1: the epilogue stages 32 x 32 blocks in shared memory and writes them with TMA (tmC is valid)
    long long* trace;       // debug: [2 CTAs][8 events][256 k-blocks] clock64 stamps of CTA 0 and its peer (pair kernel), or null
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    long long t0 = 0;
    do {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (!done) {                       // watchdog: a protocol bug must fault, not hang the GPU
            if (t0 == 0) t0 = clock64();
            else if (clock64() - t0 > 4000000000LL) __trap();
        }
    } while (!done);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// UMMA shared-memory descriptor, SWIZZLE_128B (cute::UMMA::SmemDescriptor: start>>4 [0,14), LBO>>4 [16,30),
// SBO>>4 [32,46), version=1 [46,48), layout_type=2 [61,64))
// layout_type: 2 = SWIZZLE_128B (16-byte chunks XOR row%8; K-major tiles),
//              1 = SWIZZLE_128B_BASE32B (32-byte chunks XOR row%4) — the only layout UMMA accepts for MN-major tf32
//                  operands (cutlass sm100_common.inl:92); TMA writes it with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout_type) {
    return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46) | ((uint64_t)layout_type << 61);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_c, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_c), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// A operand read from tensor memory (128 lanes x 8 columns of tf32 per instruction), B from shared memory
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_c, uint32_t tmem_a, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
        "}" ::"r"(tmem_c), "r"(tmem_a), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
                 "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
                   "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]),
                   "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]),
                   "r"(r[30]), "r"(r[31]) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                   "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
                   "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
                   "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// TMA store of one [32 rows][128 B] block (128-byte swizzled in shared memory) to coordinates (c0, c1, c2) of a 3-D tensor;
// elements outside the tensor are not written, which is what clips the ragged M / N edges.
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// Persistent kernel: every CTA walks the tile list (n fastest, so CTAs that run together share the A rows in L2)
// with five roles that only meet at mbarriers.  Two TMEM accumulators let the epilogue of tile t overlap the
// main loop of tile t+1.
template <int BN, bool A_MN, bool B_MN, bool SPLIT3, int STAGES, bool TS>
__global__ void __launch_bounds__(TC_THREADS, 1) tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                const __grid_constant__ CUtensorMap tmB,
                                                                const __grid_constant__ CUtensorMap tmC, TcArgs g) {
    constexpr uint32_t A_BYTES = TC_BM * 128, B_BYTES = BN * 128;
    static_assert(!TS || SPLIT3, "A-from-TMEM is the 3xTF32 path");
    // TS: the split A operand lives in tensor memory, so shared memory holds raw A (16 KB) + B hi/lo
    constexpr uint32_t STAGE_BYTES = TS ? A_BYTES + 2 * B_BYTES : (A_BYTES + B_BYTES) * (SPLIT3 ? 2 : 1);
    constexpr uint32_t CHUNK_BYTES = TC_BK * 128;                 // one 32-wide MN chunk of an MN-major tile
    constexpr uint32_t TMEM_NEED = 2 * BN + (TS ? 64 * STAGES : 0);   // 2 accumulators (+ A_hi/A_lo k-blocks per stage)
    constexpr uint32_t TMEM_COLS = TMEM_NEED <= 32 ? 32 : TMEM_NEED <= 64 ? 64 : TMEM_NEED <= 128 ? 128 : TMEM_NEED <= 256 ? 256 : 512;
    static_assert(TMEM_NEED <= 512, "tensor memory exhausted");
    constexpr uint32_t TMEM_A0 = 2 * BN;                          // first column of the A operand ring (TS)
    // instruction descriptor (cute::UMMA::InstrDescriptor): c=F32, a=b=TF32, majors, N>>3, M>>4
    constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | (((A_MN && !TS) ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                               ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);

    extern __shared__ unsigned char smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;          // swizzle atoms are 1024-byte aligned
    unsigned char* base_ptr = smem_raw + (base - smem_u32(smem_raw));
    constexpr uint32_t EPI_BYTES = 4 * 2 * 4096;                  // per epilogue warp: two 32 x 32 fp32 staging blocks
    const uint32_t epi_base = base + STAGES * STAGE_BYTES;
    const uint32_t bar_base = epi_base + EPI_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (TC_MAX_STAGES + s); };
    auto conv_bar = [&](int s) { return bar_base + 8u * (2 * TC_MAX_STAGES + s); };
    auto acc_full_bar = [&](int a) { return bar_base + 8u * (3 * TC_MAX_STAGES + a); };
    auto acc_empty_bar = [&](int a) { return bar_base + 8u * (3 * TC_MAX_STAGES + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (3 * TC_MAX_STAGES + 4);
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + (tmem_slot - base));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_n = (g.N + BN - 1) / BN, tiles_m = (g.M + TC_BM - 1) / TC_BM;
    const int total_kb = (g.K + TC_BK - 1) / TC_BK;
    const int splits = (total_kb + g.kb_per_split - 1) / g.kb_per_split;
    const long long num_tiles = (long long)tiles_n * tiles_m * splits;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); mbar_init(conv_bar(s), 4); }
        for (int a = 0; a < 2; ++a) { mbar_init(acc_full_bar(a), 1); mbar_init(acc_empty_bar(a), 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    auto stage_a = [&](int s) { return base + s * STAGE_BYTES; };
    auto stage_a_lo = [&](int s) { return base + s * STAGE_BYTES + A_BYTES; };
    auto stage_b = [&](int s) { return base + s * STAGE_BYTES + ((SPLIT3 && !TS) ? 2 : 1) * A_BYTES; };
    auto stage_b_lo = [&](int s) { return base + s * STAGE_BYTES + (TS ? 1 : 2) * A_BYTES + B_BYTES; };
    // tile id -> (m0, n0, first k-block, number of k-blocks)
    auto decode = [&](long long tile, int& m0, int& n0, int& kb_begin, int& nkb) {
        const int nt = (int)(tile % tiles_n);
        const long long r = tile / tiles_n;
        const int mt = (int)(r % tiles_m), z = (int)(r / tiles_m);
        m0 = mt * TC_BM; n0 = nt * BN;
        kb_begin = z * g.kb_per_split;
        nkb = min(total_kb, kb_begin + g.kb_per_split) - kb_begin;
    };

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
            uint32_t it = 0;
            for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                int m0, n0, kb_begin, nkb;
                decode(tile, m0, n0, kb_begin, nkb);
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(empty_bar(s), ph ^ 1u);
                    mbar_expect_tx(full_bar(s), A_BYTES + B_BYTES);
                    const int k0 = (kb_begin + kb) * TC_BK;
                    if (!A_MN) tma_load_2d(stage_a(s), &tmA, full_bar(s), k0, m0);
                    else
#pragma unroll
                        for (int c = 0; c < TC_BM / 32; ++c) tma_load_2d(stage_a(s) + c * CHUNK_BYTES, &tmA, full_bar(s), m0 + 32 * c, k0);
                    if (!B_MN) tma_load_2d(stage_b(s), &tmB, full_bar(s), k0, n0);
                    else
#pragma unroll
                        for (int c = 0; c < BN / 32; ++c) tma_load_2d(stage_b(s) + c * CHUNK_BYTES, &tmB, full_bar(s), n0 + 32 * c, k0);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (one thread: every instruction it executes besides the MMAs is tensor idle time) =====================
        uint32_t it = 0, tcount = 0;
        if (lane == 0)
        for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tcount) {
            int m0, n0, kb_begin, nkb;
            decode(tile, m0, n0, kb_begin, nkb);
            const uint32_t acc = tcount & 1, acc_ph = (tcount >> 1) & 1;
            mbar_wait(acc_empty_bar(acc), acc_ph ^ 1u);            // the epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t tmem_acc = tmem_base + acc * BN;
            for (int kb = 0; kb < nkb; ++kb, ++it) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                mbar_wait(SPLIT3 ? conv_bar(s) : full_bar(s), ph);
                tc_fence_after();
                {
                    if (TS)
                        for (int rep = 0; rep < g.extra_mma; ++rep)               // measurement only: extra copies of the MMAs (tensor rate)
#pragma unroll
                            for (int kk = 0; kk < TC_BK / 8; ++kk) {
                                const uint32_t b_off = B_MN ? kk * 1024u : kk * 32u;
                                const uint64_t db = umma_desc(stage_b(s) + b_off, B_MN ? CHUNK_BYTES : 16u, B_MN ? 512u : 1024u, B_MN ? 1u : 2u);
                                const uint32_t ta_hi = tmem_base + TMEM_A0 + (uint32_t)s * 64u + (uint32_t)kk * 8u;
                                umma_tf32_ts(tmem_acc, ta_hi, db, IDESC, 1u);
                                umma_tf32_ts(tmem_acc, ta_hi + 32u, db, IDESC, 1u);
                                umma_tf32_ts(tmem_acc, ta_hi, db, IDESC, 1u);
                            }
#pragma unroll
                    for (int kk = 0; kk < TC_BK / 8; ++kk) {          // UMMA_K = 8 for tf32
                        // K-major: +32 B inside the swizzle span; MN-major: next 8-row group (+1024 B)
                        const uint32_t a_off = A_MN ? kk * 1024u : kk * 32u, b_off = B_MN ? kk * 1024u : kk * 32u;
                        // K-major: 8-row groups 1024 B apart.  MN-major: 32-wide chunks CHUNK_BYTES apart (LBO),
                        // 4-row groups of the 32-byte-atom swizzle 512 B apart (SBO).
                        const uint32_t a_lbo = A_MN ? CHUNK_BYTES : 16u, b_lbo = B_MN ? CHUNK_BYTES : 16u;
                        const uint32_t a_sbo = A_MN ? 512u : 1024u, b_sbo = B_MN ? 512u : 1024u;
                        const uint32_t a_lt = A_MN ? 1u : 2u, b_lt = B_MN ? 1u : 2u;
                        const uint64_t db = umma_desc(stage_b(s) + b_off, b_lbo, b_sbo, b_lt);
                        if (TS) {
                            const uint32_t ta_hi = tmem_base + TMEM_A0 + (uint32_t)s * 64u + (uint32_t)kk * 8u, ta_lo = ta_hi + 32u;
                            const uint64_t db_lo = umma_desc(stage_b_lo(s) + b_off, b_lbo, b_sbo, b_lt);
                            umma_tf32_ts(tmem_acc, ta_hi, db, IDESC, (kb > 0 || kk > 0) ? 1u : 0u);
                            umma_tf32_ts(tmem_acc, ta_lo, db, IDESC, 1u);
                            umma_tf32_ts(tmem_acc, ta_hi, db_lo, IDESC, 1u);
                            continue;
                        }
                        const uint64_t da = umma_desc(stage_a(s) + a_off, a_lbo, a_sbo, a_lt);
                        umma_tf32(tmem_acc, da, db, IDESC, (kb > 0 || kk > 0) ? 1u : 0u);
                        if (SPLIT3) {
                            const uint64_t da_lo = umma_desc(stage_a_lo(s) + a_off, a_lbo, a_sbo, a_lt);
                            const uint64_t db_lo = umma_desc(stage_b_lo(s) + b_off, b_lbo, b_sbo, b_lt);
                            umma_tf32(tmem_acc, da_lo, db, IDESC, 1u);
                            umma_tf32(tmem_acc, da, db_lo, IDESC, 1u);
                        }
                    }
                    umma_commit(empty_bar(s));                        // frees the stage when the MMAs have read it
                    if (kb == nkb - 1) umma_commit(acc_full_bar(acc)); // accumulator complete
                }
            }
        }
    } else if (warp < 6) {
        // ===================== converters (3xTF32 split of every landed stage) =====================
        // (a second group of converter warps on alternate k-blocks was measured neutral here: at M = 128 per CTA the
        // tensor pipe, ~110 cycles per MMA, paces this kernel)
        if (SPLIT3) {
            const int t = threadIdx.x - 64;                           // 0..127
            uint32_t it = 0;
            for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                int m0, n0, kb_begin, nkb;
                decode(tile, m0, n0, kb_begin, nkb);
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(full_bar(s), ph);
                    float4* a_hi = reinterpret_cast<float4*>(base_ptr + (stage_a(s) - base));
                    float4* a_lo = reinterpret_cast<float4*>(base_ptr + (stage_a_lo(s) - base));
                    float4* b_hi = reinterpret_cast<float4*>(base_ptr + (stage_b(s) - base));
                    float4* b_lo = reinterpret_cast<float4*>(base_ptr + (stage_b_lo(s) - base));
                    const bool keep_raw = g.keep_raw_hi != 0;
                    auto split = [keep_raw](float4* hi, float4* lo, int i) {
                        float4 x = hi[i], h, l;
                        h.x = __uint_as_float(__float_as_uint(x.x) & 0xffffe000u); l.x = x.x - h.x;
                        h.y = __uint_as_float(__float_as_uint(x.y) & 0xffffe000u); l.y = x.y - h.y;
                        h.z = __uint_as_float(__float_as_uint(x.z) & 0xffffe000u); l.z = x.z - h.z;
                        h.w = __uint_as_float(__float_as_uint(x.w) & 0xffffe000u); l.w = x.w - h.w;
                        if (!keep_raw) hi[i] = h;
                        lo[i] = l;
                    };
                    if (TS) {
                        // this thread owns one row of the A tile: read it through the 128-byte swizzle, split it, and
                        // store hi / lo into the row's TMEM lane (32 columns each)
                        const int row = (warp & 3) * 32 + lane;
                        uint32_t hi[32], lo[32];
                        if (A_MN) {
                            // the tile was fetched unswizzled as 4 chunks of [32 reduction rows][32 columns]: this
                            // thread's operand row is a COLUMN of its chunk (lanes hit consecutive banks) — the
                            // transpose the MMA cannot do for an A operand in tensor memory happens here for free
                            const float* acol = reinterpret_cast<const float*>(base_ptr + (stage_a(s) - base) + (row >> 5) * CHUNK_BYTES) + (row & 31);
#pragma unroll
                            for (int kx = 0; kx < 32; ++kx) {
                                const float x = acol[kx * 32];
                                const uint32_t h = __float_as_uint(x) & 0xffffe000u;
                                hi[kx] = h;
                                lo[kx] = __float_as_uint(x - __uint_as_float(h));
                            }
                        } else {
                            const float4* arow = reinterpret_cast<const float4*>(base_ptr + (stage_a(s) - base) + row * 128);
#pragma unroll
                            for (int c = 0; c < 8; ++c) {
                                const float4 x = arow[c ^ (row & 7)];
                                const float v[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                                for (int u = 0; u < 4; ++u) {
                                    const uint32_t h = __float_as_uint(v[u]) & 0xffffe000u;
                                    hi[4 * c + u] = h;
                                    lo[4 * c + u] = __float_as_uint(v[u] - __uint_as_float(h));
                                }
                            }
                        }
                        const uint32_t ta = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + TMEM_A0 + (uint32_t)s * 64u;
                        tmem_st32(ta, hi);
                        tmem_st32(ta + 32u, lo);
                    } else {
#pragma unroll
                        for (int i = 0; i < (int)(A_BYTES / 16) / 128; ++i) split(a_hi, a_lo, t + 128 * i);
                    }
#pragma unroll
                    for (int i = 0; i < (int)(B_BYTES / 16) / 128; ++i) split(b_hi, b_lo, t + 128 * i);
                    if (TS) { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); tc_fence_before(); }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic writes -> visible to the MMA (async proxy)
                    __syncwarp();
                    if (lane == 0) mbar_arrive(conv_bar(s));
                }
            }
        }
    } else {
        // ===================== epilogue: TMEM -> registers -> HBM =====================
        const int q = warp & 3;                                   // TMEM lane quarter this warp may read
        uint32_t tcount = 0, epi_it = 0;
        for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tcount) {
            int m0, n0, kb_begin, nkb;
            decode(tile, m0, n0, kb_begin, nkb);
            const uint32_t acc = tcount & 1, acc_ph = (tcount >> 1) & 1;
            mbar_wait(acc_full_bar(acc), acc_ph);
            tc_fence_after();
            const int z = (int)(tile / ((long long)tiles_n * tiles_m));
            const int row = m0 + q * 32 + lane;
            float* out = g.partial ? g.partial + (size_t)z * g.M * g.N : g.C;
            const long long ldo = g.partial ? g.N : g.ldc;
            const bool vec_ok = (ldo % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15u) == 0);
            if (g.tma_store) {
                // Row-per-lane stores touch 32 different 128-byte lines per instruction (measured: a fifth of the L1 data pipe
                // of this kernel).  Instead each warp stages a 32 x 32 block in shared memory (its row as 8 swizzled 16-byte
                // chunks: conflict-free) and one lane hands the block to TMA, which writes full lines and clips the edges.
#pragma unroll 1
                for (int c0 = 0; c0 < BN; c0 += 32) {
                    if (n0 + c0 >= g.N) break;                        // warp-uniform
                    uint32_t r[32];
                    tmem_ld32(tmem_base + acc * BN + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, r);
                    if (!g.partial && g.bias) {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (n0 + c0 + j < g.N) r[j] = __float_as_uint(__uint_as_float(r[j]) + __ldg(g.bias + n0 + c0 + j));
                    }
                    const uint32_t buf = epi_base + (uint32_t)q * 8192u + (epi_it & 1u) * 4096u;
                    // the store that read this buffer two blocks ago must have finished reading it
                    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                    __syncwarp();
                    float4* dstrow = reinterpret_cast<float4*>(base_ptr + (buf - base) + lane * 128);
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        dstrow[c ^ (lane & 7)] = make_float4(__uint_as_float(r[4 * c]), __uint_as_float(r[4 * c + 1]),
                                                             __uint_as_float(r[4 * c + 2]), __uint_as_float(r[4 * c + 3]));
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) tma_store_3d(&tmC, buf, n0 + c0, m0 + q * 32, z);
                    ++epi_it;
                }
            } else {
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 16) {
                if (n0 + c0 >= g.N) break;                            // warp-uniform
                uint32_t r[16];
                const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                               "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                             : "r"(taddr));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (row < g.M) {
                    float* dst = out + (long long)row * ldo + n0 + c0;
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        const int col = n0 + c0 + 4 * v;
                        float4 o = make_float4(__uint_as_float(r[4 * v]), __uint_as_float(r[4 * v + 1]), __uint_as_float(r[4 * v + 2]), __uint_as_float(r[4 * v + 3]));
                        if (!g.partial && g.bias) {
                            if (col < g.N) o.x += __ldg(g.bias + col);
                            if (col + 1 < g.N) o.y += __ldg(g.bias + col + 1);
                            if (col + 2 < g.N) o.z += __ldg(g.bias + col + 2);
                            if (col + 3 < g.N) o.w += __ldg(g.bias + col + 3);
                        }
                        if (vec_ok && col + 3 < g.N) *reinterpret_cast<float4*>(dst + 4 * v) = o;
                        else {
                            if (col < g.N) dst[4 * v] = o.x;
                            if (col + 1 < g.N) dst[4 * v + 1] = o.y;
                            if (col + 2 < g.N) dst[4 * v + 2] = o.z;
                            if (col + 3 < g.N) dst[4 * v + 3] = o.w;
                        }
                    }
                }
            }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty_bar(acc));           // this warp's quarter of the accumulator is free
        }
        if (g.tma_store && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");     // staging buffers stay valid until read
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------------
// CTA-pair version of the 3xTF32 kernel (cta_group::2, UMMA_M = 256).  Measured on the single-CTA kernel above (ncu, large
// layers): the L1/shared-memory data pipe is ~85 % busy — converter loads/stores 36 %, tensor-core reads of B 27 %, epilogue
// stores 21 % — while the tensor pipe waits.  A pair of CTAs on the two SMs of a TPC computes a 256 x BN tile: each CTA
// holds its own 128 rows of A (tensor memory) and only HALF of the B tile (BN/2 rows), so per flop the converter traffic
// for B, the tensor-core reads of B and the L2 traffic for B are all halved.  Rank 0 issues the MMAs for both CTAs;
// stage release and accumulator-ready are multicast commits; "operands converted" and "accumulator drained" of the peer
// reach rank 0 as remote mbarrier arrivals.
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `cta` of the cluster (default .release.cta — the form of
// CUTLASS's ClusterBarrier::arrive; measured: .release.cluster costs the arriving thread ~1200 cycles more per arrive)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t local_bar, uint32_t cta) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_bar), "r"(cta));
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__device__ __forceinline__ void umma2_tf32_ts(uint32_t tmem_c, uint32_t tmem_a, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
        "}" ::"r"(tmem_c), "r"(tmem_a), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// completion of all MMAs issued so far -> arrive on the barrier at this offset in both CTAs of the pair
__device__ __forceinline__ void umma2_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}

// Latency, not bandwidth, bounds a ring whose slots stay occupied from the TMA request until the MMAs that read them have
// completed (commit -> refill -> land -> convert -> multiply is ~3 us, so 4 slots give one k-block per 0.75 us against
// 0.4 us of tensor work).  The pair kernel therefore uses three rings of different depth:
//   A ring   (PA_SLOTS x 16 KB): raw fp32 A tiles; a slot is released by the converters as soon as it is in tensor memory;
//   B ring   (PB_SLOTS x BN/2 rows x (raw + lo)): only 2 x 8 KB per k-block in pair mode; released by the MMAs;
//   TMEM ring (PT_SLOTS x 64 columns): A hi / lo; released by the MMAs.
// The MMA-issuing thread is the clock of the whole kernel: a tcgen05.mma occupies it for as long as the tensor pipe takes to
// accept the instruction (~54 cycles at M = 256, N = 128, measured), so every other instruction it executes is tensor idle
// time (a probe of the next hand-off's barrier issued ahead of the MMAs does not help: it queues behind them).  It therefore
// runs alone (no warp-wide waits), and signals "the MMAs of k-block it have completed" with a single
// commit per k-block on done[it % PD_SLOTS], which both the converters (tensor-memory slot it % PT_SLOTS) and the B
// producer (slot it % PB_SLOTS) wait on.
// Every hand-off between the roles costs the MMA thread a few hundred cycles (measured ~650 per hand-off against 69 per MMA),
// so one hand-off carries PK = 2 k-blocks of 32: a slot holds two [rows x 32] sub-tiles, 24 MMAs are issued per wait / commit.
constexpr int PK = 2;                // 32-wide k-blocks per slot
constexpr int PA_SLOTS = 3;
constexpr int PB_SLOTS = 3;
constexpr int PT_SLOTS = 2;
constexpr int PD_SLOTS = 4;          // >= PB_SLOTS and >= PT_SLOTS: nobody waits for a completion more than 4 slots old
constexpr int TCP_THREADS = 480;     // A producer, MMA, 2 x 4 converter warps (alternate k-blocks), 4 epilogue warps, B producer

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(TCP_THREADS, 1) tc_pair_gemm_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                      const __grid_constant__ CUtensorMap tmB,
                                                                      const __grid_constant__ CUtensorMap tmC, TcArgs g) {
    constexpr int BH = BN / 2;                                    // B rows held by one CTA
    constexpr uint32_t A_BYTES = TC_BM * 128, BH_BYTES = BH * 128;
    constexpr uint32_t A_SLOT_BYTES = PK * A_BYTES;               // PK raw [128 x 32] sub-tiles
    constexpr uint32_t B_SLOT_BYTES = 2 * PK * BH_BYTES;          // B half: PK raw (= hi) sub-tiles, then their PK lo sub-tiles
    constexpr uint32_t CHUNK_BYTES = TC_BK * 128;
    constexpr uint32_t TMEM_NEED = 2 * BN + 64 * PK * PT_SLOTS;
    constexpr uint32_t TMEM_COLS = TMEM_NEED <= 256 ? 256 : 512;
    static_assert(TMEM_NEED <= 512, "tensor memory exhausted");
    constexpr uint32_t TMEM_A0 = 2 * BN;
    // c=F32, a=b=TF32, A K-major (tensor memory), B major, N = BN, M = 256
    constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((B_MN ? 1u : 0u) << 16) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);

    extern __shared__ unsigned char smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    unsigned char* base_ptr = smem_raw + (base - smem_u32(smem_raw));
    constexpr uint32_t EPI_BYTES = 4 * 2 * 4096;
    const uint32_t b_base = base + PA_SLOTS * A_SLOT_BYTES;
    const uint32_t epi_base = b_base + PB_SLOTS * B_SLOT_BYTES;
    const uint32_t bar_base = epi_base + EPI_BYTES;
    // "Tile landed" barriers exist once per CONVERTER GROUP.  A parity wait is only meaningful on a barrier whose phases the
    // waiter observes one after the other: the two groups take alternate k-blocks, and with an odd number of slots a slot's
    // k-blocks alternate between the groups, so on a shared barrier a group would test phase n without ever having seen phase
    // n-1 complete — and the test for parity n&1 also succeeds while phase n-1 is still in flight (a load that lands late, e.g.
    // behind a TLB miss).  The group then consumed and RELEASED a slot whose load had not landed, the producer posted a second
    // expect_tx on a barrier with no pending arrival left, and the kernel faulted (seen once per ~10^5 launches).  With a
    // barrier per (group, slot) the k-blocks signalled on it (it = g mod 2, q mod SLOTS: every 2*SLOTS-th) all belong to one
    // group, which waits for each of them in turn.
    constexpr int NBAR = 3 * PA_SLOTS + 2 * PB_SLOTS + PT_SLOTS + PD_SLOTS + 4;
    static_assert(PA_SLOTS % 2 == 1 && PB_SLOTS % 2 == 1 && PT_SLOTS == 2, "phase arithmetic below: odd slot counts, one tensor-memory slot per converter group");
    auto a_full = [&](int grp, int q) { return bar_base + 8u * (grp * PA_SLOTS + q); };          // TMA bytes of the raw A tile
    auto a_empty = [&](int q) { return bar_base + 8u * (2 * PA_SLOTS + q); };                    // 4 local converter warps
    auto b_full = [&](int grp, int sb) { return bar_base + 8u * (3 * PA_SLOTS + grp * PB_SLOTS + sb); };
    auto ready = [&](int q) { return bar_base + 8u * (3 * PA_SLOTS + 2 * PB_SLOTS + q); };        // converters of both CTAs (rank 0's copy)
    auto done = [&](int d) { return bar_base + 8u * (3 * PA_SLOTS + 2 * PB_SLOTS + PT_SLOTS + d); }; // MMA commit (multicast to both CTAs)
    auto acc_full_bar = [&](int a) { return bar_base + 8u * (3 * PA_SLOTS + 2 * PB_SLOTS + PT_SLOTS + PD_SLOTS + a); };
    auto acc_empty_bar = [&](int a) { return bar_base + 8u * (3 * PA_SLOTS + 2 * PB_SLOTS + PT_SLOTS + PD_SLOTS + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * NBAR;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + (tmem_slot - base));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int tiles_n = (g.N + BN - 1) / BN, tiles_m = (g.M + TC_BM - 1) / TC_BM;
    const int pairs_m = (tiles_m + 1) / 2;
    const int total_kb = (g.K + TC_BK - 1) / TC_BK;
    const int splits = (total_kb + g.kb_per_split - 1) / g.kb_per_split;
    const long long num_units = (long long)tiles_n * pairs_m * splits;
    const long long first_unit = blockIdx.x >> 1, unit_stride = gridDim.x >> 1;
    long long* const trace = (g.trace && blockIdx.x < 2) ? g.trace + blockIdx.x * 8 * 256 : nullptr;   // CTA 0 and its peer
#define TC_TRACE(ev, it) do { if (trace && (it) < 256u) trace[(ev) * 256 + (it)] = clock64(); } while (0)

    if (threadIdx.x == 0) {
        for (int q = 0; q < PA_SLOTS; ++q) { mbar_init(a_full(0, q), 1); mbar_init(a_full(1, q), 1); mbar_init(a_empty(q), 4); }
        for (int sb = 0; sb < PB_SLOTS; ++sb) { mbar_init(b_full(0, sb), 1); mbar_init(b_full(1, sb), 1); }
        for (int q = 0; q < PT_SLOTS; ++q) mbar_init(ready(q), 8);
        for (int d = 0; d < PD_SLOTS; ++d) mbar_init(done(d), 1);
        for (int a = 0; a < 2; ++a) { mbar_init(acc_full_bar(a), 1); mbar_init(acc_empty_bar(a), 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                      // both CTAs' barriers and tensor memory exist before anything crosses the pair
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    if (trace && threadIdx.x == 0) trace[7 * 256 + 255] = clock64();        // right after the cluster barrier: clock offset between the two SMs

    auto slot_a = [&](int q) { return base + q * A_SLOT_BYTES; };
    auto slot_b = [&](int sb) { return b_base + sb * B_SLOT_BYTES; };
    auto slot_b_lo = [&](int sb) { return b_base + sb * B_SLOT_BYTES + PK * BH_BYTES; };
    // unit id -> (first row of this CTA's A tile, n0 of the pair's tile, first k-block, number of k-blocks, split)
    auto decode = [&](long long unit, int& m0, int& n0, int& kb_begin, int& nkb, int& z) {
        const int nt = (int)(unit % tiles_n);
        const long long rr = unit / tiles_n;
        const int mp = (int)(rr % pairs_m);
        z = (int)(rr / pairs_m);
        m0 = (2 * mp + (int)rank) * TC_BM; n0 = nt * BN;           // m0 >= M for the padding half of an odd tile count: TMA reads zeros
        kb_begin = z * g.kb_per_split;
        nkb = (min(total_kb, kb_begin + g.kb_per_split) - kb_begin + PK - 1) / PK;      // hand-offs of PK k-blocks (kb_per_split is a multiple
    };                                                                                   // of PK; beyond K the TMA reads zeros)

    if (warp == 0) {
        // ===================== A producer (this CTA's 128 rows) =====================
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
            uint32_t it = 0;
            for (long long unit = first_unit; unit < num_units; unit += unit_stride) {
                int m0, n0, kb_begin, nkb, z;
                decode(unit, m0, n0, kb_begin, nkb, z);
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int q = it % PA_SLOTS;
                    const uint32_t ph = (it / PA_SLOTS) & 1;
                    const uint32_t full = a_full((int)(it & 1u), q);      // the barrier of the converter group that takes this k-block
                    mbar_wait(a_empty(q), ph ^ 1u);
                    mbar_expect_tx(full, A_SLOT_BYTES);
#pragma unroll
                    for (int j = 0; j < PK; ++j) {
                        const int k0 = (kb_begin + kb * PK + j) * TC_BK;
                        if (!A_MN) tma_load_2d(slot_a(q) + j * A_BYTES, &tmA, full, k0, m0);
                        else
#pragma unroll
                            for (int c = 0; c < TC_BM / 32; ++c) tma_load_2d(slot_a(q) + j * A_BYTES + c * CHUNK_BYTES, &tmA, full, m0 + 32 * c, k0);
                    }
                }
            }
        }
    } else if (warp == 14) {
        // ===================== B producer (this CTA's half of the B tile) =====================
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
            uint32_t it = 0;
            for (long long unit = first_unit; unit < num_units; unit += unit_stride) {
                int m0, n0, kb_begin, nkb, z;
                decode(unit, m0, n0, kb_begin, nkb, z);
                const int nb0 = n0 + (int)rank * BH;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int sb = it % PB_SLOTS;
                    if (it >= (uint32_t)PB_SLOTS) {                     // the MMAs of the k-block that used this slot have completed
                        const uint32_t prev = it - PB_SLOTS;
                        mbar_wait(done(prev % PD_SLOTS), (prev / PD_SLOTS) & 1);
                    }
                    const uint32_t full = b_full((int)(it & 1u), sb);
                    mbar_expect_tx(full, PK * BH_BYTES);
#pragma unroll
                    for (int j = 0; j < PK; ++j) {
                        const int k0 = (kb_begin + kb * PK + j) * TC_BK;
                        if (!B_MN) tma_load_2d(slot_b(sb) + j * BH_BYTES, &tmB, full, k0, nb0);
                        else
#pragma unroll
                            for (int c = 0; c < BH / 32; ++c) tma_load_2d(slot_b(sb) + j * BH_BYTES + c * CHUNK_BYTES, &tmB, full, nb0 + 32 * c, k0);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (one thread of rank 0, for both CTAs) =====================
        if (leader && lane == 0) {
            uint32_t it = 0, tcount = 0;
            const uint32_t b_lbo = B_MN ? CHUNK_BYTES : 16u, b_sbo = B_MN ? 512u : 1024u, b_lt = B_MN ? 1u : 2u;
            for (long long unit = first_unit; unit < num_units; unit += unit_stride, ++tcount) {
                int m0, n0, kb_begin, nkb, z;
                decode(unit, m0, n0, kb_begin, nkb, z);
                const uint32_t acc = tcount & 1, acc_ph = (tcount >> 1) & 1;
                mbar_wait(acc_empty_bar(acc), acc_ph ^ 1u);        // both CTAs' epilogues have drained this accumulator
                tc_fence_after();
                const uint32_t tmem_acc = tmem_base + acc * BN;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int q = it % PT_SLOTS, sb = it % PB_SLOTS;
                    TC_TRACE(0, it);                                   // (rows 0/1 of the MMA thread's CTA are shared with the producers: debug only)
                    mbar_wait(ready(q), (it / PT_SLOTS) & 1);          // both CTAs: A in tensor memory, B lo written
                    TC_TRACE(1, it);
                    tc_fence_after();
                    TC_TRACE(2, it);
                    const uint64_t db0 = umma_desc(slot_b(sb), b_lbo, b_sbo, b_lt), dl0 = umma_desc(slot_b_lo(sb), b_lbo, b_sbo, b_lt);
                    const uint32_t ta0 = tmem_base + TMEM_A0 + (uint32_t)q * (64u * PK);
                    for (int rep = 0; rep < g.extra_mma; ++rep)               // measurement only: extra copies of the MMAs (tensor rate)
#pragma unroll
                        for (int kk = 0; kk < PK * (TC_BK / 8); ++kk) {
                            const uint64_t adv = (uint64_t)(((kk >> 2) * BH_BYTES + (B_MN ? (kk & 3) * 1024u : (kk & 3) * 32u)) >> 4);
                            const uint32_t ta_hi = ta0 + (uint32_t)(kk >> 2) * 64u + (uint32_t)(kk & 3) * 8u;
                            // wide_mma: N = 256 per instruction (reads past the B half and writes both accumulators: timing only)
                            const uint32_t idx = g.wide_mma ? ((IDESC & ~(0x3fu << 17)) | ((256u >> 3) << 17)) : IDESC;
                            const uint32_t tacc = g.wide_mma ? tmem_base : tmem_acc;
                            umma2_tf32_ts(tacc, ta_hi, db0 + adv, idx, 1u);
                            umma2_tf32_ts(tacc, ta_hi + 32u, db0 + adv, idx, 1u);
                            umma2_tf32_ts(tacc, ta_hi, db0 + adv, idx, 1u);
                        }
#pragma unroll
                    for (int kk = 0; kk < PK * (TC_BK / 8); ++kk) {
                        // sub-tile kk / 4; inside it the start address advances by 32 B (K-major) / one 8-row group (MN-major) per UMMA_K
                        const uint64_t adv = (uint64_t)(((kk >> 2) * BH_BYTES + (B_MN ? (kk & 3) * 1024u : (kk & 3) * 32u)) >> 4);
                        const uint32_t ta_hi = ta0 + (uint32_t)(kk >> 2) * 64u + (uint32_t)(kk & 3) * 8u, ta_lo = ta_hi + 32u;
                        umma2_tf32_ts(tmem_acc, ta_hi, db0 + adv, IDESC, (kb > 0 || kk > 0) ? 1u : 0u);
                        umma2_tf32_ts(tmem_acc, ta_lo, db0 + adv, IDESC, 1u);
                        umma2_tf32_ts(tmem_acc, ta_hi, dl0 + adv, IDESC, 1u);
                    }
                    umma2_commit(done(it % PD_SLOTS));                 // B slot and tensor-memory A slot of this k-block free in both CTAs
                    if (kb == nkb - 1) umma2_commit(acc_full_bar(acc));
                    TC_TRACE(3, it);
                }
            }
        }
    } else if (warp < 10) {
        // ===================== converters =====================
        // The chain load -> split -> tcgen05.st -> wait::st -> fence -> arrive of one k-block is longer than its 12 MMAs, and a
        // warp runs it serially: two groups of four warps take alternate k-blocks so that two chains are in flight.
        const int grp = (warp - 2) >> 2;
        const int t = (threadIdx.x - 64) & 127;
        uint32_t it = 0;
        for (long long unit = first_unit; unit < num_units; unit += unit_stride) {
            int m0, n0, kb_begin, nkb, z;
            decode(unit, m0, n0, kb_begin, nkb, z);
            for (int kb = 0; kb < nkb; ++kb, ++it) {
                if ((int)(it & 1u) != grp) continue;
                const int qa = it % PA_SLOTS, q = it % PT_SLOTS, sb = it % PB_SLOTS;
                // barrier (grp, slot) is signalled for every (2 * SLOTS)-th k-block, all of them this group's
                const uint32_t aph = (it / (2 * PA_SLOTS)) & 1, bph = (it / (2 * PB_SLOTS)) & 1;
                mbar_wait(a_full(grp, qa), aph);
                if ((warp & 3) == 2 && lane == 0) TC_TRACE(4, it);
                const int row = (warp & 3) * 32 + lane;
                if (it >= (uint32_t)PT_SLOTS) {                        // the MMAs that read this tensor-memory slot have completed
                    const uint32_t prev = it - PT_SLOTS;
                    mbar_wait(done(prev % PD_SLOTS), (prev / PD_SLOTS) & 1);
                }
                tc_fence_after();
                if ((warp & 3) == 2 && lane == 0) TC_TRACE(5, it);
#pragma unroll
                for (int j = 0; j < PK; ++j) {
                    uint32_t hi[32], lo[32];
                    if (A_MN) {
                        const float* acol = reinterpret_cast<const float*>(base_ptr + (slot_a(qa) - base) + j * A_BYTES + (row >> 5) * CHUNK_BYTES) + (row & 31);
#pragma unroll
                        for (int kx = 0; kx < 32; ++kx) {
                            const float x = acol[kx * 32];
                            const uint32_t h = __float_as_uint(x) & 0xffffe000u;
                            hi[kx] = h;
                            lo[kx] = __float_as_uint(x - __uint_as_float(h));
                        }
                    } else {
                        const float4* arow = reinterpret_cast<const float4*>(base_ptr + (slot_a(qa) - base) + j * A_BYTES + row * 128);
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const float4 x = arow[c ^ (row & 7)];
                            const float v[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                const uint32_t h = __float_as_uint(v[u]) & 0xffffe000u;
                                hi[4 * c + u] = h;
                                lo[4 * c + u] = __float_as_uint(v[u] - __uint_as_float(h));
                            }
                        }
                    }
                    const uint32_t ta = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + TMEM_A0 + (uint32_t)q * (64u * PK) + (uint32_t)j * 64u;
                    tmem_st32(ta, hi);
                    tmem_st32(ta + 32u, lo);
                }
                // The raw A slot goes back to the producer only now: the tcgen05.st above could not issue before every load
                // from the slot had returned its data (an arrive right after the loads can overtake them in the LSU queue).
                __syncwarp();
                if (lane == 0) mbar_arrive(a_empty(qa));
                // B half: lo next to the raw tile (the MMA truncates the raw fp32 words to tf32 itself)
                mbar_wait(b_full(grp, sb), bph);
                if ((warp & 3) == 2 && lane == 0) TC_TRACE(6, it);
                const float4* b_hi = reinterpret_cast<const float4*>(base_ptr + (slot_b(sb) - base));
                float4* b_lo = reinterpret_cast<float4*>(base_ptr + (slot_b_lo(sb) - base));
#pragma unroll
                for (int i = 0; i < (int)(PK * BH_BYTES / 16) / 128; ++i) {
                    const float4 x = b_hi[t + 128 * i];
                    float4 l;
                    l.x = x.x - __uint_as_float(__float_as_uint(x.x) & 0xffffe000u);
                    l.y = x.y - __uint_as_float(__float_as_uint(x.y) & 0xffffe000u);
                    l.z = x.z - __uint_as_float(__float_as_uint(x.z) & 0xffffe000u);
                    l.w = x.w - __uint_as_float(__float_as_uint(x.w) & 0xffffe000u);
                    b_lo[t + 128 * i] = l;
                }
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                tc_fence_before();
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
                    if (leader) mbar_arrive(ready(q));
                    else mbar_arrive_cluster(ready(q), 0);
                    if ((warp & 3) == 2) TC_TRACE(7, it);
                }
            }
        }
    } else if (warp < 14) {
        // ===================== epilogue: own 128 rows, TMEM -> shared memory -> TMA store =====================
        const int q = warp & 3;
        uint32_t tcount = 0, epi_it = 0;
        for (long long unit = first_unit; unit < num_units; unit += unit_stride, ++tcount) {
            int m0, n0, kb_begin, nkb, z;
            decode(unit, m0, n0, kb_begin, nkb, z);
            const uint32_t acc = tcount & 1, acc_ph = (tcount >> 1) & 1;
            mbar_wait(acc_full_bar(acc), acc_ph);
            tc_fence_after();
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                if (n0 + c0 >= g.N) break;
                uint32_t r[32];
                tmem_ld32(tmem_base + acc * BN + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, r);
                if (!g.partial && g.bias) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (n0 + c0 + j < g.N) r[j] = __float_as_uint(__uint_as_float(r[j]) + __ldg(g.bias + n0 + c0 + j));
                }
                const uint32_t buf = epi_base + (uint32_t)q * 8192u + (epi_it & 1u) * 4096u;
                if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                __syncwarp();
                float4* dstrow = reinterpret_cast<float4*>(base_ptr + (buf - base) + lane * 128);
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    dstrow[c ^ (lane & 7)] = make_float4(__uint_as_float(r[4 * c]), __uint_as_float(r[4 * c + 1]),
                                                         __uint_as_float(r[4 * c + 2]), __uint_as_float(r[4 * c + 3]));
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) tma_store_3d(&tmC, buf, n0 + c0, m0 + q * 32, z);
                ++epi_it;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (leader) mbar_arrive(acc_empty_bar(acc));
                else mbar_arrive_cluster(acc_empty_bar(acc), 0);
            }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                      // the peer may still read this CTA's shared / tensor memory through the pair's MMAs
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}
#undef TC_TRACE

// Split-K partials [splits][M*N] -> C (+ bias).  Fixed summation order for a given (shape, split count): deterministic.
// One thread per output; the loads of a round are independent so that four of them are in flight.
__global__ void tc_splitk_reduce_kernel(const float* __restrict__ partial, int splits, long long MN, int N, float* __restrict__ C,
                                        long long ldc, const float* __restrict__ bias) {
    long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (t >= MN) return;
    float s = 0.f;
    int z = 0;
    for (; z + 3 < splits; z += 4) {
        const float a = partial[(size_t)z * MN + t], b = partial[(size_t)(z + 1) * MN + t];
        const float c = partial[(size_t)(z + 2) * MN + t], d = partial[(size_t)(z + 3) * MN + t];
        s += (a + b) + (c + d);
    }
    for (; z < splits; ++z) s += partial[(size_t)z * MN + t];
    long long i = t / N; int j = (int)(t - i * N);
    if (bias) s += __ldg(bias + j);
    C[i * ldc + j] = s;
}

// Many splits of a small output (the weight gradients of the narrow layers): 8 threads share one output, each sums every 8th
// partial (four loads in flight), the 8 lane sums are combined in a fixed order through shared memory.
__global__ void __launch_bounds__(256) tc_splitk_reduce_wide_kernel(const float* __restrict__ partial, int splits, long long MN, int N,
                                                                    float* __restrict__ C, long long ldc, const float* __restrict__ bias) {
    __shared__ float sh[8][32];
    const int lane = threadIdx.x & 31, zl = threadIdx.x >> 5;
    const long long t = blockIdx.x * 32LL + lane;
    float s = 0.f;
    if (t < MN) {
        int z = zl;
        for (; z + 24 < splits; z += 32) {
            const float a = partial[(size_t)z * MN + t], b = partial[(size_t)(z + 8) * MN + t];
            const float c = partial[(size_t)(z + 16) * MN + t], d = partial[(size_t)(z + 24) * MN + t];
            s += (a + b) + (c + d);
        }
        for (; z < splits; z += 8) s += partial[(size_t)z * MN + t];
    }
    sh[zl][lane] = s;
    __syncthreads();
    if (zl == 0 && t < MN) {
        float r = ((sh[0][lane] + sh[1][lane]) + (sh[2][lane] + sh[3][lane])) + ((sh[4][lane] + sh[5][lane]) + (sh[6][lane] + sh[7][lane]));
        long long i = t / N; int j = (int)(t - i * N);
        if (bias) r += __ldg(bias + j);
        C[i * ldc + j] = r;
    }
}

// ---- host side ------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

// 2-D fp32 tensor map: inner (contiguous) extent x outer extent with row pitch ld elements; 128-byte swizzle;
// out-of-bounds elements read as zero (that is what pads ragged M / N / K edges).
// swizzle: 0 = none (tile only read by threads), 1 = 128B (K-major UMMA operand), 2 = 128B with 32-byte atoms (MN-major tf32 operand)
bool make_map(CUtensorMap* map, const float* ptr, long long inner, long long outer, long long ld, int box_outer, int swizzle) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    cuuint32_t box[2] = {32u, (cuuint32_t)box_outer};
    cuuint32_t estr[2] = {1u, 1u};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              swizzle == 0 ? CU_TENSOR_MAP_SWIZZLE_NONE : swizzle == 1 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B,
              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// 3-D fp32 tensor map for the output (N x M x Z, row pitch ld, plane pitch `plane` elements), 32 x 32 boxes, 128-byte swizzle
bool make_out_map(CUtensorMap* map, float* ptr, long long N, long long M, long long Z, long long ld, long long plane) {
    EncodeTiledFn fn = encode_fn();
    if (!fn || (ld % 4) || (plane % 4) || (reinterpret_cast<uintptr_t>(ptr) & 15u)) return false;
    cuuint64_t dims[3] = {(cuuint64_t)N, (cuuint64_t)M, (cuuint64_t)Z};
    cuuint64_t strides[2] = {(cuuint64_t)ld * 4, (cuuint64_t)plane * 4};
    cuuint32_t box[3] = {32u, 32u, 1u};
    cuuint32_t estr[3] = {1u, 1u, 1u};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int g_tc_extra_mma = 0, g_tc_wide_mma = 0;   // measurement only (dc_set_gemm_debug)
long long* g_tc_trace = nullptr;   // debug: device buffer of 8 x 256 clock stamps (dc_set_gemm_trace)
int g_tc_tma_store = 1;     // epilogue through shared memory + TMA stores (0 = row-per-lane direct stores)
int g_tc_keep_raw_hi = 1;   // B_hi is the raw fp32 tile: the MMA truncates fp32 -> tf32 on read (measured bit-identical to an explicit
                            // hi = x & 0xffffe000 rewrite, dc_set_gemm_keep_raw_hi(0) restores the rewrite)
int g_tc_no_ts = 0;      // experiment switch: 1 = keep the split A operand in shared memory (SS MMAs)

template <int BN, bool SPLIT3, bool TS> struct TcStages {
    static constexpr int per_stage = TS ? (TC_BM + 2 * BN) * 128 : (TC_BM + BN) * 128 * (SPLIT3 ? 2 : 1);
    static constexpr int by_smem = (200 * 1024) / per_stage > TC_MAX_STAGES ? TC_MAX_STAGES : (200 * 1024) / per_stage;
    static constexpr int by_tmem = TS ? (512 - 2 * BN) / 64 : TC_MAX_STAGES;      // A ring: 64 TMEM columns per stage
    static constexpr int value = by_smem < by_tmem ? by_smem : by_tmem;
};

template <int BN, bool A_MN, bool B_MN, bool SPLIT3>
int launch_tc(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const TcArgs& g, long long num_tiles, cudaStream_t st) {
    constexpr bool TS = SPLIT3;                // the split A operand lives in tensor memory instead of shared memory
    constexpr int STAGES = TcStages<BN, SPLIT3, TS>::value;
    static_assert(STAGES >= 2, "tile too large for shared memory");
    const size_t smem = (size_t)STAGES * TcStages<BN, SPLIT3, TS>::per_stage + 32768 /*epilogue staging*/ + 1024 /*alignment*/ + 8 * (3 * TC_MAX_STAGES + 4) + 16;
    if (g_tc_no_ts && TS) {
        constexpr int ST2 = TcStages<BN, SPLIT3, false>::value;
        const size_t smem2 = (size_t)ST2 * TcStages<BN, SPLIT3, false>::per_stage + 32768 + 1024 + 8 * (3 * TC_MAX_STAGES + 4) + 16;
        DC_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BN, A_MN, B_MN, SPLIT3, ST2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
        const long long sms2 = dc_sm_count();
        tc_gemm_kernel<BN, A_MN, B_MN, SPLIT3, ST2, false><<<(unsigned)(num_tiles < sms2 ? num_tiles : sms2), TC_THREADS, smem2, st>>>(tmA, tmB, tmC, g);
        return DC_OK;
    }
    DC_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BN, A_MN, B_MN, SPLIT3, STAGES, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long sms = dc_sm_count();
    const unsigned grid = (unsigned)(num_tiles < sms ? num_tiles : sms);       // persistent: one CTA per SM
    tc_gemm_kernel<BN, A_MN, B_MN, SPLIT3, STAGES, TS><<<grid, TC_THREADS, smem, st>>>(tmA, tmB, tmC, g);
    return DC_OK;
}

int g_tc_pair = 1;          // 1: CTA pairs (cta_group::2) for the 3xTF32 product when the shape allows

template <int BN, bool A_MN, bool B_MN>
int launch_tc_pair(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const TcArgs& g, int splits, cudaStream_t st) {
    const size_t smem = (size_t)PA_SLOTS * PK * TC_BM * 128 + (size_t)PB_SLOTS * PK * BN * 128 + 32768 + 1024 + 8 * (3 * PA_SLOTS + 2 * PB_SLOTS + PT_SLOTS + PD_SLOTS + 4) + 16;
    auto kern = tc_pair_gemm_kernel<BN, A_MN, B_MN>;
    DC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long tiles_n = (g.N + BN - 1) / BN, tiles_m = (g.M + TC_BM - 1) / TC_BM;
    const long long units = tiles_n * ((tiles_m + 1) / 2) * splits;
    const long long slots = dc_sm_count() / 2;                                  // persistent: one CTA pair per TPC
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((units < slots ? units : slots) * 2));
    cfg.blockDim = dim3(TCP_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    DC_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmC, g));
    return DC_OK;
}

template <bool A_MN, bool B_MN>
int dispatch_tc_pair(int BN, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& c, const TcArgs& g, int splits, cudaStream_t st) {
    if (BN == 64) return launch_tc_pair<64, A_MN, B_MN>(a, b, c, g, splits, st);
    return launch_tc_pair<128, A_MN, B_MN>(a, b, c, g, splits, st);
}

template <bool A_MN, bool B_MN>
int dispatch_tc(int BN, bool split3, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& c, const TcArgs& g, long long tiles, cudaStream_t st) {
    if (split3) {
        if (BN == 64) return launch_tc<64, A_MN, B_MN, true>(a, b, c, g, tiles, st);
        return launch_tc<128, A_MN, B_MN, true>(a, b, c, g, tiles, st);
    }
    if (BN == 64) return launch_tc<64, A_MN, B_MN, false>(a, b, c, g, tiles, st);
    return launch_tc<128, A_MN, B_MN, false>(a, b, c, g, tiles, st);
}

}  // namespace

// Precision mode of the tensor-core path: 0 = off (fp32 SIMT everywhere), 1 = TF32, 3 = 3xTF32 (default).
static int g_tc_mode = 3;
extern "C" int dc_set_gemm_mode(int mode) {
    DC_REQUIRE(mode == 0 || mode == 1 || mode == 3, "dc_set_gemm_mode: mode must be 0 (fp32 SIMT), 1 (TF32) or 3 (3xTF32)");
    g_tc_mode = mode;
    return DC_OK;
}
extern "C" int dc_get_gemm_mode(void) { return g_tc_mode; }
// experiment switch: A operand of the 3xTF32 split in shared memory (1) instead of tensor memory (0, default)
extern "C" int dc_set_gemm_no_ts(int v) { g_tc_no_ts = v; return DC_OK; }
// tuning switch: 1 (default) = do not rewrite the hi part of B in shared memory (the MMA truncates fp32 operands to tf32), 0 = rewrite
extern "C" int dc_set_gemm_keep_raw_hi(int v) { g_tc_keep_raw_hi = v; return DC_OK; }
// tuning switch: 1 (default) = epilogue through shared memory and TMA stores, 0 = row-per-lane direct stores
extern "C" int dc_set_gemm_tma_store(int v) { g_tc_tma_store = v; return DC_OK; }
// tuning switch: 1 (default) = CTA pairs (cta_group::2, 256-row tiles, B split across the pair), 0 = one CTA per tile
extern "C" int dc_set_gemm_pair(int v) { g_tc_pair = v; return DC_OK; }
// debug: device buffer (8 x 256 int64) that CTA 0 of the pair kernel fills with clock64 stamps of its pipeline events; null = off
extern "C" int dc_set_gemm_trace(void* buf) { g_tc_trace = (long long*)buf; return DC_OK; }
// measurement only: issue `extra_mma` additional copies of every k-block's MMAs (wide != 0: with N = 256 per instruction, pair
// kernel); the products are garbage while this is set — scripts/diag_tc_rate.py derives the cost of one tcgen05.mma from the slope
extern "C" int dc_set_gemm_debug(int extra_mma, int wide) {
    DC_REQUIRE(extra_mma >= 0 && extra_mma <= 64, "dc_set_gemm_debug: extra_mma must be in [0,64]");
    g_tc_extra_mma = extra_mma; g_tc_wide_mma = wide != 0;
    return DC_OK;
}

// Split-K factor for an M x N output with reduction K.  Cost model in units of one k-block (128 x BN x 32) of main loop: the
// persistent grid runs ceil(units / SMs) waves of (k-blocks per split + pipeline fill), the reduce pass costs a launch plus a
// term per split.  Picks the cheapest factor whose partials fit the workspace; the same function sizes the workspace.
int tc_choose_splits(long long M, int N, long long K, size_t ws_cap) {
    const int BN = N <= 64 ? 64 : 128;
    const long long total_kb = (K + TC_BK - 1) / TC_BK;
    const long long tiles = ((M + TC_BM - 1) / TC_BM) * ((N + BN - 1) / BN);
    const long long sms = dc_sm_count();
    if (total_kb < 16 || tiles >= 2 * sms) return 1;
    long long max_s = total_kb / 8;
    if (max_s > 256) max_s = 256;
    const double per_split = 0.2 + (double)M * N / 300000.0;
    double best_cost = 1e300;
    int best = 1;
    for (long long sp = 1; sp <= max_s; ++sp) {
        if (sp > 1 && (size_t)sp * M * N * sizeof(float) > ws_cap) break;
        const long long kb = (total_kb + sp - 1) / sp;
        const long long eff = (total_kb + kb - 1) / kb;              // splits actually produced
        const long long waves = (tiles * eff + sms - 1) / sms;
        const double cost = (double)waves * ((double)kb + 5.0) + (eff > 1 ? 8.0 + (double)eff * per_split : 0.0);
        if (cost < best_cost - 1e-9) { best_cost = cost; best = (int)sp; }
    }
    return best;
}

// Generic tensor-core product  C[M,N] = op(A) op(B)^T (+bias).  a_mn / b_mn select the operand layout:
//   a_mn = 0: A is [M, K] with K contiguous (lda = row pitch);  a_mn = 1: A is stored [K, M] with M contiguous.
//   b_mn = 0: B is [N, K] with K contiguous;                    b_mn = 1: B is stored [K, N] with N contiguous.
// Returns 1 (not an error) when the shape is not eligible, so the caller can use the SIMT kernel.
int dc_tc_gemm(const float* A, long long lda, int a_mn, const float* B, long long ldb, int b_mn, const float* bias,
               long long M, int N, long long K, float* C, long long ldc, void* ws, size_t ws_bytes, cudaStream_t st,
               int* launches) {
    *launches = 0;
    if (g_tc_mode == 0 || M < 64 || N < 16 || K < 8) return 1;
    if (lda % 4 || ldb % 4 || (reinterpret_cast<uintptr_t>(A) & 15u) || (reinterpret_cast<uintptr_t>(B) & 15u)) return 1;
    if (M >= 2147483647LL || K >= 2147483647LL) return 1;
    const int BN = N <= 64 ? 64 : 128;
    CUtensorMap tmA, tmB;
    // an MN-major A goes through tensor memory in the 3xTF32 path (the converters transpose it), so it is fetched unswizzled
    const bool a_via_tmem = g_tc_mode == 3 && !g_tc_no_ts;
    const int total_kb = (int)((K + TC_BK - 1) / TC_BK);
    const long long tiles_m_ = (M + TC_BM - 1) / TC_BM;
    const long long tiles = tiles_m_ * ((N + BN - 1) / BN);
    TcArgs g = {};
    g.C = C; g.ldc = ldc; g.bias = bias; g.M = (int)M; g.N = N; g.K = (int)K;
    g.keep_raw_hi = g_tc_keep_raw_hi;
    g.extra_mma = g_tc_extra_mma;
    g.wide_mma = g_tc_wide_mma;
    g.trace = g_tc_trace;
    CUtensorMap tmC;
    int splits = 1;
    // CTA pairs (two M tiles per unit): worth it when at most one tile in nine is padding; they need the TMA-store epilogue,
    // i.e. an output (or split-K partial buffer) with 16-byte aligned rows — otherwise the single-CTA kernel is planned
    bool pair = g_tc_pair && g_tc_tma_store && g_tc_mode == 3 && a_via_tmem && tiles_m_ >= 2 && (tiles_m_ % 2 == 0 || tiles_m_ >= 9);
    for (int attempt = 0; attempt < 2; ++attempt) {
        splits = tc_choose_splits(M, N, K, ws_bytes);
        g.kb_per_split = (total_kb + splits - 1) / splits;
        if (pair) g.kb_per_split = (g.kb_per_split + PK - 1) / PK * PK;   // the pair kernel hands over PK k-blocks at a time
        splits = (total_kb + g.kb_per_split - 1) / g.kb_per_split;
        g.partial = splits > 1 ? (float*)ws : nullptr;
        memset(&tmC, 0, sizeof(tmC));
        g.tma_store = g_tc_tma_store && (g.partial ? make_out_map(&tmC, g.partial, N, M, splits, N, M * (long long)N)
                                                   : make_out_map(&tmC, C, N, M, 1, ldc, M * ldc)) ? 1 : 0;
        if (!pair || g.tma_store) break;
        pair = false;
    }
    bool ok = a_mn ? make_map(&tmA, A, M, K, lda, TC_BK, a_via_tmem ? 0 : 2) : make_map(&tmA, A, K, M, lda, TC_BM, 1);
    ok = ok && (b_mn ? make_map(&tmB, B, N, K, ldb, TC_BK, 2) : make_map(&tmB, B, K, N, ldb, pair ? BN / 2 : BN, 1));
    if (!ok) return 1;
    const long long num_tiles = tiles * splits;
    const bool split3 = g_tc_mode == 3;
    int rc;
    if (pair) {
        if (!a_mn && !b_mn) rc = dispatch_tc_pair<false, false>(BN, tmA, tmB, tmC, g, splits, st);
        else if (!a_mn && b_mn) rc = dispatch_tc_pair<false, true>(BN, tmA, tmB, tmC, g, splits, st);
        else if (a_mn && b_mn) rc = dispatch_tc_pair<true, true>(BN, tmA, tmB, tmC, g, splits, st);
        else return 1;
    }
    else if (!a_mn && !b_mn) rc = dispatch_tc<false, false>(BN, split3, tmA, tmB, tmC, g, num_tiles, st);
    else if (!a_mn && b_mn) rc = dispatch_tc<false, true>(BN, split3, tmA, tmB, tmC, g, num_tiles, st);
    else if (a_mn && b_mn) rc = dispatch_tc<true, true>(BN, split3, tmA, tmB, tmC, g, num_tiles, st);
    else return 1;
    if (rc) return rc < 0 ? rc : -rc;
    *launches = 1;
    if (splits > 1) {
        const long long MN = M * N;
        if (splits >= 16 && MN <= (1LL << 20))
            tc_splitk_reduce_wide_kernel<<<dc_ceil_div(MN, 32), 256, 0, st>>>(g.partial, splits, MN, N, C, ldc, bias);
        else
            tc_splitk_reduce_kernel<<<dc_ceil_div(MN, 256), 256, 0, st>>>(g.partial, splits, MN, N, C, ldc, bias);
        *launches = 2;
    }
    return DC_OK;
}

// workspace for the split-K partials of a product with output M x N and reduction K
size_t dc_tc_ws_bytes(long long M, long long N, long long K) {
    if (M < 1 || N < 1 || K < 1 || N > 2147483647LL) return 0;
    const int sp = tc_choose_splits(M, (int)N, K, (size_t)-1);
    return sp > 1 ? (size_t)sp * M * N * sizeof(float) : 0;
}

// Raw access to the tensor-core product for tests and experiments (no SIMT fallback: returns 1 when the
// shape is not eligible).  Layout flags as in dc_tc_gemm above.
extern "C" int dc_gemm_tc_raw(const float* A, int64_t lda, int a_mn, const float* B, int64_t ldb, int b_mn,
                              const float* bias_or_null, int64_t M, int N, int64_t K, float* C, int64_t ldc,
                              void* ws, size_t ws_bytes, void* stream) {
    int n = 0;
    int rc = dc_tc_gemm(A, lda, a_mn, B, ldb, b_mn, bias_or_null, M, N, K, C, ldc, ws, ws_bytes, (cudaStream_t)stream, &n);
    if (rc == 1) { dc_set_error("dc_gemm_tc_raw: shape/layout not eligible for the tensor-core kernel"); return 1; }
    if (rc) return rc;
    return dc_check_launch_n("dc_gemm_tc_raw", n);
}
