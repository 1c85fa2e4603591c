// u8_batch.cuh — multi-query u8 scoring as an int8 contraction on the 5th-generation tensor cores.
//
// Q encoded queries x N encoded rows x K = actual_dim code bytes:  S[q][n] = sum_k qc[q][k] * vc[n][k]
// (the same integer impl_score_dot_* computes per pair, quantization/cpp/avx2.c:31-69), then
// score = multiplier * S + query.offset + vector_offset left to right without FMA
// (encoded_vectors_u8.rs:347) and a per-query top-k (demos/src/ann_benchmark_data.rs:150-167).
//
// One persistent CTA per SM, warp-specialised:
//   warp 0      TMA producer: 2-D tiled loads (cp.async.bulk.tensor, 128-byte swizzle) of a
//               128-query x 128-byte A block and a 256-row x 128-byte B block per pipeline stage
//   warp 1      MMA issuer: tcgen05.mma.cta_group::1.kind::i8, u8 x u8 -> s32, M=128 N=256 K=32 per
//               instruction, accumulators in TMEM (2 buffers x 256 columns), tcgen05.commit frees stages
//   warps 2..9  epilogue: tcgen05.ld (thread = TMEM lane = one query), threshold filter in the INTEGER domain
//               (one VIADDMNMX per accumulator against a per-query integer bound, see u8_batch.cu), exact
//               re-score of the few survivors in the reference's f32 order, survivors appended to the
//               query's candidate buffer in global memory
// Operands come from a device-side SoA mirror (codes[N][K] contiguous, offsets[N]) because the
// reference's AoS stride actual_dim + 4 is not a multiple of 16 bytes and cannot be a TMA stride.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace qz {

constexpr int kBatchM = 128;       // queries per tile (TMEM lanes)
constexpr int kBatchN = 256;       // corpus rows per tile (TMEM columns per accumulator buffer)
constexpr int kBatchKC = 128;      // code bytes per pipeline stage = one 128-byte swizzle row
constexpr int kBatchStages = 4;
constexpr int kEpiWarps = 8;       // epilogue warps: 2 per TMEM lane quarter, 128 accumulator columns each
constexpr int kQueueCap = 256;     // candidates an epilogue warp can stage in shared memory between flushes
constexpr int kVoffBufs = 4;       // ring of per-tile vector-offset blocks (1 KB each)
constexpr int kBatchThreads = (2 + kEpiWarps) * 32;  // producer, MMA, epilogue warps
constexpr uint32_t kStageA = kBatchM * kBatchKC;  // 16 KB
constexpr uint32_t kStageB = kBatchN * kBatchKC;  // 32 KB
constexpr uint32_t kStageBytes = kStageA + kStageB;
// stages + ring of per-tile vector offsets (f32) and their integer images (s32) + barriers + align slack + queues
constexpr size_t kBatchSmem = (size_t)kBatchStages * kStageBytes + 2 * kVoffBufs * kBatchN * 4 + 256 + 1024 + kEpiWarps * kQueueCap * 12;

struct U8BatchParams {
    uint32_t nq, k_chunks;       // queries; ceil(actual_dim / 128)
    uint32_t row_begin, row_end; // corpus rows of this launch (chunk)
    uint32_t n_qb, n_ct;         // query blocks, corpus tiles in the chunk
    const float *qoff;           // [nq]
    const float *voff;           // [count] vector offsets
    const int32_t *vq;           // [count] floor(vector offset / multiplier): the offsets in units of one integer step
    float multiplier, rmult;     // rmult = 1 / multiplier (integer filter only)
    float voff_absmax;           // max |vector offset| over the index
    uint32_t id_base, flip, cap;
    const uint64_t *thr;         // [nq] current k-th best key (0 = none yet)
    uint64_t *cand;              // [nq][cap]
    uint32_t *cnt;               // [nq]
    uint32_t *overflow;          // set when a candidate did not fit
#ifdef QZ_TUNING
    uint32_t dbg;                // experiment switches (tuning builds only): 1 = no filtering at all, 2 = no re-scoring
#endif
};

// bounded spin: a pipeline bug traps (launch failure) instead of hanging the device
__device__ __forceinline__ void mbar_wait_bounded(uint64_t *bar, uint32_t parity) {
    for (uint32_t i = 0; i < (1u << 26); i++)
        if (mbar_try_wait(bar, parity)) return;
    __trap();
}

__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int32_t c0, int32_t c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}

// shared-memory matrix descriptor: K-major, 128-byte swizzle, 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t umma_desc_sw128(const void *smem_ptr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_u32(smem_ptr) & 0x3FFFFu) >> 4);  // start address, bits [0,14)
    d |= (uint64_t)1 << 16;                                  // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                        // stride byte offset, bits [32,46)
    d |= (uint64_t)1 << 46;                                  // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                                  // SWIZZLE_128B
    return d;
}

// instruction descriptor for kind::i8: D = s32, A = B = unsigned 8-bit, both K-major, M x N
__host__ __device__ constexpr uint32_t umma_idesc_u8(uint32_t M, uint32_t N) {
    return (2u << 4) | (0u << 7) | (0u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}
// the same MMA with the descriptors given as their low words (start address, the only part that changes between
// stages and K steps) and one shared high word: a 32-bit add per instruction in the single issuing thread
__device__ __forceinline__ void umma_i8_lo(uint32_t tmem_d, uint32_t adesc_lo, uint32_t bdesc_lo, uint32_t desc_hi,
                                           uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], da, db, %4, {%6, %6, %6, %6}, p;\n\t}" ::"r"(tmem_d),
        "r"(adesc_lo), "r"(bdesc_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}
__device__ __forceinline__ uint32_t umma_desc_sw128_lo(const void *smem_ptr) {
    return ((smem_u32(smem_ptr) & 0x3FFFFu) >> 4) | (1u << 16);
}
constexpr uint32_t kUmmaDescSw128Hi = (1024u >> 4) | (1u << 14) | (2u << 29);  // SBO 1024 B, version 1, SWIZZLE_128B
// one lane of a converged warp (always the same one for the full mask)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// multicast forms for thread-block clusters: the B block of a stage is loaded once from L2 and delivered to the
// same shared-memory offset of every CTA in `mask`; a stage is released in every CTA of the cluster
__device__ __forceinline__ void tma_load_2d_mc(void *dst, const CUtensorMap *map, int32_t c0, int32_t c1, uint64_t *bar,
                                               uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, "
        "%4}], [%2], %5;" ::"r"(smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
        : "memory");
}
__device__ __forceinline__ void umma_commit_mc(uint64_t *bar, uint16_t mask) {
    asm volatile(
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
            smem_u32(bar)),
        "h"(mask)
        : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
// the registers are tied to the wait so that no use of them can be scheduled above it
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                   "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]),
                   "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]),
                   "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
                 :
                 : "memory");
}

// wait for three outstanding 32-column loads; all 96 registers are tied to the wait
__device__ __forceinline__ void tmem_ld_wait3(uint32_t (&a)[32], uint32_t (&b)[32], uint32_t (&c)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    tmem_ld_wait(a);  // (nothing is outstanding any more: these only pin the registers behind the wait)
    tmem_ld_wait(b);
    tmem_ld_wait(c);
}

}  // namespace qz
