// pq_scan2.cu — EncodedVectorsPQ full scan, streaming version.
//
// score_point for every row (encoded_vectors_pq.rs:405-494): score = sum over chunks of lut[ch*256 + code[ch]].
// The query LUT (n_chunks x 256 f32, 768 KiB at 1536-d / chunk 2) does not fit one SM's shared memory, so the
// chunk axis is cut into slices of 128 chunks (128 KiB of LUT, bank-permuted by pq_lut_permute_kernel: the four
// gathers of code word w of a row-slice fall in bank w, whatever the codes; layout at pq2_gather below).
//
// One persistent CTA per SM owns blocks of `block_rows` consecutive rows.  For a block it walks the slices in
// order: the slice's LUT is bulk-copied into shared memory, four producer lanes (one per consumer group) stream
// the block's 128-byte code sub-rows through a ring of 64-row tiles (2-D TMA tiles of the row-major code matrix,
// completing on mbarriers), sixteen consumer warps gather and add (one row per lane, pq2_rows16), and the per-row
// sums accumulate in shared memory in slice order.  After the last slice the block's scores are final: they are written out (4 B per row) for the top-k
// pass, and/or (k <= 128) reduced on the spot to the CTA's running top-k list: keys above the current threshold are
// compacted into the (idle) LUT area and reduced with the cooperative register sorting networks of topk.cuh.
// HBM traffic = the codes, once; nothing else leaves the SM but the final scores or the k best keys.
// (The first version read codes with plain loads, kept per-slice partial sums in global memory and was
// latency-bound at 2.2 TB/s)
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "qz_internal.h"
#include "topk.cuh"
#include "peer_exchange.cuh"
#include "tensor_map.h"

namespace qz {

constexpr int kPq2SliceChunks = 128;
constexpr int kPq2SliceFloats = kPq2SliceChunks * 256;  // 128 KiB
constexpr int kPq2TileRows = 64;
constexpr int kPq2TileBytes = kPq2TileRows * kPq2SliceChunks;  // 8 KiB
constexpr int kPq2Warps = 16;                                   // consumers; warp 16 = producer
constexpr int kPq2Groups = 4;                                   // 4 warps x 16 rows share a tile
constexpr int kPq2Threads = (kPq2Warps + 1) * 32;

struct PqScan2Params {
    uint64_t count;
    uint32_t n_slices, n_slots;
    uint32_t block_rows;  // multiple of 64
    uint32_t n_blocks;
    const float *perm;  // [n_slices][kPq2SliceFloats]
    float *scores;      // optional [count]
    uint64_t *partial;  // optional [gridDim.x][128]: the CTA's 128 best keys, best first (fused top-k, k <= 128)
    uint32_t k, flip, id_base;
    // in-launch final merge + optional cross-GPU exchange, as in row_scan.cuh (done != NULL: last CTA merges)
    uint32_t *done;
    MergeParams merge;
    PeerExchange px;
};

__device__ __forceinline__ void pq2_tma_tile(void *dst, const CUtensorMap *map, int32_t c0, int32_t c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}

// One LUT gather.  The slice's entry (b, code) of lane w's column sits at byte
//   (b >> 1) * 65536 + code * 256 + (b & 1) * 128 + w * 4        (pq.cu pq_lut_permute_kernel)
// so the variable part of the address, code << 8 | w * 4, is ONE byte permute of the code word with a per-lane
// constant, and the rest is the load's immediate.  FAST: the per-lane constant also carries the upper half of the
// LUT's shared-memory address and the immediate its bits 8..15 (kPq2LutBase: user shared memory starts 1 KiB into
// the CTA's window on sm_100), which leaves no add at all; the kernel checks the address and takes the other
// form (one add per gather) if the assumption does not hold.
constexpr uint32_t kPq2LutBase = 0x400;
template <bool FAST, int B>
__device__ __forceinline__ float pq2_gather(uint32_t cw, uint32_t lane_const, uint32_t lut_s) {
    float v;
    if (FAST) {
        const uint32_t a = __byte_perm(cw, lane_const, 0x7604 | (B << 4));
        asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(a), "n"(kPq2LutBase + (B >> 1) * 65536 + (B & 1) * 128));
    } else {
        const uint32_t a = lut_s + __byte_perm(cw, lane_const, 0x5504 | (B << 4));
        asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(a), "n"((B >> 1) * 65536 + (B & 1) * 128));
    }
    return v;
}

// 16 rows x 128 chunks of one warp.  Lane L works on ONE row (L & 15) and on the 16 code words 16*(L >> 4) ..+15 of
// it, in the order word = 16*(L >> 4) + ((L & 15) ^ j), j = 0..15: at every step the 32 lanes read 32 different word
// columns (code loads and LUT gathers both conflict free), and a row's 128 lookups are spread over just two lanes,
// so the whole cross-lane reduction is one shuffle (the first version gave each lane one column of all 16 rows and
// paid a 16-value transposing butterfly: 16 shuffles + 30 selects per 16 rows, on the same pipe as the gathers).
// lc0 = the lane constant of pq2_gather for j = 0; step j uses lc0 ^ 4j.  Returns the row's slice sum.
template <bool FAST>
__device__ __forceinline__ float pq2_rows16(const uint32_t (&cw)[16], uint32_t lc0, uint32_t lut_s) {
    // The words arrive in the order w = (L & 15) ^ j, which depends on the row's position.  The sum must not: a row
    // scores the same wherever it sits (a shard of the corpus and the whole corpus agree bit for bit).  A balanced
    // tree over j that pairs j with j ^ 1, then ^ 2, ^ 4, ^ 8 pairs the same WORDS whatever the xor mask, and a + b
    // is commutative, so its value is that of the same tree over w.
    float w[16];
#pragma unroll
    for (int j = 0; j < 16; j++) {
        const uint32_t lc = lc0 ^ (4u * j);
        const float g0 = pq2_gather<FAST, 0>(cw[j], lc, lut_s), g1 = pq2_gather<FAST, 1>(cw[j], lc, lut_s);
        const float g2 = pq2_gather<FAST, 2>(cw[j], lc, lut_s), g3 = pq2_gather<FAST, 3>(cw[j], lc, lut_s);
        w[j] = __fadd_rn(__fadd_rn(g0, g1), __fadd_rn(g2, g3));
    }
#pragma unroll
    for (int d = 1; d < 16; d <<= 1) {
#pragma unroll
        for (int j = 0; j < 16; j += 2 * d) w[j] = __fadd_rn(w[j], w[j + d]);
    }
    const float s = w[0];
    return __fadd_rn(s, __shfl_xor_sync(0xffffffffu, s, 16));
}

// position of a producer lane in its group's tile sequence: t = lane, lane + 4, ... of every (block, slice)
struct Pq2Cursor {
    uint32_t lb, s, t;
    __device__ __forceinline__ bool valid(uint32_t my_blocks) const { return lb < my_blocks; }
    __device__ __forceinline__ void next(uint32_t tiles_pb, uint32_t n_slices, uint32_t g) {
        t += kPq2Groups;
        if (t >= tiles_pb) {
            t = g;
            if (++s == n_slices) {
                s = 0;
                lb++;
            }
        }
    }
};

__global__ void __launch_bounds__(kPq2Threads, 1) pq_scan2_kernel(const __grid_constant__ CUtensorMap map, const PqScan2Params p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    float *slut = reinterpret_cast<float *>(smem);
    uint8_t *ring = smem + (size_t)kPq2SliceFloats * 4;
    float *acc = reinterpret_cast<float *>(ring + (size_t)p.n_slots * kPq2TileBytes);
    uint64_t *full = reinterpret_cast<uint64_t *>(acc + p.block_rows);
    uint64_t *empty = full + p.n_slots;
    uint64_t *lutbar = empty + p.n_slots;
    uint64_t *top = lutbar + 1;                                  // [128] running best keys of this CTA
    uint32_t *n_cand = reinterpret_cast<uint32_t *>(top + 128);  // append cursor of the candidate area
    const bool do_topk = p.partial != nullptr;

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < p.n_slots; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kPq2Warps / kPq2Groups);
        }
        mbar_init(lutbar, 1);
        fence_mbar_init();
    }
    if (threadIdx.x < 128) top[threadIdx.x] = 0ull;
    __syncthreads();

    const uint32_t tiles_pb = p.block_rows / kPq2TileRows;  // tiles per (block, slice)
    const uint32_t my_blocks = blockIdx.x < p.n_blocks ? (p.n_blocks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

    if (warp == kPq2Warps) {
        // ---------------- producers ----------------
        // Lane g streams the tiles of consumer group g (t = g, g + 4, ... of every (block, slice)) through the group's
        // own two slots, g and g + 4: the groups' pipelines are independent, so a group that falls behind does not
        // hold back the loads of the others, and a slot only ever serves one group, whatever the tile count.
        // (Pulling tiles into L2 ahead of the ring with cp.async.bulk.prefetch.tensor made the scan 14 % slower; launching
        // this kernel as a programmatic dependent of the LUT permutation changed nothing measurable.)
        if (lane < kPq2Groups && lane < tiles_pb) {
            Pq2Cursor cur{0, 0, lane};
            for (uint32_t n = 0; cur.valid(my_blocks); n++) {
                const uint32_t slot = lane + kPq2Groups * (n & 1), use = n >> 1;
                if (use > 0) mbar_wait(&empty[slot], (use - 1) & 1);
                mbar_arrive_expect_tx(&full[slot], kPq2TileBytes);
                // rows past `count` are out of bounds for the tensor map: zero filled, never stored
                pq2_tma_tile(ring + (size_t)slot * kPq2TileBytes, &map, (int32_t)(cur.s * kPq2SliceChunks),
                             (int32_t)((uint64_t)(blockIdx.x + cur.lb * gridDim.x) * p.block_rows + (uint64_t)cur.t * kPq2TileRows),
                             &full[slot]);
                cur.next(tiles_pb, p.n_slices, lane);
            }
        }
        return;
    }

    // ---------------- consumers ----------------
    const uint32_t group = warp / (kPq2Warps / kPq2Groups), member = warp % (kPq2Warps / kPq2Groups);
    const uint32_t lut_s = smem_u32(slut);
    const bool fast = (lut_s & 0xFFFFu) == kPq2LutBase;
    const uint32_t rsub = lane & 15u;                                  // this lane's row within the warp's 16
    const uint32_t col0 = (lane & 16u) + rsub;                          // its word column at step 0
    const uint32_t lc0 = (fast ? (lut_s & 0xFFFF0000u) : 0u) | col0 * 4u;
    uint32_t n = 0;         // tiles this warp's group has consumed (the group's slots alternate: group, group + 4)
    uint32_t lut_phase = 0;
    for (uint32_t lb = 0; lb < my_blocks; lb++) {
        const uint64_t row0 = (uint64_t)(blockIdx.x + lb * gridDim.x) * p.block_rows;
        for (uint32_t s = 0; s < p.n_slices; s++) {
            // every consumer is done with the previous slice's LUT (and, at s == 0, has stored the previous block)
            // (the top-k phase writes candidate keys into the LUT area with ordinary stores: order them before the
            // bulk copy's writes)
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            named_bar_sync(1, kPq2Warps * 32);
            if (threadIdx.x == 0) {
                mbar_arrive_expect_tx(lutbar, kPq2SliceFloats * 4);
                const char *src = reinterpret_cast<const char *>(p.perm + (size_t)s * kPq2SliceFloats);
                for (int i = 0; i < 4; i++)
                    bulk_g2s(reinterpret_cast<char *>(slut) + i * 32768, src + i * 32768, 32768, lutbar);
            }
            mbar_wait(lutbar, lut_phase);
            lut_phase ^= 1;
            for (uint32_t t = group; t < tiles_pb; t += kPq2Groups, n++) {
                const uint32_t slot = group + kPq2Groups * (n & 1), use = n >> 1;
                mbar_wait(&full[slot], use & 1);
                const uint32_t *tile = reinterpret_cast<const uint32_t *>(ring + (size_t)slot * kPq2TileBytes);
                const char *tile_row = reinterpret_cast<const char *>(tile) + (member * 16 + rsub) * 128;
                uint32_t cw[16];
#pragma unroll
                for (int j = 0; j < 16; j++)
                    cw[j] = *reinterpret_cast<const uint32_t *>(tile_row + ((lc0 ^ (4u * j)) & 0x7Cu));
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[slot]);  // the codes are in registers: the slot can be refilled
                const float total = fast ? pq2_rows16<true>(cw, lc0, lut_s) : pq2_rows16<false>(cw, lc0, lut_s);
                if (lane < 16) {
                    const uint32_t r = t * kPq2TileRows + member * 16 + rsub;
                    acc[r] = s == 0 ? total : __fadd_rn(acc[r], total);  // slice order: ((p0 + p1) + p2) + ...
                }
            }
        }
        // the block's scores are final
        named_bar_sync(1, kPq2Warps * 32);
        const uint64_t left = p.count - row0;
        const uint32_t n = left < p.block_rows ? (uint32_t)left : p.block_rows;
        if (p.scores)
            for (uint32_t i = threadIdx.x; i < n; i += kPq2Warps * 32) p.scores[row0 + i] = acc[i];
        if (do_topk) {
            // The LUT area is idle until the next slice load: candidate keys go there.  cand[0..128) = the running
            // list, then every key of this block above the current threshold (all of them for the first block).
            uint64_t *cand = reinterpret_cast<uint64_t *>(slut);
            // Rows [i_begin, i_end) are appended if they beat `thr`, then everything staged is reduced to the 128
            // best.  Without a threshold yet (the CTA's first block) the first 1920 rows go in unfiltered and give
            // one; the rest of the block is then filtered against it (a few percent pass) in a second round.
            uint32_t i_begin = 0;
            while (i_begin < n) {
                const uint64_t thr = top[p.k - 1];
                const uint32_t i_end = (thr == 0 && n - i_begin > 1920) ? i_begin + 1920 : n;
                named_bar_sync(1, kPq2Warps * 32);  // everyone has read top[] / thr
                if (threadIdx.x < 128) cand[threadIdx.x] = top[threadIdx.x];
                if (threadIdx.x == 0) *n_cand = 128;
                named_bar_sync(1, kPq2Warps * 32);
                for (uint32_t i0 = i_begin; i0 < i_end; i0 += kPq2Warps * 32) {
                    const uint32_t i = i0 + threadIdx.x;
                    uint64_t key = 0;
                    if (i < i_end) key = make_key(acc[i], p.id_base + (uint32_t)(row0 + i), p.flip);
                    const bool pass = i < i_end && key > thr;
                    const uint32_t mk = __ballot_sync(0xffffffffu, pass);
                    if (mk) {
                        uint32_t pos0 = 0;
                        if (lane == 0) pos0 = atomicAdd(n_cand, (uint32_t)__popc(mk));
                        pos0 = __shfl_sync(0xffffffffu, pos0, 0);
                        if (pass) cand[pos0 + __popc(mk & ((1u << lane) - 1))] = key;
                    }
                }
                named_bar_sync(1, kPq2Warps * 32);
                uint32_t nc = *n_cand;
                // windows of 2048 keys -> their 128 best at the window start -> packed -> again, until one is left
                for (;;) {
                    const uint32_t nw = (nc + 2047) / 2048;
                    for (uint32_t w = 0; w < nw; w++) {
                        const uint32_t nk = nc - w * 2048 < 2048 ? nc - w * 2048 : 2048;
                        coop_top128(cand + (size_t)w * 2048, nk, (nk + 127) / 128, warp, lane);
                    }
                    if (nw == 1) break;
                    for (uint32_t i = 128 + threadIdx.x; i < nw * 128; i += kPq2Warps * 32)
                        cand[i] = cand[(size_t)(i >> 7) * 2048 + (i & 127)];
                    named_bar_sync(1, kPq2Warps * 32);
                    nc = nw * 128;
                }
                if (threadIdx.x < 128) top[threadIdx.x] = cand[threadIdx.x];
                named_bar_sync(1, kPq2Warps * 32);
                i_begin = i_end;
            }
            // (the barrier at the top of the next slice / after the loop orders this against the next LUT load)
        }
    }
    if (do_topk) {
        named_bar_sync(1, kPq2Warps * 32);
        if (threadIdx.x < 128) p.partial[(size_t)blockIdx.x * 128 + threadIdx.x] = top[threadIdx.x];
        if (p.done) {
            __threadfence();
            named_bar_sync(1, kPq2Warps * 32);
            if (threadIdx.x == 0) {
                const uint32_t prev = atomicAdd(p.done, 1u);
                __threadfence();
                *n_cand = prev == gridDim.x - 1;
            }
            named_bar_sync(1, kPq2Warps * 32);
            if (*n_cand) {  // last CTA: all lists are published; the LUT area is the merge's work area
                prune_merge<1, true>(p.merge, reinterpret_cast<uint64_t *>(slut), threadIdx.x, kPq2Warps * 32);
                if (threadIdx.x == 0) *p.done = 0;
                if (p.px.world > 1) peer_exchange_merge(p.px, p.k, reinterpret_cast<uint64_t *>(slut));
            }
        }
    }
}

// Picks the geometry and launches.  Returns false when the shape is outside this kernel (row pitch not a
// multiple of 16 bytes: no tensor map) — the caller then uses the first-generation kernel.
// k == 0: scores only.  0 < k <= 128: the per-CTA lists go to ctx->d_partial ([*n_lists][128]) and d_scores is optional.
bool pq_scan2_launch(qz_ctx *ctx, const uint8_t *d_codes, uint64_t count, uint32_t n_chunks, uint32_t n_slices,
                     const float *d_perm, float *d_scores, uint32_t k, int largest, uint32_t id_base, uint32_t *n_lists,
                     const ScanFinal *fin, cudaError_t *err) {
    *err = cudaSuccess;
    if (n_chunks % 16 != 0 || count == 0 || count > 0x7FFFFFFFull) return false;
    if (ensure_smem_attr(ctx, (const void *)pq_scan2_kernel, 227 * 1024) != QZ_OK) {
        *err = cudaErrorInvalidValue;
        return true;
    }
    CUtensorMap map;
    if (!make_tensor_map_u8_2d(&map, d_codes, count, n_chunks, kPq2SliceChunks, kPq2TileRows, CU_TENSOR_MAP_SWIZZLE_NONE,
                               CU_TENSOR_MAP_L2_PROMOTION_NONE))
        return false;
    PqScan2Params p{};
    p.count = count;
    p.n_slices = n_slices;
    p.perm = d_perm;
    p.scores = d_scores;
    p.partial = k ? ctx->d_partial : nullptr;
    p.k = k ? k : 1;
    p.flip = largest ? 0u : 0xFFFFFFFFu;
    p.id_base = id_base;
    const uint32_t env_block = tune_u32("QZ_PQ_BLOCK_ROWS", 0);
    p.n_slots = 2 * kPq2Groups;  // two slots per consumer group (producer lanes in the kernel)
    const size_t fixed = (size_t)kPq2SliceFloats * 4 + (size_t)p.n_slots * kPq2TileBytes + (2 * p.n_slots + 1) * 8 + 128 * 8 + 16 + 1024;
    const uint32_t max_block = (uint32_t)((ctx->smem_optin - fixed) / 4 / kPq2TileRows * kPq2TileRows);
    // one block per CTA when the corpus is small enough for that, else blocks of the largest size that fits
    uint64_t per_cta = (count + ctx->sm_count - 1) / ctx->sm_count;
    per_cta = (per_cta + kPq2TileRows - 1) / kPq2TileRows * kPq2TileRows;
    p.block_rows = per_cta < max_block ? (uint32_t)per_cta : max_block;
    if (env_block) p.block_rows = env_block / kPq2TileRows * kPq2TileRows;
    if (p.block_rows == 0 || p.block_rows > max_block) p.block_rows = max_block;
    p.n_blocks = (uint32_t)((count + p.block_rows - 1) / p.block_rows);
    const uint32_t grid = p.n_blocks < (uint32_t)ctx->sm_count ? p.n_blocks : (uint32_t)ctx->sm_count;
    const size_t smem = fixed + (size_t)p.block_rows * 4;
    if (fin) *fin->fused = false;
    if (fin && k && grid <= kPruneMaxLists && !tune_flag("QZ_SCAN_NO_FUSED_MERGE")) {
        MergeParams &m = p.merge;
        m.in = ctx->d_partial;
        m.n_lists = grid;
        m.in_len = 128;
        m.in_stride = 128;
        m.Kp = 128;
        m.lists_per_cta = pow2_ceil(grid < 2 ? 2 : grid);
        m.out = nullptr;
        m.final_out = 1;
        m.k = k;
        m.flip = p.flip;
        m.keys = fin->d_keys;
        m.scores = fin->d_scores;
        m.ids = fin->d_ids;
        p.done = ctx->d_done;
        *fin->fused = true;
        if (fin->px) {
            p.px = *static_cast<const PeerExchange *>(fin->px);
            m.keys = p.px.gather[p.px.rank] + ((size_t)p.px.buf * p.px.world + p.px.rank) * kPeerStride;
            m.scores = nullptr;
            m.ids = nullptr;
        }
    }
    timing_begin(ctx);
    pq_scan2_kernel<<<grid, kPq2Threads, smem, ctx->stream>>>(map, p);
    timing_end(ctx);
    *err = cudaGetLastError();
    if (n_lists) *n_lists = grid;
    return true;
}

}  // namespace qz
