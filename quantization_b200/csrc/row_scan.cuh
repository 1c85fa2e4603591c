// row_scan.cuh — the single-query brute-force scan over fixed-size encoded rows.
//
// One persistent CTA per SM.  A producer thread streams tiles of `tile_rows` consecutive rows
// from HBM into a shared-memory ring with 1-D bulk async copies (TMA engine, UBLKCP) that
// complete on mbarriers; sixteen consumer warps in n_groups groups drain it (group g takes the CTA's local tiles
// g, g + n_groups, ...; local tile t lives in slot t mod n_slots and n_groups divides n_slots, so a slot always
// serves the same group).  Which GLOBAL tile a local tile is comes from a static prefix plus a device-wide tile
// counter (RowScanParams::tile_ctr), written by the producer next to the slot's mbarrier.  A
// warp works on 8 rows at a time: lane l owns 32-bit words l, l+32, ... of every row
// (bank-conflict free for any row stride, and the matching query words stay in registers),
// accumulates an exact integer per row, and a 9-shuffle butterfly leaves row r's total in a
// group of 4 lanes.  (v1 used 8 warps x 16-row batches and was consumer-latency bound at
// 5.7 TB/s: ~10 cycles between issues per warp; see DESIGN.md.)
// The epilogue applies the codec's affine map, optionally stores the score, and feeds the fused
// CTA top-k (topk.cuh).  Rows keep the reference's AoS layout byte for byte, including the
// 4-byte-misaligned u8 rows (stride = actual_dim + 4): tiles are 16-row multiples so every tile
// starts 16-byte aligned.
//
// Replaces, for the u8 codec: impl_score_dot_* / impl_score_l1_* (quantization/cpp/avx2.c:31-128,
// sse.c:23-47,108-149) and EncodedVectorsU8::score_point (encoded_vectors_u8.rs:331-384);
// for the binary codec: impl_xor_popcnt_* (sse.c:49-106) and calculate_metric
// (encoded_vectors_binary.rs:237-252); plus the caller's scan loop and heap
// (demos/src/ann_benchmark_data.rs:150-167).
#pragma once
#include "common.cuh"
#include "topk.cuh"
#include "peer_exchange.cuh"

namespace qz {

constexpr int kOpDot = 0, kOpL1 = 1, kOpXor = 2;
// kOpXorV / kOpXorV4: xor-popcount with 8 / 4 lanes per row reading 16-byte vectors (rows that are whole 16-byte
// units): 4 / 8 rows per warp step, two steps per batch, a 3- / 2-level reduction instead of the 5-level row
// butterfly.  For these mappings WPL counts 16-byte vectors per lane.  The host picks the lane count that keeps
// the LDS.128 quarter-warps bank-conflict free (4 lanes when the row stride is 64 mod 128 bytes).
constexpr int kOpXorV = 3, kOpXorV4 = 4;
constexpr int kConsumerWarps = 16;
constexpr int kScanThreads = (kConsumerWarps + 1) * 32;
constexpr int kMaxSlots = 16;
constexpr int kBatchRows = 8;    // rows one warp scores per step
constexpr int kTileRowAlign = 16;  // tiles are 16-row multiples => every tile starts 16-byte aligned

struct RowScanParams {
    const uint8_t *storage;  // 16-byte aligned, readable up to round_up(count*row_bytes, 16)
    uint32_t count;          // rows (point ids are u32, so count < 2^32)
    uint64_t total_bytes16;  // round_up(count*row_bytes, 16)
    uint32_t row_words;      // row_bytes / 4
    uint32_t hdr_words;      // 1 for u8 rows (f32 vector_offset), 0 for binary
    uint32_t code_words;     // words scored per row
    uint32_t tile_rows;      // multiple of 16
    uint32_t slot_bytes;     // >= tile_rows*row_bytes, multiple of 128
    uint32_t n_slots;        // power of two, <= 16
    uint32_t n_groups;       // consumer warp groups (power of two, divides n_slots): group g drains local tiles
                             // g, g + n_groups, ...; with n_slots = 2 * n_groups a group's next tile is in
                             // flight while it works on the current one
    uint32_t n_tiles;
    uint32_t warm_rows;      // rows per CTA scored before the first top-k threshold exists (<= 2048)
    const uint32_t *query;  // device blob: u8 [f32 offset][codes], binary [bits]
    float multiplier;       // u8
    uint32_t dim;           // binary
    int bin_zx;             // binary: 1 => zeros - xor, 0 => xor - zeros
    float *scores;          // optional [count]
    uint64_t *partial;      // optional [gridDim.x][Kp]
    uint32_t k, Kp, flip, id_base;
    // final merge inside this launch (done != NULL): the CTA that finishes last reduces all gridDim.x partial
    // lists to the final top-k with its consumer warps.  merge.in = partial, merge.n_lists = gridDim.x.
    uint32_t *done;  // zero before the launch; the last CTA leaves it zero again
    MergeParams merge;
    PeerExchange px;  // px.world > 1: merge.keys points at this GPU's own gather slot; see PeerExchange
    uint64_t *stamps;  // optional [gridDim.x][QZ_STAMP_WORDS] phase stamps (qz_ctx_set_phase_stamps)
    // 1 = the query is written by the kernel launched just before this one on the stream (the host-facing search: query
    // encode -> scan) and this launch is its programmatic dependent: the CTAs set up and the ring starts filling
    // while the encode kernel runs; the consumers wait for it before they read the query
    uint32_t query_dep;
    // Tile schedule.  Local tile t of a CTA is global tile blockIdx.x + t * gridDim.x while t < n_static (one ring's
    // worth: they are issued while the first draw is in flight); later
    // tiles are drawn from a device-wide counter (global tile gridDim.x * n_static + ticket), a few at a time,
    // so a CTA that starts late (the SM was still merging the previous scan of a chain) or runs on a slower SM simply
    // takes fewer of them.  tile_ctr == NULL: every tile is static (n_static is ignored).
    uint32_t n_static;
    uint32_t *tile_ctr;  // zero before the launch; the last CTA leaves it zero again
};
constexpr uint32_t kTileChunk = 32;  // largest draw from the tile counter
constexpr uint32_t kNoTile = 0xFFFFFFFFu;

// tiles of a CTA that go through the top-k warm-up (0: no warm-up); shared by the kernel and the host's planning
__host__ __device__ __forceinline__ uint32_t scan_warm_rounds(uint32_t warm_rows, uint32_t n_groups, uint32_t tile_rows) {
    const uint32_t cap = (uint32_t)(kConsumerWarps * kCandPerWarp), per_round = n_groups * tile_rows;
    if (per_round > cap) return 0;
    uint32_t r = (warm_rows + per_round - 1) / per_round;
    if (r * per_round > cap) r = cap / per_round;
    return r;
}

__device__ __forceinline__ uint64_t global_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// thread `who` of the CTA records phase i (no-op without a stamp buffer)
__device__ __forceinline__ void stamp(const RowScanParams &p, int i, bool who, bool wall = false) {
    if (p.stamps && who) p.stamps[(size_t)blockIdx.x * QZ_STAMP_WORDS + i] = wall ? global_ns() : (uint64_t)clock64();
}

template <int OP>
__device__ __forceinline__ uint32_t scan_op(uint32_t v, uint32_t q, uint32_t acc) {
    if (OP == kOpDot) return __dp4a(v, q, acc);
    if (OP == kOpL1) return acc + __vsadu4(v, q);
    return acc + __popc(v ^ q);
}

// WPL = 32-bit words per lane per row (ceil(code_words / 32)); FULL = code_words == 32 * WPL, so no
// lane ever needs masking.
template <int OP, int WPL, bool FULL>
__global__ void __launch_bounds__(kScanThreads, 1) row_scan_kernel(const RowScanParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *ring = smem;
    uint64_t *top = reinterpret_cast<uint64_t *>(smem + (size_t)p.n_slots * p.slot_bytes);
    uint64_t *cand_all = top + p.Kp;
    uint64_t *full = cand_all + kConsumerWarps * kCandPerWarp;
    uint64_t *empty = full + kMaxSlots;
    uint64_t *thr = empty + kMaxSlots;
    int *lock = reinterpret_cast<int *>(thr + 1);
    uint32_t *slot_tile = reinterpret_cast<uint32_t *>(thr + 2);  // [kMaxSlots] global tile held by each slot

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool do_topk = p.partial != nullptr;
    stamp(p, 0, threadIdx.x == 0, true);
    stamp(p, 1, threadIdx.x == 0);
    if (p.stamps && threadIdx.x < 2) p.stamps[(size_t)blockIdx.x * QZ_STAMP_WORDS + 5 + threadIdx.x] = 0;
    const uint32_t n_slots = p.n_slots, n_groups = p.n_groups;
    const uint32_t warps_per_slot = kConsumerWarps / n_groups;  // warps that share a tile

    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < n_slots; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], warps_per_slot);
        }
        *thr = 0;
        *lock = 0;
        fence_mbar_init();
    }
    if (do_topk)
        for (uint32_t i = threadIdx.x; i < p.Kp; i += blockDim.x) top[i] = 0;
    // with a tile counter a CTA may draw fewer tiles than the warm-up covers (it started late and the pool was empty):
    // the warm-up's staging area starts as empty keys, and the reduction always takes all of it
    if (do_topk && p.tile_ctr)
        for (uint32_t i = threadIdx.x; i < (uint32_t)(kConsumerWarps * kCandPerWarp); i += blockDim.x) cand_all[i] = 0ull;
    __syncthreads();
    // A scan launched behind this one as a programmatic dependent (row_scan.cu, chained launches) may take the SMs
    // this grid's CTAs leave while the last CTA is still merging; it waits for this grid to complete before it
    // touches anything shared with it (griddepcontrol.wait below).
    if (p.done) asm volatile("griddepcontrol.launch_dependents;");

    // tiles of this CTA: blockIdx.x, blockIdx.x + gridDim.x, ...; local tile t lives in slot t % n_slots
    const uint32_t my_tiles = (p.n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
    const uint32_t tile_bytes = p.tile_rows * p.row_words * 4;

    if (warp == kConsumerWarps) {
        // ---------------- producer: one thread feeds the ring with 1-D bulk copies ----------------
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            uint32_t slot = 0, round = 0;
            // global tile gt (or kNoTile: this CTA's stream ends here for the group that owns the slot) -> next slot
            auto issue = [&](uint32_t gt) {
                if (round > 0) mbar_wait(&empty[slot], (round - 1) & 1);
                slot_tile[slot] = gt;
                if (gt != kNoTile) {
                    const uint64_t off = (uint64_t)gt * tile_bytes;
                    uint64_t bytes = p.total_bytes16 - off;
                    if (bytes > tile_bytes) bytes = tile_bytes;
                    mbar_arrive_expect_tx(&full[slot], (uint32_t)bytes);
                    bulk_g2s_hint(ring + (size_t)slot * p.slot_bytes, p.storage + off, (uint32_t)bytes, &full[slot], pol);
                } else {
                    mbar_arrive(&full[slot]);
                }
                if (++slot == n_slots) {
                    slot = 0;
                    round++;
                }
            };
            if (p.tile_ctr) {
                // guided self-scheduling: a draw takes 1 / (2 * CTAs) of what is left (at most kTileChunk tiles, at
                // least one), so the counter sees a few dozen atomics per CTA and the last draws are single tiles;
                // the next draw is always in flight while the current one is issued
                const uint32_t base = gridDim.x * p.n_static, n_dyn = p.n_tiles - base;
                auto want = [&](uint32_t drawn) {
                    const uint32_t sz = (drawn < n_dyn ? n_dyn - drawn : 0) / (2 * gridDim.x);
                    return sz < 1 ? 1u : (sz > kTileChunk ? kTileChunk : sz);
                };
                uint32_t sz_next = want(0);
                uint32_t c_next = atomicAdd(p.tile_ctr, sz_next);  // in flight during the static part
                for (uint32_t t = 0; t < p.n_static; t++) issue(blockIdx.x + t * gridDim.x);
                for (;;) {
                    const uint32_t c = c_next, sz = sz_next;
                    if (c >= n_dyn) break;
                    sz_next = want(c + sz);
                    c_next = atomicAdd(p.tile_ctr, sz_next);
                    for (uint32_t j = 0; j < sz && c + j < n_dyn; j++) issue(base + c + j);
                }
            } else {
                for (uint32_t gt = blockIdx.x; gt < p.n_tiles; gt += gridDim.x) issue(gt);
            }
            for (uint32_t g = 0; g < n_groups; g++) issue(kNoTile);  // every group finds one at its next local tile
        }
        return;
    }

    // ---------------- consumers: warps_per_slot warps share each slot ----------------
    constexpr bool kVec = OP == kOpXorV || OP == kOpXorV4;
    constexpr int kLpr = OP == kOpXorV4 ? 4 : 8;                // lanes per row (vector mappings)
    constexpr int kRps = 32 / kLpr;                             // rows per warp step
    constexpr int kRowsB = kVec ? 2 * kRps : kBatchRows;        // rows per batch
    const uint32_t lsub = lane & (kLpr - 1), lgrp = lane / kLpr;
    constexpr int kQ = kVec ? 4 * WPL : WPL;  // query words a lane keeps in registers
    uint32_t q[kQ];
    bool jv[WPL];
    if (p.query_dep) asm volatile("griddepcontrol.wait;" ::: "memory");
#pragma unroll
    for (int j = 0; j < WPL; j++) {
        if (kVec) {
            const uint32_t v = lsub + kLpr * j;  // 16-byte vector of the row this lane owns
            jv[j] = FULL || 4 * v < p.code_words;
#pragma unroll
            for (int c = 0; c < 4; c++) q[(4 * j + c) % kQ] = jv[j] ? __ldcg(p.query + 4 * v + c) : 0u;
        } else {
            const uint32_t w = lane + 32 * j;
            jv[j] = FULL || w < p.code_words;
            q[j] = jv[j] ? __ldcg(p.query + p.hdr_words + w) : 0u;
        }
    }
    const float q_off = p.hdr_words ? __uint_as_float(__ldcg(p.query)) : 0.0f;

    CtaTopK st{top, thr, lock, p.k, p.Kp};
    uint64_t *cand = cand_all + warp * kCandPerWarp;
    int cnt = 0;
    const uint32_t rw = p.row_words;
    const bool upper16 = lane & 16, upper8 = lane & 8, upper4 = lane & 4;
    // row (of the 8-row batch) this lane reports, and whether it reports at all (8 reporters per warp)
    const uint32_t rr = kVec ? lgrp + (lsub ? kRps : 0)
                             : ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
    const bool reporter = kVec ? lsub < 2 : !(lane & 3);
    const uint32_t group = warp / warps_per_slot, member = warp - group * warps_per_slot;
    const uint32_t slot_mask = n_slots - 1, slot_shift = __ffs(n_slots) - 1;
    const uint32_t batch_stride = warps_per_slot * kRowsB;

    // Top-k warm-up (k <= 128): the keys of the CTA's first `warm_tiles` tiles go unfiltered into the
    // staging area (one slot per row, no atomics); then the consumer warps together reduce them to
    // the CTA's first sorted list and threshold.  Without this every warp starts with threshold 0,
    // fills its 128-entry buffer within 16 steps and the 16 lock-serialised flushes that follow
    // stall the whole ring (measured: +80 us on a 218 us scan).
    const bool coop = do_topk && p.Kp == (uint32_t)kCandPerWarp;
    const uint32_t warm_rounds = coop ? scan_warm_rounds(p.warm_rows, n_groups, p.tile_rows) : 0;

    uint32_t round = 0;
    bool warm = warm_rounds > 0, ended = false;
    for (uint32_t t = group;; t += n_groups, round++) {
        if (warm && round == warm_rounds) {
            // ---- end of warm-up: cooperative reduction of the staged keys (all consumer warps) ----
            // (with a tile counter: all of the zero-initialised staging area, whatever this CTA got to draw)
            uint32_t tiles_w = warm_rounds * n_groups;
            if (!p.tile_ctr && tiles_w > my_tiles) tiles_w = my_tiles;
            const uint32_t n_keys = tiles_w * p.tile_rows;
            const uint32_t n_chunks = (n_keys + kCandPerWarp - 1) / kCandPerWarp;
            named_bar_sync(1, kConsumerWarps * 32);
            coop_top128(cand_all, n_keys, n_chunks, warp, lane);
            if (warp == 0) {
                for (int i = lane; i < kCandPerWarp; i += 32) top[i] = cand_all[i];
                __syncwarp();
                if (lane == 0) *thr = top[p.k - 1];
            }
            named_bar_sync(1, kConsumerWarps * 32);
            stamp(p, 2, threadIdx.x == 0);
            warm = false;
        }
        if (ended) {
            if (warm) continue;  // (fewer tiles than warm-up rounds: keep counting rounds up to the barrier)
            break;
        }
        const uint32_t slot = t & slot_mask;
        const uint32_t *sl = reinterpret_cast<const uint32_t *>(ring + (size_t)slot * p.slot_bytes);
        mbar_wait(&full[slot], (t >> slot_shift) & 1);
        const uint32_t gt = slot_tile[slot];
        if (gt == kNoTile) {
            ended = true;
            if (warm) continue;
            break;
        }
        const uint32_t row0 = gt * p.tile_rows;
        const uint32_t left = p.count - row0;
        const uint32_t rows_valid = left < p.tile_rows ? left : p.tile_rows;
        const uint32_t rows_iter = warm ? p.tile_rows : rows_valid;  // warm-up also writes the empty keys of missing rows

        for (uint32_t b = member * kRowsB; b < rows_iter; b += batch_stride) {
            uint32_t total;
            if (kVec) {
                // rows b + lgrp and b + kRps + lgrp; lsub picks the 16-byte vectors
                const uint4 *r0 = reinterpret_cast<const uint4 *>(sl + (b + lgrp) * rw) + lsub;
                const uint4 *r1 = reinterpret_cast<const uint4 *>(sl + (b + kRps + lgrp) * rw) + lsub;
                uint32_t a0 = 0, a1 = 0;
#pragma unroll
                for (int j = 0; j < WPL; j++) {
                    if (FULL || jv[j]) {
                        const uint4 v0 = r0[kLpr * j], v1 = r1[kLpr * j];
                        const int qb = (4 * j) % kQ;
                        a0 += __popc(v0.x ^ q[qb]) + __popc(v0.y ^ q[qb + 1]) + __popc(v0.z ^ q[qb + 2]) +
                              __popc(v0.w ^ q[qb + 3]);
                        a1 += __popc(v1.x ^ q[qb]) + __popc(v1.y ^ q[qb + 1]) + __popc(v1.z ^ q[qb + 2]) +
                              __popc(v1.w ^ q[qb + 3]);
                    }
                }
#pragma unroll
                for (int o = kLpr / 2; o > 0; o >>= 1) {
                    a0 += __shfl_xor_sync(0xffffffffu, a0, o);
                    a1 += __shfl_xor_sync(0xffffffffu, a1, o);
                }
                total = lsub ? a1 : a0;
            } else {
            uint32_t acc[kBatchRows];
#pragma unroll
            for (int r = 0; r < kBatchRows; r++) acc[r] = 0;
            const uint32_t *base = sl + b * rw + p.hdr_words + lane;
#pragma unroll
            for (int j = 0; j < WPL; j++) {
                if (FULL || jv[j]) {
#pragma unroll
                    for (int r = 0; r < kBatchRows; r++) acc[r] = scan_op<OP>(base[r * rw + 32 * j], q[j], acc[r]);
                }
            }
            // butterfly: 8 rows x 32 lanes -> the 4 lanes with (lane >> 2) pattern rr hold row rr
#pragma unroll
            for (int i = 0; i < 4; i++) {
                uint32_t keep = upper16 ? acc[i + 4] : acc[i], send = upper16 ? acc[i] : acc[i + 4];
                acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
            }
#pragma unroll
            for (int i = 0; i < 2; i++) {
                uint32_t keep = upper8 ? acc[i + 2] : acc[i], send = upper8 ? acc[i] : acc[i + 2];
                acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
            }
            {
                uint32_t keep = upper4 ? acc[1] : acc[0], send = upper4 ? acc[0] : acc[1];
                acc[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            acc[0] += __shfl_xor_sync(0xffffffffu, acc[0], 2);
            total = acc[0] + __shfl_xor_sync(0xffffffffu, acc[0], 1);
            }

            // epilogue: 8 reporting lanes per warp, one row each
            const uint32_t rit = b + rr;
            const bool valid = reporter && rit < rows_valid;
            float score;
            if (OP == kOpXor || kVec) {
                // calculate_metric, encoded_vectors_binary.rs:237-252: zeros = dim - xor, then zeros - xor or
                // xor - zeros in f32.  Every value is an integer below 2^24, so the f32 steps are exact and equal
                // to converting the integer difference once.
                const int d2 = (int)p.dim - 2 * (int)total;
                score = __int2float_rn(p.bin_zx ? d2 : -d2);
            } else {
                // multiplier * S + query.offset + vector_offset, left to right, no FMA (u8.rs:347)
                const float vo = __uint_as_float(sl[rit * rw]);
                score = __fadd_rn(__fadd_rn(__fmul_rn(p.multiplier, __uint2float_rn(total)), q_off), vo);
            }
            const uint32_t grow = row0 + rit;
            if (p.scores && valid) p.scores[grow] = score;
            if (do_topk) {
                const uint64_t key = make_key(score, p.id_base + grow, p.flip);
                if (warm) {
                    if (reporter) cand_all[t * p.tile_rows + rit] = valid ? key : 0ull;
                } else {
                    const bool pass = valid && key > *thr;
                    const uint32_t m = __ballot_sync(0xffffffffu, pass);
                    if (m) {
                        const int n = __popc(m);
                        if (cnt + n > kCandPerWarp) {
                            // staging full.  The threshold has usually risen since these keys were staged (another
                            // warp flushed): drop what no longer passes, and only take the CTA lock if that did
                            // not free half of the buffer
                            const uint64_t th = *thr;
                            int kept = 0;
                            for (int base = 0; base < cnt; base += 32) {
                                const uint64_t c = base + (int)lane < cnt ? cand[base + lane] : 0ull;
                                const bool keep = c > th;
                                const uint32_t mk = __ballot_sync(0xffffffffu, keep);
                                if (keep) cand[kept + __popc(mk & ((1u << lane) - 1))] = c;
                                kept += __popc(mk);
                                __syncwarp();
                            }
                            cnt = kept;
                            if (cnt + n > kCandPerWarp / 2) {
                                cta_topk_flush(st, cand, cnt, lane);
                                cnt = 0;
                            }
                        }
                        if (pass) cand[cnt + __popc(m & ((1u << lane) - 1))] = key;
                        cnt += n;
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
    }

    if (do_topk) {
        __syncwarp();
        stamp(p, 3, threadIdx.x == 0);
        if (coop) {
            // Leftovers.  Much of what is still staged no longer beats the CTA's k-th best: every warp first drops
            // those keys, then all warps pack the rest contiguously.  One chunk (<= 128 keys): one warp sorts and
            // merges it into the list.  More: only the occupied chunks go through the cooperative reduction (the first
            // version always reduced 16 padded chunks: 4 merge levels, 8 us).
            const uint64_t th = *thr;
            int kept = 0;
            for (int base = 0; base < cnt; base += 32) {
                const uint64_t c = base + (int)lane < cnt ? cand[base + lane] : 0ull;
                const bool keep = c > th;
                const uint32_t mk = __ballot_sync(0xffffffffu, keep);
                __syncwarp();
                if (keep) cand[kept + __popc(mk & ((1u << lane) - 1))] = c;
                kept += __popc(mk);
                __syncwarp();
            }
            cnt = kept;
            uint32_t *n_left = reinterpret_cast<uint32_t *>(lock);  // the flush lock is idle from here on
            named_bar_sync(1, kConsumerWarps * 32);                 // (all flushes done, `lock` is 0)
            uint32_t pos0 = 0;
            if (lane == 0 && cnt) pos0 = atomicAdd(n_left, (uint32_t)cnt);
            pos0 = __shfl_sync(0xffffffffu, pos0, 0);
            named_bar_sync(1, kConsumerWarps * 32);
            const uint32_t total = *n_left;
            {
                // the packed keys start in warp 0's staging area, which may still hold leftovers of its owner (and so
                // on): everybody reads its keys into registers, then (barrier) writes them to their place
                uint64_t mine[4];
#pragma unroll
                for (int r = 0; r < 4; r++) mine[r] = (int)(lane + 32 * r) < cnt ? cand[lane + 32 * r] : 0ull;
                named_bar_sync(1, kConsumerWarps * 32);  // (and everyone has read `total`)
                if (threadIdx.x == 0) *n_left = 0;       // it is the flush lock again
#pragma unroll
                for (int r = 0; r < 4; r++)
                    if ((int)(lane + 32 * r) < cnt) cand_all[pos0 + lane + 32 * r] = mine[r];
                named_bar_sync(1, kConsumerWarps * 32);
                if (total <= (uint32_t)kCandPerWarp) {
                    if (warp == 0 && total) cta_topk_flush(st, cand_all, (int)total, lane);
                } else {
                    // a few chunks: sorted in parallel, max-merge tree of log2(chunks) levels, then into the list
                    coop_top128(cand_all, total, (total + kCandPerWarp - 1) / kCandPerWarp, warp, lane);
                    if (warp == 0) {
                        uint64_t t4[4];
#pragma unroll
                        for (int r = 0; r < 4; r++) {
                            const uint64_t a = top[lane * 4 + r], b = cand_all[kCandPerWarp - 1 - (lane * 4 + r)];
                            t4[r] = a > b ? a : b;
                        }
                        reg128_merge_desc(t4, lane);
#pragma unroll
                        for (int r = 0; r < 4; r++) top[lane * 4 + r] = t4[r];
                    }
                }
            }
        } else if (cnt > 0) {
            cta_topk_flush(st, cand, cnt, lane);
        }
        named_bar_sync(1, kConsumerWarps * 32);
        // the per-CTA lists, the arrival counter, the output buffers and the peer slots are shared with the previous
        // scan of a chain: it must be complete (no-op for an ordinary launch)
        asm volatile("griddepcontrol.wait;" ::: "memory");
        uint64_t *out = p.partial + (size_t)blockIdx.x * p.Kp;
        for (uint32_t i = threadIdx.x; i < p.Kp; i += kConsumerWarps * 32) out[i] = top[i];
        stamp(p, 4, threadIdx.x == 0);
        if (p.done) {
            // last-CTA-done: publish this list, count, and let the last arrival merge them all.  Every bulk copy
            // of this CTA has been consumed, so the ring is free to serve as the merge's work area.
            __threadfence();
            named_bar_sync(1, kConsumerWarps * 32);
            if (threadIdx.x == 0) {
                const uint32_t prev = atomicAdd(p.done, 1u);
                __threadfence();
                *lock = prev == gridDim.x - 1;
            }
            named_bar_sync(1, kConsumerWarps * 32);
            // (the merge's work area starts at shared offset 0 and may overlay `lock` when the ring is small:
            // everyone reads the flag into a register before anyone enters the merge)
            const int is_last = *lock;
            named_bar_sync(1, kConsumerWarps * 32);
            if (is_last) {
                prune_merge<1, true>(p.merge, reinterpret_cast<uint64_t *>(ring), threadIdx.x, kConsumerWarps * 32);
                if (threadIdx.x == 0) {
                    *p.done = 0;
                    if (p.tile_ctr) *p.tile_ctr = 0;
                }
                stamp(p, 5, threadIdx.x == 0);
                if (p.px.world > 1) peer_exchange_merge(p.px, p.k, reinterpret_cast<uint64_t *>(ring));
                stamp(p, 6, threadIdx.x == 0);
            }
        }
    }
    stamp(p, 7, threadIdx.x == 0, true);
}

}  // namespace qz
