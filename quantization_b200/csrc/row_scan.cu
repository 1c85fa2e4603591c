// row_scan.cu — host-side planning + launch of row_scan_kernel and merge_lists_kernel
#include <stdlib.h>

#include "qz_internal.h"
#include "row_scan.cuh"

namespace qz {

struct ScanPlan {
    uint32_t tile_rows, slot_bytes, n_slots, n_groups, wpl, lpr;
    bool vec;  // binary rows scored with the 8-lanes-per-row vector mapping (kOpXorV)
    uint32_t n_tiles;
    bool full;
    size_t smem;
};

cudaError_t launch_row_scan_dot(qz_ctx *, uint32_t, bool, const RowScanParams &, uint32_t, size_t);
cudaError_t launch_row_scan_l1(qz_ctx *, uint32_t, bool, const RowScanParams &, uint32_t, size_t);
cudaError_t launch_row_scan_xor(qz_ctx *, uint32_t, bool, const RowScanParams &, uint32_t, size_t);
cudaError_t launch_row_scan_xorv(qz_ctx *, uint32_t, uint32_t, bool, const RowScanParams &, uint32_t, size_t);

static const int kWplSet[] = {1, 2, 3, 4, 6, 8, 12, 16, 24, 32};

// Shapes the TMA ring kernel handles: rows that are whole 32-bit words, at most 32*32 code words.
// Ring geometry: slots of ~24 KB (16 rows of the 1536-d u8 layout); the slot count is a power of two
// (16 / n_slots consumer warps are bound to each slot).  hbm_probe (profiles/r1_hbm_probe.txt): 4-8
// slots of 24 KB per SM sustain 7.0-7.3 TB/s of bulk copies; many small slots do not.
// `small_rows` (binary codec): the scan is instruction-heavy per byte, so a warp group must not wait for its own
// refill: 16 slots of ~12 KB, 8 groups, each group alternating between two slots.  (u8 rows: 8 slots of ~24 KB,
// one per group, measured at the HBM ceiling.)
static bool plan_row_scan(const qz_ctx *ctx, uint64_t count, uint32_t row_bytes, uint32_t code_words, uint32_t Kp,
                          bool small_rows, ScanPlan *pl) {
    if (row_bytes == 0 || row_bytes % 4 != 0 || code_words == 0 || code_words > 32 * 32) return false;
    if (count >= 0xFFFFFFFFull) return false;
    const uint32_t env_target = tune_u32("QZ_SCAN_TILE_BYTES", 0), env_align = tune_u32("QZ_SCAN_ROW_ALIGN", 0);
    const uint32_t env_slots = tune_u32("QZ_SCAN_SLOTS", 0), env_groups = tune_u32("QZ_SCAN_GROUPS", 0);
    const uint32_t target = env_target ? env_target : (small_rows ? 12416 : 24832);
    const uint32_t align = env_align ? env_align : kTileRowAlign;
    uint32_t tile_rows = (target / row_bytes) / align * align;
    if (tile_rows < align) tile_rows = align;
    // small corpora: shrink tiles until every SM has ~2 of them
    const uint64_t want_tiles = 2ull * ctx->sm_count;
    while (tile_rows > align && (count + tile_rows - 1) / tile_rows < want_tiles) tile_rows -= align;
    uint32_t slot_bytes = (tile_rows * row_bytes + 127) / 128 * 128;
    const size_t fixed = (size_t)Kp * 8 + (size_t)kConsumerWarps * kCandPerWarp * 8 + 2 * kMaxSlots * 8 + 16 + kMaxSlots * 4;
    if (ctx->smem_optin < fixed + (size_t)slot_bytes) return false;
    uint32_t fit = (uint32_t)((ctx->smem_optin - fixed) / slot_bytes);
    uint32_t n_slots = 1;
    const uint32_t max_slots = env_slots ? env_slots : (small_rows ? 16 : 8);
    while (n_slots * 2 <= fit && n_slots * 2 <= max_slots && n_slots * 2 <= (uint32_t)kMaxSlots) n_slots *= 2;
    uint32_t n_groups = env_groups ? pow2_ceil(env_groups) : 8;
    if (n_groups > (uint32_t)kConsumerWarps) n_groups = kConsumerWarps;
    if (n_groups > n_slots) n_groups = n_slots;
    pl->vec = small_rows && row_bytes % 16 == 0 && row_bytes <= 512 && !tune_flag("QZ_SCAN_NO_VEC");
    pl->lpr = 8;
    uint32_t need = (code_words + 31) / 32;
    uint32_t wpl = 0;
    for (int w : kWplSet)
        if ((uint32_t)w >= need) {
            wpl = w;
            break;
        }
    if (!wpl) return false;
    pl->full = code_words == 32 * wpl;
    pl->lpr = 8;
    if (pl->vec) {
        const uint32_t row_vec = row_bytes / 16;
        // 4 lanes per row when two consecutive rows then fall into disjoint bank halves (stride = 64 mod 128 bytes)
        const uint32_t env_lpr = tune_u32("QZ_SCAN_LPR", 0);
        pl->lpr = env_lpr ? env_lpr : ((row_vec % 8 == 4 && row_vec <= 16) ? 4 : 8);
        wpl = (row_vec + pl->lpr - 1) / pl->lpr;  // 16-byte vectors per lane
        pl->full = row_vec == pl->lpr * wpl;
        if (wpl > 4) pl->vec = false;
    }
    if (!pl->vec) {
        wpl = 0;
        for (int w : kWplSet)
            if ((uint32_t)w >= need) {
                wpl = w;
                break;
            }
        pl->full = code_words == 32 * wpl;
    }
    pl->tile_rows = tile_rows;
    pl->slot_bytes = slot_bytes;
    pl->n_slots = n_slots;
    pl->n_groups = n_groups;
    pl->wpl = wpl;
    pl->n_tiles = (uint32_t)((count + tile_rows - 1) / tile_rows);
    pl->smem = fixed + (size_t)n_slots * slot_bytes;
    return true;
}

// Launch the scan of a u8 or binary index.  d_scores (optional) receives every score;
// if k > 0 the per-CTA top-k lists land in ctx->d_partial ([*n_lists][*Kp_out]).
// Returns QZ_ARGUMENTS with a message if the row shape is not supported by this kernel.
// With `fin` (k > 0) the kernel also merges the lists and writes the final results, when the merge's work
// area fits the ring; *fin->fused tells the caller whether it still has to call launch_merge.
qz_status launch_row_scan(qz_index *ix, const void *d_query, float *d_scores, uint32_t k, int largest,
                          uint32_t *n_lists, uint32_t *Kp_out, const ScanFinal *fin) {
    qz_ctx *ctx = ix->ctx;
    const uint32_t Kp = k ? topk_pad(k) : 128;
    const uint32_t hdr = ix->codec == QZ_CODEC_U8 ? 1 : 0;
    const uint32_t code_words = ix->row_bytes / 4 - hdr;
    ScanPlan pl;
    if (!plan_row_scan(ctx, ix->count, ix->row_bytes, code_words, Kp, ix->codec == QZ_CODEC_BIN, &pl))
        return fail(ctx, QZ_ARGUMENTS, "row shape not supported by the streaming scan kernel");
    RowScanParams p{};
    p.storage = ix->d_storage;
    p.count = (uint32_t)ix->count;
    p.total_bytes16 = (ix->count * (uint64_t)ix->row_bytes + 15) / 16 * 16;
    p.row_words = ix->row_bytes / 4;
    p.hdr_words = hdr;
    p.code_words = code_words;
    p.tile_rows = pl.tile_rows;
    p.slot_bytes = pl.slot_bytes;
    p.n_slots = pl.n_slots;
    p.n_groups = pl.n_groups;
    // u8 CTAs see a few thousand rows: 512 unfiltered rows buy a threshold good enough that no warp ever fills its
    // staging buffer.  Binary CTAs see ~70k rows: a longer warm-up keeps the buffers from filling all at once.
    const uint32_t env_warm = tune_u32("QZ_SCAN_WARM", 0);
    p.warm_rows = env_warm ? env_warm : (ix->codec == QZ_CODEC_BIN ? 2048 : 512);
    p.n_tiles = pl.n_tiles;
    p.query = reinterpret_cast<const uint32_t *>(d_query);
    p.multiplier = ix->multiplier;
    p.dim = ix->dim;
    const bool dot = ix->distance == QZ_DOT;
    p.bin_zx = (dot && !ix->invert) || (!dot && ix->invert);
    p.scores = d_scores;
    p.partial = k ? ctx->d_partial : nullptr;
    p.k = k ? k : 1;
    p.Kp = Kp;
    p.flip = largest ? 0u : 0xFFFFFFFFu;
    p.id_base = ix->id_base;
    p.stamps = ctx->d_stamps;
    uint32_t grid = pl.n_tiles < (uint32_t)ctx->sm_count ? pl.n_tiles : (uint32_t)ctx->sm_count;
    if (fin) *fin->fused = false;
    const bool env_nofuse = tune_flag("QZ_SCAN_NO_FUSED_MERGE");
    if (fin && k && !env_nofuse && grid <= kPruneMaxLists) {
        const uint32_t lpc = pow2_ceil(grid < 2 ? 2 : grid);
        size_t need = (size_t)(lpc / 2) * Kp * 8;
        if (need < kPruneSmem) need = kPruneSmem;
        // the merge works in the kernel's shared memory from offset 0 (ring, then the by then dead top-k areas)
        if (need <= ctx->smem_optin) {
            if (pl.smem < need) pl.smem = need;
            MergeParams &m = p.merge;
            m.in = ctx->d_partial;
            m.n_lists = grid;
            m.in_len = Kp;
            m.in_stride = Kp;
            m.Kp = Kp;
            m.lists_per_cta = lpc;
            m.out = nullptr;
            m.final_out = 1;
            m.k = k;
            m.flip = p.flip;
            m.keys = fin->d_keys;
            m.scores = fin->d_scores;
            m.ids = fin->d_ids;
            p.done = ctx->d_done;
            *fin->fused = true;
            // tile counter (row_scan.cuh): one CTA per SM; the static part is one ring of tiles, everything else is drawn,
            // so a CTA that starts late (its SM was still merging and exchanging the previous scan of a chain: 10-25 us)
            // does little more than that ring.  Two counters alternate: a chained scan starts while its predecessor's
            // last CTA is still merging, but never before the scan before that one has completed (with one CTA per SM
            // the predecessor cannot be fully resident, hence cannot let ITS dependents go, until then).
            const uint32_t share = pl.n_tiles / grid;
            const uint32_t n_static = p.n_slots;
            if (grid == (uint32_t)ctx->sm_count && share >= 2 * n_static && !tune_flag("QZ_SCAN_STATIC_TILES")) {
                p.n_static = n_static;
                p.tile_ctr = ctx->d_done + 16 + 16 * (ctx->scan_seq++ & 1);
            }
            if (fin->px) {
                p.px = *static_cast<const PeerExchange *>(fin->px);
                m.keys = p.px.gather[p.px.rank] + ((size_t)p.px.buf * p.px.world + p.px.rank) * kPeerStride;
                m.scores = nullptr;
                m.ids = nullptr;
            }
        }
    }
    if (fin && fin->px && !*fin->fused)
        return fail(ctx, QZ_ARGUMENTS, "peer exchange needs the in-launch merge (k <= 128, streaming row shape)");
    // chained launch: see qz_ctx::chain_tail.  The kernel reads nothing a previous scan writes before its
    // griddepcontrol.wait (query and storage are inputs of both), so only the tail of the previous scan overlaps.
    const bool chainable = fin && *fin->fused && !ctx->timing && (ctx->own_stream || ctx->chain_ok);
    ctx->pdl_next = chainable && ctx->chain_tail != 0 && ctx->chain_tail == ctx->launches;
    // host-facing search: the query-encode kernel is the previous launch on the stream and triggers its dependents at
    // once, so this kernel's set-up and first bulk copies overlap it (nothing but events may sit between the two)
    if (ctx->query_dep_next && !ctx->timing && !ctx->d_stamps) {
        p.query_dep = 1;
        ctx->pdl_next = true;
    }
    ctx->query_dep_next = false;
    cudaError_t e;
    timing_begin(ctx);
    if (ix->codec == QZ_CODEC_BIN && pl.vec) e = launch_row_scan_xorv(ctx, pl.lpr, pl.wpl, pl.full, p, grid, pl.smem);
    else if (ix->codec == QZ_CODEC_BIN) e = launch_row_scan_xor(ctx, pl.wpl, pl.full, p, grid, pl.smem);
    else if (ix->distance == QZ_L1) e = launch_row_scan_l1(ctx, pl.wpl, pl.full, p, grid, pl.smem);
    else e = launch_row_scan_dot(ctx, pl.wpl, pl.full, p, grid, pl.smem);
    timing_end(ctx);
    ctx->pdl_next = false;
    ctx->chain_tail = 0;
    if (e == cudaErrorInvalidValue && !ctx->err.empty() && ctx->err.find("cudaFuncSetAttribute") != std::string::npos)
        return QZ_ENCODING;  // ensure_smem_attr failed and left its message
    if (e != cudaSuccess) return cuda_fail(ctx, e, "row_scan_kernel launch");
    ctx->launches++;
    if (chainable) ctx->chain_tail = ctx->launches;
    if (n_lists) *n_lists = grid;
    if (Kp_out) *Kp_out = Kp;
    return QZ_OK;
}

// Reduce sorted lists to the final top-k.  One launch when everything fits one CTA's shared
// memory (<= 32768/Kp lists), otherwise a tree of launches through ctx->d_merge_tmp.
qz_status launch_merge(qz_ctx *ctx, const uint64_t *d_in, uint32_t n_lists, uint32_t in_len, uint32_t in_stride,
                       uint32_t k, int largest, uint64_t *d_keys, float *d_scores, uint32_t *d_ids) {
    QZ_TRY(ensure_smem_attr(ctx, (const void *)merge_lists_kernel<0>, 128 * 1024));
    QZ_TRY(ensure_smem_attr(ctx, (const void *)prune_merge_kernel<0>, 128 * 1024));
    const uint32_t Kp = topk_pad(in_len > k ? in_len : k);
    const uint32_t max_lists = 32768 / Kp;  // (lists/2)*Kp keys <= 16384
    const uint64_t *in = d_in;
    uint64_t *bufs[2] = {ctx->d_merge_tmp, ctx->d_merge_tmp + (size_t)128 * QZ_MAX_K};
    int which = 0;
    for (;;) {
        uint32_t lpc = pow2_ceil(n_lists < 2 ? 2 : n_lists);
        if (lpc > max_lists) lpc = max_lists;
        const uint32_t grid = n_lists ? (n_lists + lpc - 1) / lpc : 1;
        const bool final_level = grid == 1;
        MergeParams p{};
        p.in = in;
        p.n_lists = n_lists;
        p.in_len = in_len;
        p.in_stride = in_stride;
        p.Kp = Kp;
        p.lists_per_cta = lpc;
        p.out = final_level ? nullptr : bufs[which];
        p.out_stride = Kp;
        p.out_len = Kp;
        p.final_out = final_level;
        p.k = k;
        p.flip = largest ? 0u : 0xFFFFFFFFu;
        p.keys = d_keys;
        p.scores = d_scores;
        p.ids = d_ids;
        size_t smem = (size_t)(lpc / 2) * Kp * 8;
        if (final_level && n_lists <= kPruneMaxLists) {
            if (smem < kPruneSmem) smem = kPruneSmem;
            prune_merge_kernel<0><<<1, 1024, smem, ctx->stream>>>(p);
        } else {
            merge_lists_kernel<0><<<grid, 1024, smem, ctx->stream>>>(p);
        }
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail(ctx, e, "merge_lists_kernel launch");
        ctx->launches++;
        if (final_level) break;
        in = bufs[which];
        n_lists = grid;
        in_len = Kp;
        in_stride = Kp;
        which ^= 1;
    }
    return QZ_OK;
}

}  // namespace qz
