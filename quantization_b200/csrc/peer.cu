// peer.cu — multi-GPU top-k exchange over peer memory (NVLink / NVSwitch), fused into the scan launch.
//
// Shards are contiguous point-id ranges, one process per GPU.  The per-query exchange is tiny (k keys
// per GPU), so it is latency- not bandwidth-bound: instead of a host-launched collective plus a merge launch, the
// scan kernel's last CTA stores its keys straight into every peer's buffer, flags them, waits for the peers' flags
// and merges (row_scan.cuh, PeerExchange).  This file owns the symmetric buffers: allocation, CUDA IPC handles
// (exchanged by the caller with whatever transport it has — torch.distributed in sharded.py), peer mapping.
#include <string.h>

#include "qz_internal.h"
#include "row_scan.cuh"

struct qz_peer {
    qz_ctx *ctx = nullptr;
    uint32_t world = 0, rank = 0, seq = 0;
    uint32_t last_read[2][qz::kMaxPeers] = {};  // per buffer half and rank: the last step in which that rank read the half
    cudaEvent_t staged = nullptr;     // rooted host form on a non-root rank: the query's H2D copy out of the pinned
    bool staged_pending = false;      // staging buffer (the call returns without synchronising; the next one must
                                      // not overwrite the staging buffer before that copy has run)
    uint8_t *local = nullptr;            // gather[2][world][kPeerStride] u64, then flags[2][world] u32
    uint8_t *mapped[qz::kMaxPeers] = {};  // peers' buffers in this process (mapped[rank] = local)
    bool connected = false;
};

namespace qz {
static size_t gather_bytes(uint32_t world) { return (size_t)2 * world * kPeerStride * sizeof(uint64_t); }
static size_t peer_bytes(uint32_t world) { return gather_bytes(world) + 256; }
}  // namespace qz

namespace qz {
qz_status pq_scan(qz_index *ix, const void *d_lut, float *d_scores, uint32_t k, int largest, uint32_t *n_lists,
                  uint32_t *Kp_out, const ScanFinal *fin);
}

using namespace qz;

extern "C" {

qz_status qz_peer_create(qz_ctx *ctx, uint32_t world, uint32_t rank, uint8_t *handle_out, qz_peer **out) {
    if (!ctx || !out || !handle_out) return QZ_ARGUMENTS;
    *out = nullptr;
    if (world < 1 || world > (uint32_t)kMaxPeers || rank >= world)
        return fail(ctx, QZ_ARGUMENTS, "qz_peer_create: world must be 1..8 and rank < world");
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    qz_peer *pe = new qz_peer();
    pe->ctx = ctx;
    pe->world = world;
    pe->rank = rank;
    cudaError_t e = cudaMalloc((void **)&pe->local, peer_bytes(world));
    if (e == cudaSuccess) e = cudaMemset(pe->local, 0, peer_bytes(world));
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, pe->local);
    if (e != cudaSuccess) {
        cudaFree(pe->local);
        delete pe;
        return cuda_fail(ctx, e, "qz_peer_create");
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == QZ_PEER_HANDLE_BYTES, "handle size");
    memcpy(handle_out, &h, sizeof h);
    pe->mapped[rank] = pe->local;
    ctx->refs.fetch_add(1);
    *out = pe;
    return QZ_OK;
}

qz_status qz_peer_connect(qz_peer *pe, const uint8_t *all_handles) {
    if (!pe || !all_handles) return QZ_ARGUMENTS;
    qz_ctx *ctx = pe->ctx;
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    for (uint32_t r = 0; r < pe->world; r++) {
        if (r == pe->rank || pe->mapped[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, all_handles + (size_t)r * QZ_PEER_HANDLE_BYTES, sizeof h);
        void *ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return cuda_fail(ctx, e, "qz_peer_connect: cudaIpcOpenMemHandle");
        pe->mapped[r] = static_cast<uint8_t *>(ptr);
    }
    pe->connected = true;
    return QZ_OK;
}

void qz_peer_destroy(qz_peer *pe) {
    if (!pe) return;
    qz_ctx *ctx = pe->ctx;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (uint32_t r = 0; r < pe->world; r++)
        if (r != pe->rank && pe->mapped[r]) cudaIpcCloseMemHandle(pe->mapped[r]);
    cudaFree(pe->local);
    if (pe->staged) cudaEventDestroy(pe->staged);
    delete pe;
    ctx_release(ctx);
}

static qz_status scan_topk_exchange(qz_index *ix, qz_peer *pe, int32_t root, const void *d_enc_query, uint32_t k,
                                    int32_t largest, uint64_t *d_keys, float *d_scores, uint32_t *d_ids) {
    if (!ix || !ix->ctx || !pe) return QZ_ARGUMENTS;
    qz_ctx *ctx = ix->ctx;
    if (pe->ctx != ctx) return fail(ctx, QZ_ARGUMENTS, "scan_topk_peer: index and peer group live on different contexts");
    if (!pe->connected && pe->world > 1) return fail(ctx, QZ_ARGUMENTS, "scan_topk_peer: qz_peer_connect first");
    if (k == 0 || k > kPeerStride) return fail(ctx, QZ_ARGUMENTS, "scan_topk_peer: k must be in [1, 128]");
    if (ix->count == 0) return fail(ctx, QZ_ARGUMENTS, "scan_topk_peer: empty shard");
    if (!d_enc_query) return fail(ctx, QZ_ARGUMENTS, "scan_topk_peer: NULL query");
    if (root >= (int32_t)pe->world) return fail(ctx, QZ_ARGUMENTS, "scan_topk_gather: root must be a rank of the group");
    if (pe->world == 1) return qz_scan_topk_dev(ix, d_enc_query, k, largest, d_keys, d_scores, d_ids);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    PeerExchange x{};
    x.world = pe->world;
    x.rank = pe->rank;
    x.seq = ++pe->seq;
    x.buf = x.seq & 1;
    x.root = root;
    for (uint32_t r = 0; r < pe->world; r++) {
        x.need_ack[r] = pe->last_read[x.buf][r];  // (see PeerExchange: flow control)
        x.gather[r] = reinterpret_cast<uint64_t *>(pe->mapped[r]);
        x.flags[r] = reinterpret_cast<uint32_t *>(pe->mapped[r] + gather_bytes(pe->world));
    }
    MergeParams &m = x.merge;
    m.in = x.gather[x.rank] + (size_t)x.buf * x.world * kPeerStride;
    m.n_lists = x.world;
    m.in_len = k;
    m.in_stride = kPeerStride;
    m.Kp = 128;
    m.lists_per_cta = pow2_ceil(x.world < 2 ? 2 : x.world);
    m.out = nullptr;
    m.final_out = 1;
    m.k = k;
    m.flip = largest ? 0u : 0xFFFFFFFFu;
    m.keys = d_keys;
    m.scores = d_scores;
    m.ids = d_ids;
    bool fused = false;
    ScanFinal fin{nullptr, nullptr, nullptr, &fused, &x};
    uint32_t n_lists = 0, Kp = 0;
    qz_status st;
    if (ix->codec == QZ_CODEC_PQ) {
        st = pq_scan(ix, d_enc_query, nullptr, k, largest, &n_lists, &Kp, &fin);
        if (st == QZ_OK && !fused) st = fail(ctx, QZ_ARGUMENTS, "scan_topk_peer: PQ row pitch must be a multiple of 16 bytes");
    } else {
        st = launch_row_scan(ix, d_enc_query, nullptr, k, largest, &n_lists, &Kp, &fin);
    }
    if (st != QZ_OK) {
        pe->seq--;  // the exchange did not happen
        return st;
    }
    for (uint32_t r = 0; r < pe->world; r++)
        if (root < 0 || (uint32_t)root == r) pe->last_read[x.buf][r] = x.seq;
    return st;
}

qz_status qz_scan_topk_peer_dev(qz_index *ix, qz_peer *pe, const void *d_enc_query, uint32_t k, int32_t largest,
                                uint64_t *d_keys, float *d_scores, uint32_t *d_ids) {
    return scan_topk_exchange(ix, pe, -1, d_enc_query, k, largest, d_keys, d_scores, d_ids);
}

qz_status qz_scan_topk_gather_dev(qz_index *ix, qz_peer *pe, uint32_t root, const void *d_enc_query, uint32_t k,
                                  int32_t largest, uint64_t *d_keys, float *d_scores, uint32_t *d_ids) {
    return scan_topk_exchange(ix, pe, (int32_t)root, d_enc_query, k, largest, d_keys, d_scores, d_ids);
}

// encode_query + scan + cross-GPU exchange, host buffers, one synchronisation (the multi-GPU form of qz_search_topk).
// *n_out = number of results = min(k, sum of the shards' rows) — unfilled slots are dropped; rooted form: 0 on a
// rank that is not the root (its outputs are not written and it does not wait for anyone).
static qz_status search_topk_exchange(qz_index *ix, qz_peer *pe, int32_t root, const float *query, uint32_t k,
                                      int32_t largest, float *scores, uint32_t *ids, uint32_t *n_out) {
    if (!ix || !ix->ctx || !pe) return QZ_ARGUMENTS;
    qz_ctx *ctx = ix->ctx;
    if (!scores || !ids || !query) return fail(ctx, QZ_ARGUMENTS, "search_topk_peer: NULL argument");
    if (k == 0 || k > kPeerStride) return fail(ctx, QZ_ARGUMENTS, "search_topk_peer: k must be in [1, 128]");
    const size_t qb = qz_query_bytes(ix), fb = (size_t)ix->dim * sizeof(float);
    if (!fb || !qb) return fail(ctx, QZ_ARGUMENTS, "search_topk_peer: dim 0");
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t fb_pad = (fb + 255) / 256 * 256, qb_pad = (qb + 255) / 256 * 256;
    QZ_TRY(ensure_query(ctx, fb_pad + qb_pad + 16));
    QZ_TRY(ensure_pin(ctx, fb_pad + (size_t)2 * QZ_MAX_K * 4));
    uint8_t *d_blob = ctx->d_query + fb_pad;
    if (pe->staged_pending) {  // a non-root call may still be reading the staged query
        QZ_CUDA(ctx, cudaEventSynchronize(pe->staged));
        pe->staged_pending = false;
    }
    // no copy operations: the pinned staging buffer is read (f32 query) and written (k scores, k ids) by the kernels
    // in place, as in qz_search_topk
    memcpy(ctx->h_pin, query, fb);
    QZ_TRY(qz_encode_query_dev(ix, reinterpret_cast<const float *>(ctx->h_pin), 1, d_blob));
    if (root >= 0 && pe->world > 1 && (uint32_t)root != pe->rank) {
        if (!pe->staged) QZ_CUDA(ctx, cudaEventCreateWithFlags(&pe->staged, cudaEventDisableTiming));
        QZ_CUDA(ctx, cudaEventRecord(pe->staged, ctx->stream));  // (after the kernel that reads the staged query)
        pe->staged_pending = true;
    } else {
        // nothing sits between the encode kernel and the scan: programmatic dependent launch, as in qz_search_topk
        ctx->query_dep_next = ix->codec == QZ_CODEC_U8;
    }
    float *hs = reinterpret_cast<float *>(ctx->h_pin + fb_pad);
    uint32_t *hi = reinterpret_cast<uint32_t *>(hs + k);
    const qz_status scan_st = scan_topk_exchange(ix, pe, root, d_blob, k, largest, nullptr, hs, hi);
    ctx->query_dep_next = false;
    QZ_TRY(scan_st);
    if (root >= 0 && pe->world > 1 && (uint32_t)root != pe->rank) {  // not the root: nothing comes back
        if (n_out) *n_out = 0;
        return QZ_OK;  // (asynchronous: the launch is on the stream; the caller's next call or qz_ctx_synchronize orders it)
    }
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    uint32_t m = 0;
    while (m < k && hi[m] != 0xFFFFFFFFu) m++;  // empty slots decode to id 0xFFFFFFFF
    memcpy(scores, hs, (size_t)m * 4);
    memcpy(ids, hi, (size_t)m * 4);
    if (n_out) *n_out = m;
    return QZ_OK;
}

qz_status qz_search_topk_peer(qz_index *ix, qz_peer *pe, const float *query, uint32_t k, int32_t largest, float *scores,
                              uint32_t *ids, uint32_t *n_out) {
    return search_topk_exchange(ix, pe, -1, query, k, largest, scores, ids, n_out);
}

qz_status qz_search_topk_gather(qz_index *ix, qz_peer *pe, uint32_t root, const float *query, uint32_t k, int32_t largest,
                                float *scores, uint32_t *ids, uint32_t *n_out) {
    return search_topk_exchange(ix, pe, (int32_t)root, query, k, largest, scores, ids, n_out);
}

}  // extern "C"
