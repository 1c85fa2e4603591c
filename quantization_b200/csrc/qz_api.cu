// qz_api.cu — index handles and the scan / score entry points of include/qz_cuda.h
#include <string.h>

#include "qz_internal.h"

using namespace qz;

namespace qz {
// pq.cu
qz_status pq_scan(qz_index *ix, const void *d_lut, float *d_scores, uint32_t k, int largest, uint32_t *n_lists,
                  uint32_t *Kp_out, const ScanFinal *fin);
qz_status pq_encode_query(qz_index *ix, const float *d_queries, uint32_t nq, void *d_out);
qz_status pq_score_points(qz_index *ix, const void *d_lut, const uint32_t *d_rows, uint64_t n, float *d_out);
qz_status pq_score_pairs(qz_index *ix, const uint32_t *d_a, const uint32_t *d_b, uint64_t n, float *d_out);
void pq_index_free(qz_index *ix);
// encode.cu
qz_status bin_encode_rows(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, uint32_t row_bytes,
                          uint8_t *d_out);

// allocate the device copy of an encoded storage: padded so that whole 16-byte chunks past the
// last row are readable (the streaming scan rounds its last tile up to 16 bytes)
qz_status index_alloc_storage(qz_index *ix, const uint8_t *host_storage) {
    qz_ctx *ctx = ix->ctx;
    const size_t bytes = (size_t)ix->count * ix->row_bytes;
    ix->alloc_bytes = (bytes + 15) / 16 * 16 + 256;
    QZ_CUDA(ctx, cudaMalloc((void **)&ix->d_storage, ix->alloc_bytes));
    if (host_storage && bytes) {
        QZ_CUDA(ctx, cudaMemsetAsync(ix->d_storage + bytes, 0, ix->alloc_bytes - bytes, ctx->stream));
        QZ_CUDA(ctx, cudaMemcpyAsync(ix->d_storage, host_storage, bytes, cudaMemcpyHostToDevice, ctx->stream));
        QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    } else {
        QZ_CUDA(ctx, cudaMemsetAsync(ix->d_storage, 0, ix->alloc_bytes, ctx->stream));
    }
    return QZ_OK;
}

// scan of any codec: every score to d_scores (optional) and/or per-CTA top-k lists in
// ctx->d_partial
static qz_status scan_any(qz_index *ix, const void *d_query, float *d_scores, uint32_t k, int largest,
                          uint32_t *n_lists, uint32_t *Kp, const ScanFinal *fin = nullptr) {
    qz_ctx *ctx = ix->ctx;
    if (ix->codec == QZ_CODEC_PQ) return pq_scan(ix, d_query, d_scores, k, largest, n_lists, Kp, fin);
    qz_status st = launch_row_scan(ix, d_query, d_scores, k, largest, n_lists, Kp, fin);
    if (st != QZ_ARGUMENTS) return st;
    // row shape outside the streaming kernel (rows that are not whole words, or huge rows):
    // warp-per-row gather kernel, then top-k over the score array
    ctx->err.clear();
    float *sc = d_scores;
    if (!sc) {
        QZ_TRY(ensure_io(ctx, ix->count * sizeof(float)));
        sc = reinterpret_cast<float *>(ctx->d_io);
    }
    QZ_TRY(launch_gather(ix, d_query, nullptr, nullptr, ix->count, sc));
    if (k) QZ_TRY(launch_scores_topk(ctx, sc, 1, 0, ix->count, nullptr, k, largest, ix->id_base, n_lists, Kp));
    return QZ_OK;
}
}  // namespace qz

static bool bad_index(const qz_index *ix) { return !ix || !ix->ctx; }

extern "C" {

uint32_t qz_u8_actual_dim(uint32_t dim) { return dim + (16 - dim % 16) % 16; }
uint32_t qz_u8_row_bytes(uint32_t dim) { return qz_u8_actual_dim(dim) + 4; }

qz_status qz_u8_index_create(qz_ctx *ctx, const uint8_t *storage, uint64_t count, uint32_t dim, int32_t distance,
                             int32_t invert, float alpha, float offset, float multiplier, uint32_t id_base,
                             qz_index **out) {
    if (!ctx || !out) return fail(ctx, QZ_ARGUMENTS, "qz_u8_index_create: NULL argument");
    *out = nullptr;
    if (distance < QZ_DOT || distance > QZ_L2) return fail(ctx, QZ_ARGUMENTS, "qz_u8_index_create: bad distance");
    if (dim == 0) return fail(ctx, QZ_ARGUMENTS, "qz_u8_index_create: dim is 0");
    if (count + id_base > 0xFFFFFFFEull) return fail(ctx, QZ_ARGUMENTS, "qz_u8_index_create: point ids exceed u32");
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    qz_index *ix = new qz_index();
    ix->ctx = ctx;
    ctx->refs.fetch_add(1);
    ix->codec = QZ_CODEC_U8;
    ix->count = count;
    ix->dim = dim;
    ix->actual_dim = qz_u8_actual_dim(dim);
    ix->row_bytes = ix->actual_dim + 4;
    ix->distance = distance;
    ix->invert = invert != 0;
    ix->alpha = alpha;
    ix->offset = offset;
    ix->multiplier = multiplier;
    ix->id_base = id_base;
    qz_status st = index_alloc_storage(ix, storage);
    if (st != QZ_OK) {
        cudaFree(ix->d_storage);
        ctx->refs.fetch_sub(1);
        delete ix;
        return st;
    }
    *out = ix;
    return QZ_OK;
}

qz_status qz_u8_index_set_params(qz_index *ix, float alpha, float offset, float multiplier) {
    if (bad_index(ix) || ix->codec != QZ_CODEC_U8) return QZ_ARGUMENTS;
    ix->alpha = alpha;
    ix->offset = offset;
    ix->multiplier = multiplier;
    ix->soa_valid = false;  // the batched path's mirror (built from the storage and the multiplier) is stale
    return QZ_OK;
}

void qz_index_destroy(qz_index *ix) {
    if (!ix) return;
    if (ix->ctx) {
        {
            std::lock_guard<std::mutex> g(ix->ctx->mu);
            cudaSetDevice(ix->ctx->device);
            cudaStreamSynchronize(ix->ctx->stream);
            cudaFree(ix->d_storage);
            cudaFree(ix->d_codes_soa);
            cudaFree(ix->d_voff_soa);
            cudaFree(ix->d_vq_soa);
            cudaFree(ix->d_batch);
            pq_index_free(ix);
        }
        ctx_release(ix->ctx);  // the context outlives its indexes even if qz_ctx_destroy came first
    }
    delete ix;
}

int32_t qz_index_codec(const qz_index *ix) { return ix ? ix->codec : -1; }
uint64_t qz_index_count(const qz_index *ix) { return ix ? ix->count : 0; }
uint32_t qz_index_row_bytes(const qz_index *ix) { return ix ? ix->row_bytes : 0; }
uint32_t qz_query_bytes(const qz_index *ix) {
    if (!ix) return 0;
    if (ix->codec == QZ_CODEC_PQ) return ix->n_chunks * 256 * sizeof(float);
    return ix->row_bytes;
}
void *qz_index_storage_dev(qz_index *ix) { return ix ? ix->d_storage : nullptr; }

qz_status qz_index_download(qz_index *ix, uint8_t *storage) {
    if (bad_index(ix) || !storage) return QZ_ARGUMENTS;
    qz_ctx *ctx = ix->ctx;
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    QZ_CUDA(ctx, cudaMemcpyAsync(storage, ix->d_storage, (size_t)ix->count * ix->row_bytes, cudaMemcpyDeviceToHost,
                                 ctx->stream));
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return QZ_OK;
}

qz_status qz_encode_query_dev(qz_index *ix, const float *d_queries, uint32_t nq, void *d_out) {
    if (bad_index(ix) || !d_queries || !d_out) return fail(ix ? ix->ctx : nullptr, QZ_ARGUMENTS, "encode_query: NULL");
    QZ_CUDA(ix->ctx, cudaSetDevice(ix->ctx->device));
    if (ix->codec == QZ_CODEC_U8) return launch_u8_encode_query(ix, d_queries, nq, d_out);
    if (ix->codec == QZ_CODEC_BIN) return bin_encode_rows(ix->ctx, d_queries, nq, ix->dim, ix->row_bytes, (uint8_t *)d_out);
    return pq_encode_query(ix, d_queries, nq, d_out);
}

qz_status qz_encode_query(qz_index *ix, const float *query, void *out) {
    if (bad_index(ix)) return QZ_ARGUMENTS;
    if (qz_query_bytes(ix) == 0) return QZ_OK;  // dim 0: nothing to encode
    if (!query || !out) return fail(ix->ctx, QZ_ARGUMENTS, "encode_query: NULL");
    qz_ctx *ctx = ix->ctx;
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t qb = qz_query_bytes(ix), fb = (size_t)ix->dim * sizeof(float);
    const size_t fb_pad = (fb + 255) / 256 * 256;
    if (qb == 0) return QZ_OK;
    QZ_TRY(ensure_query(ctx, fb_pad + qb));
    QZ_TRY(ensure_pin(ctx, fb_pad + qb));
    memcpy(ctx->h_pin, query, fb);
    QZ_CUDA(ctx, cudaMemcpyAsync(ctx->d_query, ctx->h_pin, fb, cudaMemcpyHostToDevice, ctx->stream));
    QZ_TRY(qz_encode_query_dev(ix, reinterpret_cast<const float *>(ctx->d_query), 1, ctx->d_query + fb_pad));
    QZ_CUDA(ctx, cudaMemcpyAsync(ctx->h_pin + fb_pad, ctx->d_query + fb_pad, qb, cudaMemcpyDeviceToHost, ctx->stream));
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    memcpy(out, ctx->h_pin + fb_pad, qb);
    return QZ_OK;
}

qz_status qz_scan_topk_dev(qz_index *ix, const void *d_enc_query, uint32_t k, int32_t largest, uint64_t *d_keys,
                           float *d_scores, uint32_t *d_ids) {
    if (bad_index(ix)) return QZ_ARGUMENTS;
    qz_ctx *ctx = ix->ctx;
    if (k == 0 || k > QZ_MAX_K) return fail(ctx, QZ_ARGUMENTS, "scan_topk: k must be in [1, QZ_MAX_K]");
    if (!d_enc_query && qz_query_bytes(ix)) return fail(ctx, QZ_ARGUMENTS, "scan_topk: NULL query");
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t n_lists = 0, Kp = topk_pad(k);
    bool fused = false;
    const ScanFinal fin{d_keys, d_scores, d_ids, &fused};
    if (ix->count) QZ_TRY(scan_any(ix, d_enc_query, nullptr, k, largest, &n_lists, &Kp, &fin));
    if (fused) return QZ_OK;
    return launch_merge(ctx, ctx->d_partial, n_lists, Kp, Kp, k, largest, d_keys, d_scores, d_ids);
}

qz_status qz_scan_topk(qz_index *ix, const void *enc_query, uint32_t k, int32_t largest, float *scores, uint32_t *ids,
                       uint32_t *n_out) {
    if (bad_index(ix)) return QZ_ARGUMENTS;
    qz_ctx *ctx = ix->ctx;
    if (!scores || !ids) return fail(ctx, QZ_ARGUMENTS, "scan_topk: NULL output");
    if (k == 0 || k > QZ_MAX_K) return fail(ctx, QZ_ARGUMENTS, "scan_topk: k must be in [1, QZ_MAX_K]");
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t qb = qz_query_bytes(ix), qb_pad = (qb + 255) / 256 * 256;
    if (qb && !enc_query) return fail(ctx, QZ_ARGUMENTS, "scan_topk: NULL query");
    QZ_TRY(ensure_query(ctx, qb_pad + 16));
    QZ_TRY(ensure_pin(ctx, qb_pad + (size_t)2 * QZ_MAX_K * 4));
    if (qb) {
        memcpy(ctx->h_pin, enc_query, qb);
        QZ_CUDA(ctx, cudaMemcpyAsync(ctx->d_query, ctx->h_pin, qb, cudaMemcpyHostToDevice, ctx->stream));
    }
    // the k ids go right behind the k scores: one copy of 2 * k * 4 bytes brings both back
    QZ_TRY(qz_scan_topk_dev(ix, ctx->d_query, k, largest, nullptr, ctx->d_scores_k,
                            reinterpret_cast<uint32_t *>(ctx->d_scores_k + k)));
    float *hs = reinterpret_cast<float *>(ctx->h_pin + qb_pad);
    uint32_t *hi = reinterpret_cast<uint32_t *>(hs + k);
    QZ_CUDA(ctx, cudaMemcpyAsync(hs, ctx->d_scores_k, (size_t)2 * k * 4, cudaMemcpyDeviceToHost, ctx->stream));
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const uint32_t m = (uint32_t)(ix->count < k ? ix->count : k);
    memcpy(scores, hs, (size_t)m * 4);
    memcpy(ids, hi, (size_t)m * 4);
    if (n_out) *n_out = m;
    return QZ_OK;
}

qz_status qz_search_topk(qz_index *ix, const float *query, uint32_t k, int32_t largest, float *scores, uint32_t *ids,
                         uint32_t *n_out) {
    if (bad_index(ix)) return QZ_ARGUMENTS;
    qz_ctx *ctx = ix->ctx;
    if (!scores || !ids) return fail(ctx, QZ_ARGUMENTS, "search_topk: NULL output");
    if (k == 0 || k > QZ_MAX_K) return fail(ctx, QZ_ARGUMENTS, "search_topk: k must be in [1, QZ_MAX_K]");
    const size_t qb = qz_query_bytes(ix), fb = (size_t)ix->dim * sizeof(float);
    if (fb && !query) return fail(ctx, QZ_ARGUMENTS, "search_topk: NULL query");
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t fb_pad = (fb + 255) / 256 * 256, qb_pad = (qb + 255) / 256 * 256;
    QZ_TRY(ensure_query(ctx, fb_pad + qb_pad + 16));
    QZ_TRY(ensure_pin(ctx, fb_pad + (size_t)2 * QZ_MAX_K * 4));
    // No copy operations: the staging buffer is pinned host memory, which the device reads and writes in place (unified
    // addressing).  The encode kernel reads the f32 query from it over PCIe (6 KB), the scan's last thread block
    // stores the k scores and ids into it, and the only host-side wait is the stream synchronisation: two
    // cudaMemcpyAsync (~5 us each of launch + DMA latency) less per query.
    uint8_t *d_blob = ctx->d_query + fb_pad;
    if (fb && qb) {
        memcpy(ctx->h_pin, query, fb);
        QZ_TRY(qz_encode_query_dev(ix, reinterpret_cast<const float *>(ctx->h_pin), 1, d_blob));
        // the scan is launched as the programmatic dependent of the encode kernel (u8: the kernel that triggers early)
        ctx->query_dep_next = ix->codec == QZ_CODEC_U8;
    }
    float *hs = reinterpret_cast<float *>(ctx->h_pin + fb_pad);
    uint32_t *hi = reinterpret_cast<uint32_t *>(hs + k);
    const qz_status scan_st = qz_scan_topk_dev(ix, d_blob, k, largest, nullptr, hs, hi);
    ctx->query_dep_next = false;
    QZ_TRY(scan_st);
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const uint32_t m = (uint32_t)(ix->count < k ? ix->count : k);
    memcpy(scores, hs, (size_t)m * 4);
    memcpy(ids, hi, (size_t)m * 4);
    if (n_out) *n_out = m;
    return QZ_OK;
}

qz_status qz_scan_scores_dev(qz_index *ix, const void *d_enc_query, float *d_scores) {
    if (bad_index(ix)) return QZ_ARGUMENTS;
    if (ix->count == 0) return QZ_OK;
    if (!d_scores) return fail(ix->ctx, QZ_ARGUMENTS, "scan_scores: NULL output");
    QZ_CUDA(ix->ctx, cudaSetDevice(ix->ctx->device));
    return scan_any(ix, d_enc_query, d_scores, 0, 1, nullptr, nullptr);
}

qz_status qz_scan_scores(qz_index *ix, const void *enc_query, float *scores) {
    if (bad_index(ix)) return QZ_ARGUMENTS;
    qz_ctx *ctx = ix->ctx;
    if (ix->count == 0) return QZ_OK;
    if (!scores) return fail(ctx, QZ_ARGUMENTS, "scan_scores: NULL output");
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t qb = qz_query_bytes(ix), qb_pad = (qb + 255) / 256 * 256;
    if (qb && !enc_query) return fail(ctx, QZ_ARGUMENTS, "scan_scores: NULL query");
    const size_t sb = ix->count * sizeof(float);
    QZ_TRY(ensure_query(ctx, qb_pad + 16));
    QZ_TRY(ensure_pin(ctx, qb_pad));
    float *d_sc = nullptr;
    QZ_CUDA(ctx, cudaMalloc((void **)&d_sc, sb));
    if (qb) {
        memcpy(ctx->h_pin, enc_query, qb);
        QZ_CUDA(ctx, cudaMemcpyAsync(ctx->d_query, ctx->h_pin, qb, cudaMemcpyHostToDevice, ctx->stream));
    }
    qz_status st = scan_any(ix, ctx->d_query, d_sc, 0, 1, nullptr, nullptr);
    cudaError_t e = cudaSuccess;
    if (st == QZ_OK) e = cudaMemcpyAsync(scores, d_sc, sb, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_sc);
    if (st != QZ_OK) return st;
    if (e != cudaSuccess) return cuda_fail(ctx, e, "scan_scores");
    return QZ_OK;
}

qz_status qz_merge_topk_dev(qz_ctx *ctx, const uint64_t *d_keys_in, uint32_t n_lists, uint32_t k, int32_t largest,
                            uint64_t *d_keys, float *d_scores, uint32_t *d_ids) {
    if (!ctx) return QZ_ARGUMENTS;
    if (k == 0 || k > QZ_MAX_K) return fail(ctx, QZ_ARGUMENTS, "merge_topk: k must be in [1, QZ_MAX_K]");
    if (n_lists && !d_keys_in) return fail(ctx, QZ_ARGUMENTS, "merge_topk: NULL input");
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_merge(ctx, d_keys_in, n_lists, k, k, k, largest, d_keys, d_scores, d_ids);
}

}  // extern "C"

// stage an id list on the device, run `body(d_ids...)`, bring n floats back
static qz_status score_ids_common(qz_index *ix, const void *enc_query, const uint32_t *rows_a, const uint32_t *rows_b,
                                  uint64_t n, float *scores, bool pairs) {
    qz_ctx *ctx = ix->ctx;
    if (n == 0) return QZ_OK;
    if (!rows_b || !scores || (pairs && !rows_a)) return fail(ctx, QZ_ARGUMENTS, "score: NULL argument");
    for (uint64_t i = 0; i < n; i++)
        if (rows_b[i] >= ix->count || (pairs && rows_a[i] >= ix->count))
            return fail(ctx, QZ_ARGUMENTS, "score: row index out of range");
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t qb = pairs ? 0 : qz_query_bytes(ix), qb_pad = (qb + 255) / 256 * 256;
    if (qb && !enc_query) return fail(ctx, QZ_ARGUMENTS, "score: NULL query");
    const size_t ib = n * sizeof(uint32_t), ib_pad = (ib + 255) / 256 * 256;
    const size_t tot = 2 * ib_pad + n * sizeof(float);
    QZ_TRY(ensure_query(ctx, qb_pad + 16));
    QZ_TRY(ensure_io(ctx, tot));
    QZ_TRY(ensure_pin(ctx, qb_pad > tot ? qb_pad : tot));
    if (qb) {
        memcpy(ctx->h_pin, enc_query, qb);
        QZ_CUDA(ctx, cudaMemcpyAsync(ctx->d_query, ctx->h_pin, qb, cudaMemcpyHostToDevice, ctx->stream));
        QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    memcpy(ctx->h_pin, rows_b, ib);
    if (pairs) memcpy(ctx->h_pin + ib_pad, rows_a, ib);
    QZ_CUDA(ctx, cudaMemcpyAsync(ctx->d_io, ctx->h_pin, pairs ? ib_pad + ib : ib, cudaMemcpyHostToDevice, ctx->stream));
    const uint32_t *d_b = reinterpret_cast<const uint32_t *>(ctx->d_io);
    const uint32_t *d_a = pairs ? reinterpret_cast<const uint32_t *>(ctx->d_io + ib_pad) : nullptr;
    float *d_out = reinterpret_cast<float *>(ctx->d_io + 2 * ib_pad);
    if (ix->codec == QZ_CODEC_PQ) {
        if (pairs) QZ_TRY(pq_score_pairs(ix, d_a, d_b, n, d_out));
        else QZ_TRY(pq_score_points(ix, ctx->d_query, d_b, n, d_out));
    } else {
        QZ_TRY(launch_gather(ix, pairs ? nullptr : ctx->d_query, d_a, d_b, n, d_out));
    }
    QZ_CUDA(ctx, cudaMemcpyAsync(ctx->h_pin, d_out, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    memcpy(scores, ctx->h_pin, n * sizeof(float));
    return QZ_OK;
}

extern "C" {

qz_status qz_score_points(qz_index *ix, const void *enc_query, const uint32_t *rows, uint64_t n, float *scores) {
    if (bad_index(ix)) return QZ_ARGUMENTS;
    return score_ids_common(ix, enc_query, nullptr, rows, n, scores, false);
}

qz_status qz_score_pairs(qz_index *ix, const uint32_t *rows_a, const uint32_t *rows_b, uint64_t n, float *scores) {
    if (bad_index(ix)) return QZ_ARGUMENTS;
    return score_ids_common(ix, nullptr, rows_a, rows_b, n, scores, true);
}

// device-pointer forms: ids, query and outputs already on the device, asynchronous, ids are NOT range-checked
qz_status qz_score_points_dev(qz_index *ix, const void *d_enc_query, const uint32_t *d_rows, uint64_t n,
                              float *d_scores) {
    if (bad_index(ix)) return QZ_ARGUMENTS;
    if (n == 0) return QZ_OK;
    if (!d_rows || !d_scores || (!d_enc_query && qz_query_bytes(ix)))
        return fail(ix->ctx, QZ_ARGUMENTS, "score_points_dev: NULL argument");
    QZ_CUDA(ix->ctx, cudaSetDevice(ix->ctx->device));
    if (ix->codec == QZ_CODEC_PQ) return pq_score_points(ix, d_enc_query, d_rows, n, d_scores);
    return launch_gather(ix, d_enc_query, nullptr, d_rows, n, d_scores);
}

qz_status qz_score_pairs_dev(qz_index *ix, const uint32_t *d_rows_a, const uint32_t *d_rows_b, uint64_t n,
                             float *d_scores) {
    if (bad_index(ix)) return QZ_ARGUMENTS;
    if (n == 0) return QZ_OK;
    if (!d_rows_a || !d_rows_b || !d_scores) return fail(ix->ctx, QZ_ARGUMENTS, "score_pairs_dev: NULL argument");
    QZ_CUDA(ix->ctx, cudaSetDevice(ix->ctx->device));
    if (ix->codec == QZ_CODEC_PQ) return pq_score_pairs(ix, d_rows_a, d_rows_b, n, d_scores);
    return launch_gather(ix, nullptr, d_rows_a, d_rows_b, n, d_scores);
}

}  // extern "C"

extern "C" {

/* get_storage_size * size_of::<T>(), encoded_vectors_binary.rs:99-116 (u8 words), :152-159 (u128 words) */
uint32_t qz_bin_row_bytes(uint32_t dim, int32_t store_bytes) {
    if (store_bytes == 16) return ((dim + 127) / 128) * 16;
    const uint32_t bytes = dim > 128 ? 16 : (dim > 64 ? 8 : (dim > 32 ? 4 : 1));
    const uint32_t bits = 8 * bytes;
    return ((dim + bits - 1) / bits) * bytes;
}

qz_status qz_bin_index_create(qz_ctx *ctx, const uint8_t *storage, uint64_t count, uint32_t dim, int32_t store_bytes,
                              int32_t distance, int32_t invert, uint32_t id_base, qz_index **out) {
    if (!ctx || !out) return fail(ctx, QZ_ARGUMENTS, "qz_bin_index_create: NULL argument");
    *out = nullptr;
    if (distance < QZ_DOT || distance > QZ_L2) return fail(ctx, QZ_ARGUMENTS, "qz_bin_index_create: bad distance");
    if (store_bytes != 1 && store_bytes != 16)
        return fail(ctx, QZ_ARGUMENTS, "qz_bin_index_create: store_bytes must be 1 (u8) or 16 (u128)");
    if (count + id_base > 0xFFFFFFFEull) return fail(ctx, QZ_ARGUMENTS, "qz_bin_index_create: point ids exceed u32");
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    qz_index *ix = new qz_index();
    ix->ctx = ctx;
    ctx->refs.fetch_add(1);
    ix->codec = QZ_CODEC_BIN;
    ix->count = count;
    ix->dim = dim;
    ix->store = store_bytes;
    ix->row_bytes = qz_bin_row_bytes(dim, store_bytes);
    ix->distance = distance;
    ix->invert = invert != 0;
    ix->id_base = id_base;
    qz_status st = index_alloc_storage(ix, storage);
    if (st != QZ_OK) {
        cudaFree(ix->d_storage);
        ctx->refs.fetch_sub(1);
        delete ix;
        return st;
    }
    *out = ix;
    return QZ_OK;
}

}  // extern "C"
