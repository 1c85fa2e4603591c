// chunked.cu — out-of-core scan: the encoded storage stays in host memory and is streamed through two device
// buffers (EncodedStorage backends that are not one resident buffer, encoded_storage.rs:7-25;
// corpora larger than HBM).  Copy of chunk c+1 (a second stream) overlaps the scan of chunk c; every chunk leaves
// its k best keys (global point ids) in a list and one merge ends the query.  PCIe-bound: the kernels are the
// resident ones, untouched.
#include <string.h>

#include <algorithm>
#include <string>

#include "qz_internal.h"

using namespace qz;

extern "C" {

qz_status qz_host_pin(qz_ctx *ctx, const void *ptr, uint64_t bytes) {
    if (!ctx) return QZ_ARGUMENTS;
    if (!ptr || !bytes) return QZ_OK;
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    QZ_CUDA(ctx, cudaHostRegister(const_cast<void *>(ptr), bytes, cudaHostRegisterDefault));
    return QZ_OK;
}

qz_status qz_host_unpin(qz_ctx *ctx, const void *ptr) {
    if (!ctx) return QZ_ARGUMENTS;
    if (!ptr) return QZ_OK;
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    QZ_CUDA(ctx, cudaHostUnregister(const_cast<void *>(ptr)));
    return QZ_OK;
}

qz_status qz_scan_topk_chunked(qz_index *a, qz_index *b, const uint8_t *host_storage, uint64_t count,
                               const void *enc_query, uint32_t k, int32_t largest, float *scores, uint32_t *ids,
                               uint32_t *n_out) {
    if (!a || !b || !a->ctx || a->ctx != b->ctx) return QZ_ARGUMENTS;
    qz_ctx *ctx = a->ctx;
    if (a == b || a->codec != b->codec || a->row_bytes != b->row_bytes || a->count != b->count || a->count == 0)
        return fail(ctx, QZ_ARGUMENTS, "scan_topk_chunked: needs two distinct buffer indexes of the same shape, capacity > 0");
    if (!scores || !ids) return fail(ctx, QZ_ARGUMENTS, "scan_topk_chunked: NULL output");
    if (k == 0 || k > QZ_MAX_K) return fail(ctx, QZ_ARGUMENTS, "scan_topk_chunked: k must be in [1, QZ_MAX_K]");
    if (count && !host_storage) return fail(ctx, QZ_ARGUMENTS, "scan_topk_chunked: NULL storage");
    if (count > 0xFFFFFFFEull) return fail(ctx, QZ_ARGUMENTS, "scan_topk_chunked: point ids exceed u32");
    const size_t qb = qz_query_bytes(a);
    if (qb && !enc_query) return fail(ctx, QZ_ARGUMENTS, "scan_topk_chunked: NULL query");
    if (n_out) *n_out = 0;
    if (count == 0) return QZ_OK;
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_tail = 0;  // host-buffer entry points copy and synchronise: never part of a scan chain
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint64_t cap = a->count, row = a->row_bytes;
    const uint64_t n_chunks = (count + cap - 1) / cap;
    {   // launch_merge reduces the per-chunk lists through ctx->d_merge_tmp (two halves of 128 * QZ_MAX_K keys): the
        // first tree level leaves ceil(n_chunks / (32768 / Kp)) lists of Kp keys there
        const uint64_t Kp = topk_pad(k), per_cta = 32768 / Kp;
        const uint64_t max_chunks = std::min<uint64_t>(32768, per_cta * ((128ull * QZ_MAX_K) / Kp));
        if (n_chunks > max_chunks)
            return fail(ctx, QZ_ARGUMENTS, "scan_topk_chunked: too many chunks for this k (" + std::to_string(n_chunks) +
                                               " > " + std::to_string(max_chunks) + "); use bigger buffers");
    }
    const size_t qb_pad = (qb + 255) / 256 * 256;
    QZ_TRY(ensure_query(ctx, qb_pad + 16));
    QZ_TRY(ensure_pin(ctx, qb_pad > (size_t)k * 8 ? qb_pad : (size_t)k * 8));
    if (qb) {
        memcpy(ctx->h_pin, enc_query, qb);
        QZ_CUDA(ctx, cudaMemcpyAsync(ctx->d_query, ctx->h_pin, qb, cudaMemcpyHostToDevice, ctx->stream));
    }
    uint64_t *d_lists = nullptr;
    cudaStream_t copy = nullptr;
    cudaEvent_t ready[2] = {nullptr, nullptr}, freed[2] = {nullptr, nullptr};
    qz_index *buf[2] = {a, b};
    const uint32_t base0[2] = {a->id_base, b->id_base};
    qz_status st = QZ_OK;
    cudaError_t e = cudaMalloc((void **)&d_lists, n_chunks * k * sizeof(uint64_t));
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&copy, cudaStreamNonBlocking);
    for (int i = 0; i < 2 && e == cudaSuccess; i++) {
        e = cudaEventCreateWithFlags(&ready[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&freed[i], cudaEventDisableTiming);
    }
    // the buffers may still be in use by earlier work on the context stream
    if (e == cudaSuccess) e = cudaEventRecord(freed[0], ctx->stream);
    if (e == cudaSuccess) e = cudaEventRecord(freed[1], ctx->stream);
    for (uint64_t c = 0; c < n_chunks && e == cudaSuccess && st == QZ_OK; c++) {
        qz_index *ix = buf[c & 1];
        const uint64_t r0 = c * cap, rows = count - r0 < cap ? count - r0 : cap;
        e = cudaStreamWaitEvent(copy, freed[c & 1], 0);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(ix->d_storage, host_storage + r0 * row, rows * row, cudaMemcpyHostToDevice, copy);
        if (e == cudaSuccess) e = cudaEventRecord(ready[c & 1], copy);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stream, ready[c & 1], 0);
        if (e != cudaSuccess) break;
        ix->count = rows;
        ix->id_base = (uint32_t)r0;
        st = qz_scan_topk_dev(ix, ctx->d_query, k, largest, d_lists + c * k, nullptr, nullptr);
        ix->count = cap;
        if (st == QZ_OK) e = cudaEventRecord(freed[c & 1], ctx->stream);
    }
    a->id_base = base0[0];
    b->id_base = base0[1];
    if (e == cudaSuccess && st == QZ_OK)
        st = launch_merge(ctx, d_lists, (uint32_t)n_chunks, k, k, k, largest, nullptr, ctx->d_scores_k, ctx->d_ids_k);
    if (e == cudaSuccess && st == QZ_OK) {
        float *hs = reinterpret_cast<float *>(ctx->h_pin);
        uint32_t *hi = reinterpret_cast<uint32_t *>(ctx->h_pin + (size_t)k * 4);
        e = cudaMemcpyAsync(hs, ctx->d_scores_k, (size_t)k * 4, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(hi, ctx->d_ids_k, (size_t)k * 4, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e == cudaSuccess) {
            const uint32_t m = (uint32_t)(count < k ? count : k);
            memcpy(scores, hs, (size_t)m * 4);
            memcpy(ids, hi, (size_t)m * 4);
            if (n_out) *n_out = m;
        }
    }
    cudaStreamSynchronize(ctx->stream);
    if (copy) {
        cudaStreamSynchronize(copy);
        cudaStreamDestroy(copy);
    }
    for (int i = 0; i < 2; i++) {
        if (ready[i]) cudaEventDestroy(ready[i]);
        if (freed[i]) cudaEventDestroy(freed[i]);
    }
    cudaFree(d_lists);
    if (e != cudaSuccess) return cuda_fail(ctx, e, "scan_topk_chunked");
    return st;
}

}  // extern "C"
