// qz_ctx.cu — context, error plumbing, scratch management (C-ABI of include/qz_cuda.h)
#include <string.h>

#include "qz_internal.h"

static thread_local std::string g_global_err;

namespace qz {

qz_status fail(qz_ctx *ctx, qz_status st, const std::string &msg) {
    if (ctx) ctx->err = msg;
    else g_global_err = msg;
    return st;
}

qz_status cuda_fail(qz_ctx *ctx, cudaError_t e, const char *what) {
    std::string m = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    return fail(ctx, QZ_ENCODING, m);
}

static qz_status grow_dev(qz_ctx *ctx, uint8_t **p, size_t *have, size_t want) {
    if (*have >= want) return QZ_OK;
    size_t n = want < 4096 ? 4096 : want + want / 4;
    if (*p) {
        QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        QZ_CUDA(ctx, cudaFree(*p));
        *p = nullptr;
        *have = 0;
    }
    QZ_CUDA(ctx, cudaMalloc((void **)p, n));
    *have = n;
    return QZ_OK;
}

qz_status ensure_query(qz_ctx *ctx, size_t bytes) { return grow_dev(ctx, &ctx->d_query, &ctx->d_query_bytes, bytes); }
qz_status ensure_io(qz_ctx *ctx, size_t bytes) { return grow_dev(ctx, &ctx->d_io, &ctx->d_io_bytes, bytes); }

qz_status ensure_pin(qz_ctx *ctx, size_t bytes) {
    if (ctx->h_pin_bytes >= bytes) return QZ_OK;
    size_t n = bytes < 65536 ? 65536 : bytes + bytes / 4;
    if (ctx->h_pin) {
        QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        QZ_CUDA(ctx, cudaFreeHost(ctx->h_pin));
        ctx->h_pin = nullptr;
        ctx->h_pin_bytes = 0;
    }
    QZ_CUDA(ctx, cudaMallocHost((void **)&ctx->h_pin, n));
    ctx->h_pin_bytes = n;
    return QZ_OK;
}

qz_status ensure_smem_attr(qz_ctx *ctx, const void *fn, int bytes) {
    std::lock_guard<std::mutex> g(ctx->attr_mu);
    if (ctx->attr_done.count(fn)) return QZ_OK;
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    QZ_CUDA(ctx, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    ctx->attr_done.insert(fn);
    return QZ_OK;
}

static void timing_drain(qz_ctx *ctx, bool wait) {
    while (ctx->ev_tail < ctx->ev_head) {
        const int i = (int)(ctx->ev_tail % qz_ctx::kTimingRing);
        if (wait) cudaEventSynchronize(ctx->ev_stop[i]);
        else if (cudaEventQuery(ctx->ev_stop[i]) != cudaSuccess) break;
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, ctx->ev_start[i], ctx->ev_stop[i]) == cudaSuccess) {
            ctx->timed_ms += ms;
            ctx->timed_n++;
        }
        ctx->ev_tail++;
    }
}

void timing_begin(qz_ctx *ctx) {
    if (!ctx->timing) return;
    if (ctx->ev_head - ctx->ev_tail >= (uint64_t)qz_ctx::kTimingRing) timing_drain(ctx, true);
    const int i = (int)(ctx->ev_head % qz_ctx::kTimingRing);
    if (!ctx->ev_start[i]) {
        cudaEventCreate(&ctx->ev_start[i]);
        cudaEventCreate(&ctx->ev_stop[i]);
    }
    cudaEventRecord(ctx->ev_start[i], ctx->stream);
}

void timing_end(qz_ctx *ctx) {
    if (!ctx->timing) return;
    const int i = (int)(ctx->ev_head % qz_ctx::kTimingRing);
    cudaEventRecord(ctx->ev_stop[i], ctx->stream);
    ctx->ev_head++;
}

static void ctx_free(qz_ctx *ctx);
void ctx_release(qz_ctx *ctx) {
    if (ctx && ctx->refs.fetch_sub(1) == 1) ctx_free(ctx);
}
}  // namespace qz

using namespace qz;

extern "C" {

qz_status qz_ctx_create(int32_t device, void *stream, qz_ctx **out) {
    if (!out) return fail(nullptr, QZ_ARGUMENTS, "qz_ctx_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(nullptr, QZ_ENCODING,
                    std::string("qz_ctx_create: no CUDA device (") + cudaGetErrorString(e) +
                        "); libqz_cuda has no CPU fallback");
    if (device < 0 || device >= n) return fail(nullptr, QZ_ARGUMENTS, "qz_ctx_create: bad device ordinal");
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDeviceProperties");
    if (prop.major != 10)
        return fail(nullptr, QZ_ENCODING,
                    "qz_ctx_create: device is sm_" + std::to_string(prop.major * 10 + prop.minor) +
                        ", this library contains sm_100a code only");
    qz_ctx *ctx = new qz_ctx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    e = cudaSetDevice(device);
    if (e == cudaSuccess) {
        if (stream) {
            ctx->stream = (cudaStream_t)stream;
        } else {
            e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
            ctx->own_stream = true;
        }
    }
    const size_t lists = ctx->sm_count > 256 ? ctx->sm_count : 256;
    if (e == cudaSuccess) e = cudaMalloc((void **)&ctx->d_partial, lists * QZ_MAX_K * sizeof(uint64_t));
    if (e == cudaSuccess) e = cudaMalloc((void **)&ctx->d_merge_tmp, (size_t)256 * QZ_MAX_K * sizeof(uint64_t));
    if (e == cudaSuccess) e = cudaMalloc((void **)&ctx->d_done, 256);
    if (e == cudaSuccess) e = cudaMemset(ctx->d_done, 0, 256);
    if (e == cudaSuccess) e = cudaMalloc((void **)&ctx->d_keys, QZ_MAX_K * sizeof(uint64_t));
    // one allocation, ids right behind the scores: a host-facing search reads both back with ONE copy
    if (e == cudaSuccess) e = cudaMalloc((void **)&ctx->d_scores_k, 2 * QZ_MAX_K * sizeof(float));
    if (e == cudaSuccess) ctx->d_ids_k = reinterpret_cast<uint32_t *>(ctx->d_scores_k + QZ_MAX_K);
    if (e != cudaSuccess) {
        qz_status st = cuda_fail(nullptr, e, "qz_ctx_create");
        qz_ctx_destroy(ctx);
        return st;
    }
    *out = ctx;
    return QZ_OK;
}

void qz_ctx_destroy(qz_ctx *ctx) {
    if (!ctx) return;
    qz::ctx_release(ctx);
}

}  // extern "C"

namespace qz {
static void ctx_free(qz_ctx *ctx) {
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    cudaFree(ctx->d_partial);
    cudaFree(ctx->d_merge_tmp);
    cudaFree(ctx->d_done);
    cudaFree(ctx->d_keys);
    cudaFree(ctx->d_scores_k);  // d_ids_k lives in the same allocation
    cudaFree(ctx->d_query);
    cudaFree(ctx->d_io);
    if (ctx->h_pin) cudaFreeHost(ctx->h_pin);
    for (int i = 0; i < qz_ctx::kTimingRing; i++) {
        if (ctx->ev_start[i]) cudaEventDestroy(ctx->ev_start[i]);
        if (ctx->ev_stop[i]) cudaEventDestroy(ctx->ev_stop[i]);
    }
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}
}  // namespace qz

extern "C" {

const char *qz_last_error(const qz_ctx *ctx) { return ctx ? ctx->err.c_str() : g_global_err.c_str(); }
const char *qz_last_global_error(void) { return g_global_err.c_str(); }

qz_status qz_ctx_synchronize(qz_ctx *ctx) {
    if (!ctx) return QZ_ARGUMENTS;
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return QZ_OK;
}

qz_status qz_ctx_set_timing(qz_ctx *ctx, int32_t enabled) {
    if (!ctx) return QZ_ARGUMENTS;
    timing_drain(ctx, true);
    ctx->timing = enabled != 0;
    ctx->timed_ms = 0.0;
    ctx->timed_n = 0;
    return QZ_OK;
}

qz_status qz_ctx_set_scan_chaining(qz_ctx *ctx, int32_t enabled) {
    if (!ctx) return QZ_ARGUMENTS;
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->chain_ok = enabled != 0;
    ctx->chain_tail = 0;
    return QZ_OK;
}

qz_status qz_ctx_set_phase_stamps(qz_ctx *ctx, uint64_t *d_stamps) {
    if (!ctx) return QZ_ARGUMENTS;
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->d_stamps = d_stamps;
    return QZ_OK;
}

qz_status qz_ctx_get_timing(qz_ctx *ctx, double *total_ms, uint64_t *launches) {
    if (!ctx) return QZ_ARGUMENTS;
    timing_drain(ctx, true);
    if (total_ms) *total_ms = ctx->timed_ms;
    if (launches) *launches = ctx->timed_n;
    return QZ_OK;
}

void *qz_ctx_stream(const qz_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
uint64_t qz_ctx_launch_count(const qz_ctx *ctx) { return ctx ? ctx->launches : 0; }
int32_t qz_ctx_sm_count(const qz_ctx *ctx) { return ctx ? ctx->sm_count : 0; }

}  // extern "C"
