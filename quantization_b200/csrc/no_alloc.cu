// no_alloc.cu — qz_merge_topk_static_dev: the multi-level top-k merge of qz_merge_topk_dev with its intermediate
// lists in a file-scope __device__ array instead of cudaMalloc'ed context scratch.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The entry point calls
// no cudaMalloc / cudaFree / allocating library routine; the array below is part of the loaded module, so every GPU
// that loads libqz_cuda.so holds its 2 MiB until the process exits.
//
// The array is shared by every call on a device and keeps the last call's contents.  Nothing in it has to start at
// zero: each merge level writes all Kp keys of every list it produces before the next level (a later launch on the
// same stream) reads exactly those lists, so stale keys of an earlier call are never read.  Calls that overlap on
// the GPU (two streams / two contexts on one device) race on it; calls on one stream are ordered and safe.
#include <mutex>

#include "qz_internal.h"
#include "topk.cuh"

namespace qz {

// two ping-pong halves of QZ_MERGE_STATIC_MAX_LISTS / (32768 / QZ_MAX_K) = 128 lists of QZ_MAX_K keys: the largest
// first-level output (4096 lists of 1024 keys reduce 32 at a time)
constexpr size_t kStaticHalfKeys = (size_t)128 * QZ_MAX_K;
static __device__ __align__(256) unsigned long long g_merge_scratch[2 * kStaticHalfKeys];

// The name g_merge_scratch is a host-side stand-in in host code, not a device address: the address the kernel gets
// comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kMaxDevices = 64;
static std::mutex g_setup_mu;
static uint64_t *g_scratch_ptr[kMaxDevices];  // zero-initialised: not set up yet

static qz_status static_setup(qz_ctx *ctx, uint64_t **out) {
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return fail(ctx, QZ_ARGUMENTS, "merge_topk_static: device ordinal out of range");
    std::lock_guard<std::mutex> g(g_setup_mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!g_scratch_ptr[ctx->device]) {
        void *ptr = nullptr;
        QZ_CUDA(ctx, cudaGetSymbolAddress(&ptr, g_merge_scratch));
        if (!ptr) return fail(ctx, QZ_ENCODING, "merge_topk_static: cudaGetSymbolAddress returned NULL");
        // this file's own instance of the merge kernel (<1>); the attribute is per device as well
        QZ_CUDA(ctx, cudaFuncSetAttribute((const void *)merge_lists_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          128 * 1024));
        g_scratch_ptr[ctx->device] = reinterpret_cast<uint64_t *>(ptr);
    }
    *out = g_scratch_ptr[ctx->device];
    return QZ_OK;
}

}  // namespace qz

using namespace qz;

extern "C" qz_status qz_merge_topk_static_dev(qz_ctx *ctx, const uint64_t *d_keys_in, uint32_t n_lists, uint32_t k,
                                              int32_t largest, uint64_t *d_keys, float *d_scores, uint32_t *d_ids) {
    if (!ctx) return QZ_ARGUMENTS;
    if (k == 0 || k > QZ_MAX_K) return fail(ctx, QZ_ARGUMENTS, "merge_topk_static: k must be in [1, QZ_MAX_K]");
    if (n_lists > QZ_MERGE_STATIC_MAX_LISTS)
        return fail(ctx, QZ_ARGUMENTS, "merge_topk_static: n_lists = " + std::to_string(n_lists) + " exceeds QZ_MERGE_STATIC_MAX_LISTS (" +
                                       std::to_string(QZ_MERGE_STATIC_MAX_LISTS) + "), the capacity of the built-in scratch array");
    if (n_lists && !d_keys_in) return fail(ctx, QZ_ARGUMENTS, "merge_topk_static: NULL input");
    uint64_t *scratch = nullptr;
    QZ_TRY(static_setup(ctx, &scratch));
    const uint32_t Kp = topk_pad(k);
    const uint32_t max_lists = 32768 / Kp;  // (lists / 2) * Kp keys <= 16384 = 128 KB of shared memory
    const uint64_t *in = d_keys_in;
    uint32_t in_len = k, in_stride = k;
    uint64_t *bufs[2] = {scratch, scratch + kStaticHalfKeys};
    int which = 0;
    for (;;) {
        uint32_t lpc = pow2_ceil(n_lists < 2 ? 2 : n_lists);
        if (lpc > max_lists) lpc = max_lists;
        const uint32_t grid = n_lists ? (n_lists + lpc - 1) / lpc : 1;
        const bool final_level = grid == 1;
        if (!final_level && (size_t)grid * Kp > kStaticHalfKeys)
            return fail(ctx, QZ_ENCODING, "merge_topk_static: level output exceeds the scratch array");
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
        merge_lists_kernel<1><<<grid, 1024, (size_t)(lpc / 2) * Kp * 8, ctx->stream>>>(p);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail(ctx, e, "merge_topk_static: merge_lists_kernel launch");
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
