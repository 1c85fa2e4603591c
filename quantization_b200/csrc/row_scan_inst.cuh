// row_scan_inst.cuh — template dispatch of row_scan_kernel for one OP (included by
// row_scan_{dot,l1,xor}.cu so the three families compile in parallel).
#pragma once
#include "qz_internal.h"
#include "row_scan.cuh"

namespace qz {

template <int OP, int WPL, bool FULL>
static cudaError_t launch_one(qz_ctx *ctx, const RowScanParams &p, uint32_t grid, size_t smem) {
    if (ensure_smem_attr(ctx, (const void *)row_scan_kernel<OP, WPL, FULL>, 227 * 1024) != QZ_OK)
        return cudaErrorInvalidValue;  // (text already in ctx->err)
    if (!ctx->pdl_next) {
        row_scan_kernel<OP, WPL, FULL><<<grid, kScanThreads, smem, ctx->stream>>>(p);
        return cudaGetLastError();
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kScanThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, row_scan_kernel<OP, WPL, FULL>, p);
    return e != cudaSuccess ? e : cudaGetLastError();
}

// wpl must come from row_scan_wpl(); `full` = code_words == 32 * wpl
template <int OP>
static cudaError_t launch_wpl(qz_ctx *s, uint32_t wpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem) {
#define QZ_WPL_CASE(W)                                             \
    case W:                                                        \
        return full ? launch_one<OP, W, true>(s, p, grid, smem)    \
                    : launch_one<OP, W, false>(s, p, grid, smem);
    switch (wpl) {
        QZ_WPL_CASE(1)
        QZ_WPL_CASE(2)
        QZ_WPL_CASE(3)
        QZ_WPL_CASE(4)
        QZ_WPL_CASE(6)
        QZ_WPL_CASE(8)
        QZ_WPL_CASE(12)
        QZ_WPL_CASE(16)
        QZ_WPL_CASE(24)
        default:
            return full ? launch_one<OP, 32, true>(s, p, grid, smem) : launch_one<OP, 32, false>(s, p, grid, smem);
    }
#undef QZ_WPL_CASE
}

}  // namespace qz
