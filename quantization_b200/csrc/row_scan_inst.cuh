// row_scan_inst.cuh — template dispatch of row_scan_kernel for one OP (included by
// row_scan_{dot,l1,xor}.cu so the three families compile in parallel).
#pragma once
#include "row_scan.cuh"

namespace qz {

template <int OP, int WPL, bool FULL>
static cudaError_t launch_one(const RowScanParams &p, uint32_t grid, size_t smem, cudaStream_t stream) {
    static bool attr_done = false;  // per instantiation
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(row_scan_kernel<OP, WPL, FULL>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)227 * 1024);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    row_scan_kernel<OP, WPL, FULL><<<grid, kScanThreads, smem, stream>>>(p);
    return cudaGetLastError();
}

// wpl must come from row_scan_wpl(); `full` = code_words == 32 * wpl
template <int OP>
static cudaError_t launch_wpl(uint32_t wpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem,
                              cudaStream_t s) {
#define QZ_WPL_CASE(W)                                             \
    case W:                                                        \
        return full ? launch_one<OP, W, true>(p, grid, smem, s)    \
                    : launch_one<OP, W, false>(p, grid, smem, s);
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
            return full ? launch_one<OP, 32, true>(p, grid, smem, s) : launch_one<OP, 32, false>(p, grid, smem, s);
    }
#undef QZ_WPL_CASE
}

}  // namespace qz
