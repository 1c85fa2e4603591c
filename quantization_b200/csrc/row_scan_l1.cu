// row_scan_l1.cu — instantiations of row_scan_kernel<kOpL1, *, *>
#include "row_scan_inst.cuh"

namespace qz {
cudaError_t launch_row_scan_l1(qz_ctx *ctx, uint32_t wpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem) {
    return launch_wpl<kOpL1>(ctx, wpl, full, p, grid, smem);
}
}  // namespace qz
