// row_scan_dot.cu — instantiations of row_scan_kernel<kOpDot, *, *>
#include "row_scan_inst.cuh"

namespace qz {
cudaError_t launch_row_scan_dot(qz_ctx *ctx, uint32_t wpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem) {
    return launch_wpl<kOpDot>(ctx, wpl, full, p, grid, smem);
}
}  // namespace qz
