// row_scan_dot.cu — instantiations of row_scan_kernel<kOpDot, *, *>
#include "row_scan_inst.cuh"

namespace qz {
cudaError_t launch_row_scan_dot(uint32_t wpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem,
                               cudaStream_t s) {
    return launch_wpl<kOpDot>(wpl, full, p, grid, smem, s);
}
}  // namespace qz
