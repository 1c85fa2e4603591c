// row_scan_l1.cu — instantiations of row_scan_kernel<kOpL1, *, *>
#include "row_scan_inst.cuh"

namespace qz {
cudaError_t launch_row_scan_l1(uint32_t wpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem,
                               cudaStream_t s) {
    return launch_wpl<kOpL1>(wpl, full, p, grid, smem, s);
}
}  // namespace qz
