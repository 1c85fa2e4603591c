// row_scan_xor.cu — instantiations of row_scan_kernel<kOpXor, *, *>
#include "row_scan_inst.cuh"

namespace qz {
cudaError_t launch_row_scan_xor(uint32_t wpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem,
                               cudaStream_t s) {
    return launch_wpl<kOpXor>(wpl, full, p, grid, smem, s);
}
// vector mappings: lpr = lanes per row (8 or 4), vpl = 16-byte vectors per lane (1..4), full = no lane idles
template <int OP>
static cudaError_t launch_vpl(uint32_t vpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem, cudaStream_t s) {
    switch (vpl) {
        case 1: return full ? launch_one<OP, 1, true>(p, grid, smem, s) : launch_one<OP, 1, false>(p, grid, smem, s);
        case 2: return full ? launch_one<OP, 2, true>(p, grid, smem, s) : launch_one<OP, 2, false>(p, grid, smem, s);
        case 3: return full ? launch_one<OP, 3, true>(p, grid, smem, s) : launch_one<OP, 3, false>(p, grid, smem, s);
        default: return full ? launch_one<OP, 4, true>(p, grid, smem, s) : launch_one<OP, 4, false>(p, grid, smem, s);
    }
}
cudaError_t launch_row_scan_xorv(uint32_t lpr, uint32_t vpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem,
                                cudaStream_t s) {
    return lpr == 4 ? launch_vpl<kOpXorV4>(vpl, full, p, grid, smem, s) : launch_vpl<kOpXorV>(vpl, full, p, grid, smem, s);
}
}  // namespace qz
