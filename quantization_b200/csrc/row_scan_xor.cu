// row_scan_xor.cu — instantiations of row_scan_kernel<kOpXor, *, *>
#include "row_scan_inst.cuh"

namespace qz {
cudaError_t launch_row_scan_xor(qz_ctx *ctx, uint32_t wpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem) {
    return launch_wpl<kOpXor>(ctx, wpl, full, p, grid, smem);
}
// vector mappings: lpr = lanes per row (8 or 4), vpl = 16-byte vectors per lane (1..4), full = no lane idles
template <int OP>
static cudaError_t launch_vpl(qz_ctx *s, uint32_t vpl, bool full, const RowScanParams &p, uint32_t grid, size_t smem) {
    switch (vpl) {
        case 1: return full ? launch_one<OP, 1, true>(s, p, grid, smem) : launch_one<OP, 1, false>(s, p, grid, smem);
        case 2: return full ? launch_one<OP, 2, true>(s, p, grid, smem) : launch_one<OP, 2, false>(s, p, grid, smem);
        case 3: return full ? launch_one<OP, 3, true>(s, p, grid, smem) : launch_one<OP, 3, false>(s, p, grid, smem);
        default: return full ? launch_one<OP, 4, true>(s, p, grid, smem) : launch_one<OP, 4, false>(s, p, grid, smem);
    }
}
cudaError_t launch_row_scan_xorv(qz_ctx *ctx, uint32_t lpr, uint32_t vpl, bool full, const RowScanParams &p, uint32_t grid,
                                size_t smem) {
    return lpr == 4 ? launch_vpl<kOpXorV4>(ctx, vpl, full, p, grid, smem) : launch_vpl<kOpXorV>(ctx, vpl, full, p, grid, smem);
}
}  // namespace qz
