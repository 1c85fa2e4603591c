// tensor_map.h — host-side construction of TMA tensor maps (cuTensorMapEncodeTiled through the runtime's
// driver entry point lookup, so the library links cudart only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

namespace qz {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline PFN_encodeTiled get_encode_tiled() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(p);
    }
    return fn;
}

// 2-D u8 tensor [rows][pitch] (pitch contiguous bytes, a multiple of 16), box = box_bytes x box_rows,
// out-of-bounds reads give 0
inline bool make_tensor_map_u8_2d(CUtensorMap *map, const void *base, uint64_t rows, uint32_t pitch, uint32_t box_bytes,
                                  uint32_t box_rows, CUtensorMapSwizzle swizzle,
                                  CUtensorMapL2promotion promo = CU_TENSOR_MAP_L2_PROMOTION_L2_256B) {
    PFN_encodeTiled enc = get_encode_tiled();
    if (!enc) return false;
    cuuint64_t dims[2] = {pitch, rows};
    cuuint64_t strides[1] = {pitch};
    cuuint32_t box[2] = {box_bytes, box_rows};
    cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, promo,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace qz
