"""ctypes binding of libqz_cuda.so (include/qz_cuda.h).  No fallback: if the library is
missing or no B200 is present, using the package raises."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libqz_cuda.so")

QZ_OK, QZ_ARGUMENTS, QZ_ENCODING, QZ_STOPPED, QZ_IO = 0, 1, 2, 3, 4
QZ_MAX_K = 1024

vp = C.c_void_p
u8p = C.POINTER(C.c_uint8)
f32p = C.POINTER(C.c_float)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
i32p = C.POINTER(C.c_int32)

_SIGS = {
    "qz_ctx_create": (C.c_int32, [C.c_int32, vp, C.POINTER(vp)]),
    "qz_ctx_destroy": (None, [vp]),
    "qz_last_error": (C.c_char_p, [vp]),
    "qz_last_global_error": (C.c_char_p, []),
    "qz_ctx_synchronize": (C.c_int32, [vp]),
    "qz_ctx_stream": (vp, [vp]),
    "qz_ctx_launch_count": (C.c_uint64, [vp]),
    "qz_ctx_sm_count": (C.c_int32, [vp]),
    "qz_ctx_set_timing": (C.c_int32, [vp, C.c_int32]),
    "qz_ctx_get_timing": (C.c_int32, [vp, C.POINTER(C.c_double), u64p]),
    "qz_ctx_set_phase_stamps": (C.c_int32, [vp, vp]),
    "qz_ctx_set_scan_chaining": (C.c_int32, [vp, C.c_int32]),
    "qz_u8_actual_dim": (C.c_uint32, [C.c_uint32]),
    "qz_u8_row_bytes": (C.c_uint32, [C.c_uint32]),
    "qz_u8_index_create": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_int32, C.c_int32, C.c_float, C.c_float,
                                       C.c_float, C.c_uint32, C.POINTER(vp)]),
    "qz_u8_index_set_params": (C.c_int32, [vp, C.c_float, C.c_float, C.c_float]),
    "qz_u8_encode": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_int32, C.c_int32, f32p, vp, C.c_uint64,
                                 C.c_uint64, vp, vp, f32p, f32p, f32p]),
    "qz_u8_encode_dev": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_int32, C.c_int32, f32p, vp, C.c_uint64,
                                     C.c_uint64, vp, vp, f32p, f32p, f32p]),
    "qz_u8_minmax_dev": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, f32p, f32p]),
    "qz_u8_quantile_dev": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_uint64, C.c_float, C.POINTER(C.c_int32), f32p, f32p]),
    "qz_u8_encode_range_dev": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_int32, C.c_int32, C.c_float, C.c_float, vp, vp,
                                           f32p, f32p, f32p]),
    "qz_u8_search_batch": (C.c_int32, [vp, vp, C.c_uint32, C.c_uint32, C.c_int32, vp, vp]),
    "qz_u8_search_batch_dev": (C.c_int32, [vp, vp, C.c_uint32, C.c_uint32, C.c_int32, vp, vp]),
    "qz_search_topk": (C.c_int32, [vp, vp, C.c_uint32, C.c_int32, vp, vp, u32p]),
    "qz_merge_topk_batch_dev": (C.c_int32, [vp, vp, vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int32, vp, vp]),
    "qz_bin_row_bytes": (C.c_uint32, [C.c_uint32, C.c_int32]),
    "qz_bin_index_create": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_int32, C.c_int32, C.c_int32, C.c_uint32,
                                        C.POINTER(vp)]),
    "qz_bin_encode": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_int32, vp, vp]),
    "qz_bin_encode_dev": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_int32, vp, vp]),
    "qz_pq_index_create": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_uint32, vp, C.c_int32, C.c_int32,
                                       C.c_uint32, C.POINTER(vp)]),
    "qz_pq_train": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_uint32, vp, C.c_uint64, C.c_uint64, C.c_uint32,
                                C.c_float, vp, vp, vp]),
    "qz_pq_train_dev": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_uint32, vp, C.c_uint64, C.c_uint64,
                                    C.c_uint32, C.c_float, vp, vp, vp]),
    "qz_pq_assign": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_uint32, vp, vp, vp]),
    "qz_pq_assign_dev": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint32, C.c_uint32, vp, vp, vp]),
    "qz_index_destroy": (None, [vp]),
    "qz_index_codec": (C.c_int32, [vp]),
    "qz_index_count": (C.c_uint64, [vp]),
    "qz_index_row_bytes": (C.c_uint32, [vp]),
    "qz_query_bytes": (C.c_uint32, [vp]),
    "qz_index_storage_dev": (vp, [vp]),
    "qz_index_download": (C.c_int32, [vp, vp]),
    "qz_encode_query": (C.c_int32, [vp, vp, vp]),
    "qz_encode_query_dev": (C.c_int32, [vp, vp, C.c_uint32, vp]),
    "qz_scan_topk": (C.c_int32, [vp, vp, C.c_uint32, C.c_int32, vp, vp, u32p]),
    "qz_scan_topk_dev": (C.c_int32, [vp, vp, C.c_uint32, C.c_int32, vp, vp, vp]),
    "qz_scan_scores": (C.c_int32, [vp, vp, vp]),
    "qz_scan_scores_dev": (C.c_int32, [vp, vp, vp]),
    "qz_score_points": (C.c_int32, [vp, vp, vp, C.c_uint64, vp]),
    "qz_score_pairs": (C.c_int32, [vp, vp, vp, C.c_uint64, vp]),
    "qz_score_points_dev": (C.c_int32, [vp, vp, vp, C.c_uint64, vp]),
    "qz_score_pairs_dev": (C.c_int32, [vp, vp, vp, C.c_uint64, vp]),
    "qz_merge_topk_dev": (C.c_int32, [vp, vp, C.c_uint32, C.c_uint32, C.c_int32, vp, vp, vp]),
    "qz_merge_topk_static_dev": (C.c_int32, [vp, vp, C.c_uint32, C.c_uint32, C.c_int32, vp, vp, vp]),
    "qz_host_pin": (C.c_int32, [vp, vp, C.c_uint64]),
    "qz_host_unpin": (C.c_int32, [vp, vp]),
    "qz_scan_topk_chunked": (C.c_int32, [vp, vp, vp, C.c_uint64, vp, C.c_uint32, C.c_int32, vp, vp, u32p]),
    "qz_peer_create": (C.c_int32, [vp, C.c_uint32, C.c_uint32, vp, C.POINTER(vp)]),
    "qz_peer_connect": (C.c_int32, [vp, vp]),
    "qz_peer_destroy": (None, [vp]),
    "qz_scan_topk_peer_dev": (C.c_int32, [vp, vp, vp, C.c_uint32, C.c_int32, vp, vp, vp]),
    "qz_search_topk_peer": (C.c_int32, [vp, vp, vp, C.c_uint32, C.c_int32, vp, vp, u32p]),
    "qz_scan_topk_gather_dev": (C.c_int32, [vp, vp, C.c_uint32, vp, C.c_uint32, C.c_int32, vp, vp, vp]),
    "qz_search_topk_gather": (C.c_int32, [vp, vp, C.c_uint32, vp, C.c_uint32, C.c_int32, vp, vp, u32p]),
}

_lib = None


def declared_symbols():
    return sorted(_SIGS)


def lib():
    """Load libqz_cuda.so (must have been built in-tree: python -m quantization_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m quantization_b200.build` "
                "(quantization_b200 has no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib
