/* abi_smoke.c — include/qz_cuda.h consumed by a plain C99 compiler and linked against libqz_cuda.so, the way a C
 * (or cgo / Rust-bindgen) host would: proves the header is C (no C++ leaks), that every prototype it declares has a
 * definition in the library (the table below takes the address of each one), and that the pure-host entry points
 * answer.  Built and run by tests/test_abi.py::test_header_compiles_as_c99_and_links:
 *     gcc -std=c99 -Wall -Wextra -Werror -pedantic tests/abi_smoke.c -Iinclude -Lquantization_b200 -lqz_cuda
 * Without a GPU qz_ctx_create must fail with a message — never fall back. */
#include <stdio.h>
#include <string.h>

#include "qz_cuda.h"

typedef void (*fn)(void);
#define F(x) {#x, (fn)x}
static const struct {
    const char *name;
    fn f;
} table[] = {
    F(qz_ctx_create), F(qz_ctx_destroy), F(qz_last_error), F(qz_last_global_error), F(qz_ctx_synchronize),
    F(qz_ctx_stream), F(qz_ctx_launch_count), F(qz_ctx_sm_count), F(qz_ctx_set_timing), F(qz_ctx_get_timing), F(qz_ctx_set_phase_stamps), F(qz_ctx_set_scan_chaining),
    F(qz_u8_actual_dim), F(qz_u8_row_bytes), F(qz_u8_index_create), F(qz_u8_index_set_params), F(qz_u8_encode),
    F(qz_u8_encode_dev), F(qz_u8_minmax_dev), F(qz_u8_quantile_dev), F(qz_u8_encode_range_dev), F(qz_u8_search_batch),
    F(qz_u8_search_batch_dev), F(qz_merge_topk_batch_dev), F(qz_bin_row_bytes), F(qz_bin_index_create), F(qz_bin_encode),
    F(qz_bin_encode_dev), F(qz_pq_index_create), F(qz_pq_train), F(qz_pq_train_dev), F(qz_pq_assign), F(qz_pq_assign_dev),
    F(qz_index_destroy), F(qz_index_codec), F(qz_index_count), F(qz_index_row_bytes), F(qz_query_bytes),
    F(qz_index_storage_dev), F(qz_index_download), F(qz_encode_query), F(qz_encode_query_dev), F(qz_scan_topk),
    F(qz_scan_topk_dev), F(qz_search_topk), F(qz_scan_scores), F(qz_scan_scores_dev), F(qz_score_points), F(qz_score_pairs),
    F(qz_score_points_dev), F(qz_score_pairs_dev), F(qz_merge_topk_dev), F(qz_merge_topk_static_dev), F(qz_host_pin), F(qz_host_unpin),
    F(qz_scan_topk_chunked), F(qz_peer_create), F(qz_peer_connect), F(qz_peer_destroy), F(qz_scan_topk_peer_dev),
    F(qz_search_topk_peer), F(qz_scan_topk_gather_dev), F(qz_search_topk_gather),
};

int main(void) {
    size_t i, n = sizeof table / sizeof table[0];
    qz_ctx *ctx = NULL;
    qz_status st;
    for (i = 0; i < n; i++)
        if (!table[i].f) {
            printf("missing %s\n", table[i].name);
            return 1;
        }
    /* get_actual_dim / get_quantized_vector_size (encoded_vectors_u8.rs:252-259), get_storage_size (binary.rs:99-159) */
    if (qz_u8_actual_dim(65) != 80 || qz_u8_row_bytes(1536) != 1540 || qz_bin_row_bytes(1536, 16) != 192 ||
        qz_bin_row_bytes(33, 1) != 8) {
        printf("pure-host entry points disagree with the reference's size formulae\n");
        return 1;
    }
    if (QZ_OK != 0 || QZ_ARGUMENTS != 1 || QZ_ENCODING != 2 || QZ_STOPPED != 3 || QZ_IO != 4) return 1; /* lib.rs:18-24 */
    st = qz_ctx_create(0, NULL, &ctx);
    if (st == QZ_OK) {
        printf("abi_smoke: %u symbols, context on device 0 with %d SMs\n", (unsigned)n, (int)qz_ctx_sm_count(ctx));
        qz_ctx_destroy(ctx);
    } else {
        const char *msg = qz_last_global_error();
        if (ctx != NULL || msg == NULL || strlen(msg) == 0) {
            printf("qz_ctx_create failed without a message\n");
            return 1;
        }
        printf("abi_smoke: %u symbols, no device: status %d, \"%s\"\n", (unsigned)n, (int)st, msg);
    }
    return 0;
}
