/*
 * qz_cuda.h — C-ABI of libqz_cuda.so: the sm_100a replacement for qdrant/quantization's
 * native scorers (quantization/cpp/{sse,avx2,neon}.c, built by quantization/build.rs:18-40 and
 * bound in encoded_vectors_u8.rs:476-489 / encoded_vectors_binary.rs:317-332) plus the scan
 * loop, top-k, and batch-encode steps that surround them.
 *
 * Plain C: opaque handles, pointers and sizes only.  Every function returns a qz_status
 * (0 = OK) unless stated; nothing throws or aborts across this boundary.  Status codes map
 * 1:1 onto the reference's EncodingError (quantization/src/lib.rs:18-24).
 *
 * Memory contract: `const T* host` arguments are borrowed for the duration of the call.
 * `*_dev` entry points take DEVICE pointers and are asynchronous on the context's stream
 * (they exist so that callers with device-resident data, e.g. the benchmark's timed region,
 * pay no copies).  An index owns a private device copy of the encoded storage; the caller's
 * EncodedStorage keeps owning the host bytes (get_vector_data / save keep working).
 *
 * Encoded queries are byte blobs of qz_query_bytes(index) bytes:
 *   u8     : [f32 query offset][actual_dim codes]      (EncodedQueryU8, u8.rs:19-22)
 *   binary : row_bytes packed sign bits                (EncodedBinVector, binary.rs:17-19)
 *   pq     : n_chunks*256 f32 lookup table             (EncodedQueryPQ, pq.rs:35-37)
 *
 * Ordering contract of every top-k in this library: best score first; "best" = largest when
 * `largest != 0`, else smallest; ties go to the LOWER point id.  Scores compare by their IEEE
 * bit pattern mapped monotonically (so -0.0 < +0.0).
 */
#ifndef QZ_CUDA_H
#define QZ_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_MSC_VER)
#define QZ_API __declspec(dllexport)
#else
#define QZ_API __attribute__((visibility("default")))
#endif

typedef int32_t qz_status;
enum {
    QZ_OK = 0,
    QZ_ARGUMENTS = 1, /* EncodingError::ArgumentsError */
    QZ_ENCODING = 2,  /* EncodingError::EncodingError — CUDA / NCCL failure, text in qz_last_error */
    QZ_STOPPED = 3,   /* EncodingError::Stopped */
    QZ_IO = 4         /* EncodingError::IOError */
};

/* DistanceType, quantization/src/encoded_vectors.rs:6-11 */
enum { QZ_DOT = 0, QZ_L1 = 1, QZ_L2 = 2 };

/* codec of an index */
enum { QZ_CODEC_U8 = 0, QZ_CODEC_BIN = 1, QZ_CODEC_PQ = 2 };

typedef struct qz_ctx qz_ctx;     /* one CUDA device + stream + scratch */
typedef struct qz_index qz_index; /* one shard of an encoded corpus resident in HBM */

/* ---- context ------------------------------------------------------------------------------ */

/* `stream` = an existing cudaStream_t to run on (e.g. the caller's), or NULL to create one. */
QZ_API qz_status qz_ctx_create(int32_t device, void *stream, qz_ctx **out);
QZ_API void qz_ctx_destroy(qz_ctx *ctx);
/* last error text of this context (never NULL); thread-compatible, not thread-safe */
QZ_API const char *qz_last_error(const qz_ctx *ctx);
/* text of a failure that happened before a context existed (qz_ctx_create itself) */
QZ_API const char *qz_last_global_error(void);
QZ_API qz_status qz_ctx_synchronize(qz_ctx *ctx);
/* the cudaStream_t every `*_dev` entry point of this context enqueues on (the caller's, or the one created by
 * qz_ctx_create): a caller that mixes its own streams with `*_dev` calls orders them against this one with events */
QZ_API void *qz_ctx_stream(const qz_ctx *ctx);
/* number of kernels of this library launched through `ctx` so far (bench's gpu_launches) */
QZ_API uint64_t qz_ctx_launch_count(const qz_ctx *ctx);
QZ_API int32_t qz_ctx_sm_count(const qz_ctx *ctx);
/* Optional: bracket the dominant scan kernel of every scan call with CUDA events on the context
 * stream.  qz_ctx_get_timing synchronises and returns the summed kernel time and launch count
 * since qz_ctx_set_timing (the benchmark's roofline measurement). */
QZ_API qz_status qz_ctx_set_timing(qz_ctx *ctx, int32_t enabled);
QZ_API qz_status qz_ctx_get_timing(qz_ctx *ctx, double *total_ms, uint64_t *launches);
/* Chained scans.  Consecutive qz_scan_topk_dev / qz_scan_topk_peer_dev / qz_scan_topk_gather_dev calls (u8 and binary,
 * k <= 128) on a context are launched so that the streaming of query i+1 overlaps the final merge of query i
 * (programmatic dependent launch; results and their order on the stream are unchanged).  That is only safe when no
 * other work can sit between two scans on the stream, so it is automatic for a context that owns its stream
 * (qz_ctx_create with stream = NULL) and needs this promise for a caller-provided stream: while it is on, anything
 * the caller enqueues on that stream between two scans that WRITES the encoded query or the index storage must be
 * followed by an event wait or another non-kernel operation (which ends a chain).  Default off. */
QZ_API qz_status qz_ctx_set_scan_chaining(qz_ctx *ctx, int32_t enabled);
/* Optional: per-CTA phase stamps of the u8 / binary scan kernel.  d_stamps = device buffer of
 * QZ_STAMP_WORDS u64 per SM (qz_ctx_sm_count), or NULL to switch the stamps off.  Every CTA of the next scans
 * writes, at [blockIdx][i]: 0 globaltimer (ns) at kernel entry, 1 clock64 at entry, 2 clock64 after the top-k
 * warm-up, 3 after its last tile, 4 after its list is published, 5 / 6 (last CTA only, else 0) after the final
 * merge / after the peer exchange, 7 globaltimer at exit.  Profiling aid (tools/phase_stamps.py); no reference
 * counterpart. */
#define QZ_STAMP_WORDS 8
QZ_API qz_status qz_ctx_set_phase_stamps(qz_ctx *ctx, uint64_t *d_stamps);

/* ---- u8 scalar quantizer (EncodedVectorsU8, encoded_vectors_u8.rs) ------------------------- */

/* get_actual_dim / get_quantized_vector_size, u8.rs:252-259 */
QZ_API uint32_t qz_u8_actual_dim(uint32_t dim);
QZ_API uint32_t qz_u8_row_bytes(uint32_t dim);

/* Make a device-resident index from encoded storage = count rows of
 * [f32 vector_offset][actual_dim codes] (u8.rs:78-117) and the Metadata scalars (u8.rs:24-31).
 * `id_base` is added to local row numbers to form point ids (shard of a larger corpus).
 * `storage` may be NULL: the index is allocated and zero-filled, to be written through
 * qz_index_storage_dev (e.g. by qz_u8_encode_dev). */
QZ_API qz_status qz_u8_index_create(qz_ctx *ctx, const uint8_t *storage, uint64_t count, uint32_t dim,
                                    int32_t distance, int32_t invert, float alpha, float offset, float multiplier,
                                    uint32_t id_base, qz_index **out);
/* update the Metadata scalars of an existing u8 index (after a device-side encode) */
QZ_API qz_status qz_u8_index_set_params(qz_index *index, float alpha, float offset, float multiplier);

/* EncodedVectorsU8::encode (u8.rs:34-140): min/max (quantile.rs:5-19), optional quantile cut
 * (quantile.rs:21-71), codes + per-vector header, and the three Metadata scalars.
 *   data        count x dim f32, row-major (the reference's orig_data iterator, materialised)
 *   quantile    NULL = None
 *   sample_ids  the vectors the quantile cut is computed on; NULL = all when count <= 100000,
 *               else 100000 distinct ids from an affine permutation keyed by `seed` (the
 *               reference draws them from a randomly keyed Permutor, quantile.rs:32-34)
 *   stop        optional flag polled between row chunks; non-zero => QZ_STOPPED (u8.rs:74-76)
 *   out_storage count * qz_u8_row_bytes(dim) bytes
 * The host variant streams `data` through the device in chunks; the _dev variant takes device
 * pointers for data and out (e.g. qz_index_storage_dev) and returns after synchronising. */
QZ_API qz_status qz_u8_encode(qz_ctx *ctx, const float *data, uint64_t count, uint32_t dim, int32_t distance,
                              int32_t invert, const float *quantile, const uint64_t *sample_ids, uint64_t n_sample,
                              uint64_t seed, const volatile int32_t *stop, uint8_t *out_storage, float *alpha,
                              float *offset, float *multiplier);
QZ_API qz_status qz_u8_encode_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, int32_t distance,
                                  int32_t invert, const float *quantile, const uint64_t *sample_ids,
                                  uint64_t n_sample, uint64_t seed, const volatile int32_t *stop, uint8_t *d_out,
                                  float *alpha, float *offset, float *multiplier);

/* Sharded encode: the rows of one corpus live on several GPUs, one process each.  encode's global
 * quantities are split out so the caller can combine them across ranks (an all-reduce of two floats; an
 * all-gather of the sampled rows when a quantile is asked for) and every rank then encodes its own rows with the
 * same (min, max) — the resulting shards are byte-identical to slices of a single-GPU qz_u8_encode.
 *   qz_u8_minmax_dev        find_min_max_from_iter (quantile.rs:5-19) over this rank's rows; count == 0 returns
 *                           (f32::MAX, f32::MIN), the fold's identity
 *   qz_u8_quantile_dev      find_quantile_interval's two order statistics (quantile.rs:53-70) over `n_rows` sampled
 *                           rows resident on the device; `slice_size` = the number of sampled vectors of the WHOLE
 *                           corpus (min(count, 100000), quantile.rs:30); *found = 0 keeps (min, max)
 *   qz_u8_encode_range_dev  the second half of encode (u8.rs:73-128) with (min, max) given */
QZ_API qz_status qz_u8_minmax_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, float *min_out,
                                  float *max_out);
QZ_API qz_status qz_u8_quantile_dev(qz_ctx *ctx, const float *d_sample, uint64_t n_rows, uint32_t dim,
                                    uint64_t slice_size, float quantile, int32_t *found, float *lo, float *hi);
QZ_API qz_status qz_u8_encode_range_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim,
                                        int32_t distance, int32_t invert, float min_value, float max_value,
                                        const volatile int32_t *stop, uint8_t *d_out, float *alpha, float *offset,
                                        float *multiplier);

/* Batched multi-query search (BASELINE config 5): nq f32 queries [nq][dim] against every row, the k best
 * per query.  One call = encode_query for each query (u8.rs:290-329), the nq x count x actual_dim integer
 * contraction on the tensor cores (tcgen05 kind::i8, the same sums impl_score_dot_* returns per pair,
 * cpp/avx2.c:31-69), the score map of u8.rs:347 and a per-query top-k with the library's ordering contract.
 * Dot and L2 only (L1 is not a contraction).  Outputs are [nq][k] row-major; slots past `count` hold
 * NaN / 0xFFFFFFFF.  Both variants return after the work has completed.  The first call builds a
 * device-side copy of the codes with a 16-byte-aligned pitch (count * actual_dim bytes). */
QZ_API qz_status qz_u8_search_batch(qz_index *index, const float *queries, uint32_t nq, uint32_t k, int32_t largest,
                                    float *scores, uint32_t *ids);
QZ_API qz_status qz_u8_search_batch_dev(qz_index *index, const float *d_queries, uint32_t nq, uint32_t k,
                                        int32_t largest, float *d_scores, uint32_t *d_ids);

/* Cross-shard merge for query batches: d_scores_in / d_ids_in = [n_lists][nq][k] results of
 * qz_u8_search_batch_dev on n_lists shards (e.g. after an all-gather); outputs [nq][k].  Device pointers,
 * asynchronous on the context stream.  n_lists * k <= 16384. */
QZ_API qz_status qz_merge_topk_batch_dev(qz_ctx *ctx, const float *d_scores_in, const uint32_t *d_ids_in,
                                         uint32_t n_lists, uint32_t nq, uint32_t k, int32_t largest, float *d_scores,
                                         uint32_t *d_ids);

/* ---- binary quantizer (EncodedVectorsBin, encoded_vectors_binary.rs) ------------------------- */

/* bytes per row for BitsStoreType u8 (store_bytes = 1, binary.rs:99-116) or u128 (16, :152-159) */
QZ_API uint32_t qz_bin_row_bytes(uint32_t dim, int32_t store_bytes);
/* storage = count rows of packed sign bits, bit i at byte i/8 bit i%8 (binary.rs:193-208) */
QZ_API qz_status qz_bin_index_create(qz_ctx *ctx, const uint8_t *storage, uint64_t count, uint32_t dim,
                                     int32_t store_bytes, int32_t distance, int32_t invert, uint32_t id_base,
                                     qz_index **out);

/* EncodedVectorsBin::encode (binary.rs:165-208): bit = (v > 0.0) */
QZ_API qz_status qz_bin_encode(qz_ctx *ctx, const float *data, uint64_t count, uint32_t dim, int32_t store_bytes,
                               const volatile int32_t *stop, uint8_t *out_storage);
QZ_API qz_status qz_bin_encode_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim,
                                   int32_t store_bytes, const volatile int32_t *stop, uint8_t *d_out);

/* ---- product quantizer (EncodedVectorsPQ, encoded_vectors_pq.rs) ------------------------------ */

/* codes = count rows of n_chunks centroid ids (n_chunks = ceil(dim / chunk_size), pq.rs:109-121);
 * centroids = Metadata.centroids flattened [256][dim] f32 (pq.rs:39-44). */
QZ_API qz_status qz_pq_index_create(qz_ctx *ctx, const uint8_t *codes, uint64_t count, uint32_t dim,
                                    uint32_t chunk_size, const float *centroids, int32_t distance, int32_t invert,
                                    uint32_t id_base, qz_index **out);

/* find_centroids + kmeans (pq.rs:278-342, kmeans.rs:7-136): 256 centroids per chunk from a sample of
 * min(10000, count) vectors; all chunks are trained side by side on the device.
 *   sample_ids     NULL = all ids when count <= 10000, else an affine permutation keyed by `seed`
 *                  (the reference draws them from a randomly keyed Permutor, pq.rs:300-302)
 *   max_iterations KMEANS_MAX_ITERATIONS = 100, accuracy KMEANS_ACCURACY = 1e-5 (pq.rs:23-24)
 *   centroids      out, [256][dim] f32 (host) — Metadata.centroids flattened
 *   iterations     optional out, [n_chunks]: Lloyd iterations each chunk ran
 * f64 sums run in point order (= the reference with max_kmeans_threads = 1); an empty cluster is
 * reseeded from a pseudo-random sample point keyed by `seed` (thread_rng in the reference,
 * kmeans.rs:111-118).  count <= 256: the points themselves, zero padded (pq.rs:290-297). */
QZ_API qz_status qz_pq_train(qz_ctx *ctx, const float *data, uint64_t count, uint32_t dim, uint32_t chunk_size,
                             const uint64_t *sample_ids, uint64_t n_sample, uint64_t seed, uint32_t max_iterations,
                             float accuracy, const volatile int32_t *stop, float *centroids, uint32_t *iterations);
QZ_API qz_status qz_pq_train_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, uint32_t chunk_size,
                                 const uint64_t *sample_ids, uint64_t n_sample, uint64_t seed,
                                 uint32_t max_iterations, float accuracy, const volatile int32_t *stop,
                                 float *centroids, uint32_t *iterations);
/* encode_vector for every row (pq.rs:237-265): per chunk the nearest centroid, strict <, lowest
 * index wins.  codes = count * n_chunks bytes.  `centroids` is a host pointer in both variants. */
QZ_API qz_status qz_pq_assign(qz_ctx *ctx, const float *data, uint64_t count, uint32_t dim, uint32_t chunk_size,
                              const float *centroids, const volatile int32_t *stop, uint8_t *codes);
QZ_API qz_status qz_pq_assign_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim,
                                  uint32_t chunk_size, const float *centroids, const volatile int32_t *stop,
                                  uint8_t *d_codes);

/* ---- any index ------------------------------------------------------------------------------ */

QZ_API void qz_index_destroy(qz_index *index);
QZ_API int32_t qz_index_codec(const qz_index *index);
QZ_API uint64_t qz_index_count(const qz_index *index);
QZ_API uint32_t qz_index_row_bytes(const qz_index *index);
QZ_API uint32_t qz_query_bytes(const qz_index *index);
/* device pointer of the encoded storage (count * row_bytes bytes, 16-byte aligned) */
QZ_API void *qz_index_storage_dev(qz_index *index);
/* copy the encoded storage back to the host (count * row_bytes bytes) */
QZ_API qz_status qz_index_download(qz_index *index, uint8_t *storage);

/* encode_query (u8.rs:290-329 / binary.rs:288-291 / pq.rs:525-547): `query` = dim f32 on the
 * host, `out` = qz_query_bytes(index) bytes on the host.  Runs on the device. */
QZ_API qz_status qz_encode_query(qz_index *index, const float *query, void *out);
/* device variant: nq queries [nq][dim] f32 -> nq blobs, both device pointers, asynchronous */
QZ_API qz_status qz_encode_query_dev(qz_index *index, const float *d_queries, uint32_t nq, void *d_out);

/* The fused scan: score every point against one encoded query and keep the k best.
 * Replaces the caller-side `for i in 0..count { score_point(q, i) }` + BinaryHeap of
 * demos/src/ann_benchmark_data.rs:150-167.  Host buffers; writes min(k, count) results and
 * their number to *n_out.  k <= QZ_MAX_K. */
#define QZ_MAX_K 1024
QZ_API qz_status qz_scan_topk(qz_index *index, const void *enc_query, uint32_t k, int32_t largest, float *scores,
                              uint32_t *ids, uint32_t *n_out);
/* device variant, asynchronous on the context stream: d_enc_query device blob; results as
 * k sorted u64 keys (d_keys, may be NULL) and/or decoded d_scores/d_ids (may be NULL).
 * Unfilled slots (count < k) hold key 0 / score NaN / id 0xFFFFFFFF. */
QZ_API qz_status qz_scan_topk_dev(qz_index *index, const void *d_enc_query, uint32_t k, int32_t largest,
                                  uint64_t *d_keys, float *d_scores, uint32_t *d_ids);

/* encode_query + the fused scan in one call with one synchronisation: `query` = dim f32 on the host; results as
 * qz_scan_topk.  The end-to-end path of a caller that holds raw f32 queries. */
QZ_API qz_status qz_search_topk(qz_index *index, const float *query, uint32_t k, int32_t largest, float *scores,
                                uint32_t *ids, uint32_t *n_out);

/* score_point for every point: out[i] = score_point(q, i), i in [0,count). */
QZ_API qz_status qz_scan_scores(qz_index *index, const void *enc_query, float *scores);
QZ_API qz_status qz_scan_scores_dev(qz_index *index, const void *d_enc_query, float *d_scores);

/* score_point on an id list (local row numbers): backs EncodedVectors::score_point, n = 1 legal
 * (u8.rs:331-384, binary.rs:293-300, pq.rs:549-561) */
QZ_API qz_status qz_score_points(qz_index *index, const void *enc_query, const uint32_t *rows, uint64_t n,
                                 float *scores);
/* score_internal on id pairs: backs EncodedVectors::score_internal
 * (u8.rs:386-453, binary.rs:302-314, pq.rs:566-593) */
QZ_API qz_status qz_score_pairs(qz_index *index, const uint32_t *rows_a, const uint32_t *rows_b, uint64_t n,
                                float *scores);
/* device-pointer forms for callers that keep candidate lists on the device (graph search: the id-list scoring of
 * demos/benches/encode.rs:91-102, binary.rs:62-72, pq.rs:41-46): asynchronous on the context stream, ids are not
 * range-checked */
QZ_API qz_status qz_score_points_dev(qz_index *index, const void *d_enc_query, const uint32_t *d_rows, uint64_t n,
                                     float *d_scores);
QZ_API qz_status qz_score_pairs_dev(qz_index *index, const uint32_t *d_rows_a, const uint32_t *d_rows_b, uint64_t n,
                                    float *d_scores);

/* Merge `n_lists` best-first top-k lists (e.g. one per shard after an all-gather) into one.
 * Device pointers, asynchronous: d_keys_in = [n_lists][k] u64 keys as produced by
 * qz_scan_topk_dev; outputs as in qz_scan_topk_dev. */
QZ_API qz_status qz_merge_topk_dev(qz_ctx *ctx, const uint64_t *d_keys_in, uint32_t n_lists, uint32_t k,
                                   int32_t largest, uint64_t *d_keys, float *d_scores, uint32_t *d_ids);

/* qz_merge_topk_dev without any allocation behind it.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same results bit for bit; the intermediate lists of the multi-level
 * merge live in a 2 MiB file-scope __device__ array of the library (csrc/no_alloc.cu; held by every GPU that loads
 * the library until the process exits) instead of the context's cudaMalloc'ed scratch, and the call itself makes no
 * cudaMalloc / cudaFree / allocating library call.  n_lists > QZ_MERGE_STATIC_MAX_LISTS is refused with
 * QZ_ARGUMENTS (text in qz_last_error).  Every call on a device shares the array: calls on one stream are ordered
 * and safe, calls that overlap on the GPU (two streams or two contexts on one device) race on it. */
#define QZ_MERGE_STATIC_MAX_LISTS 4096
QZ_API qz_status qz_merge_topk_static_dev(qz_ctx *ctx, const uint64_t *d_keys_in, uint32_t n_lists, uint32_t k,
                                          int32_t largest, uint64_t *d_keys, float *d_scores, uint32_t *d_ids);

/* ---- out-of-core scan ---------------------------------------------------------------------------
 * The encoded storage stays in host memory (any EncodedStorage backend that can hand out contiguous runs of rows,
 * encoded_storage.rs:7-25; corpora larger than HBM) and is streamed through two device buffers: `buf_a` / `buf_b`
 * are two indexes of the same codec and parameters created with storage = NULL and count = the chunk capacity in
 * rows.  Copies (a second stream) overlap the scans; results are those of one resident index over all `count`
 * rows, ids global.  PCIe-bound; pin the storage (qz_host_pin) for full speed.  enc_query: host blob. */
QZ_API qz_status qz_host_pin(qz_ctx *ctx, const void *ptr, uint64_t bytes);
QZ_API qz_status qz_host_unpin(qz_ctx *ctx, const void *ptr);
QZ_API qz_status qz_scan_topk_chunked(qz_index *buf_a, qz_index *buf_b, const uint8_t *host_storage, uint64_t count,
                                      const void *enc_query, uint32_t k, int32_t largest, float *scores,
                                      uint32_t *ids, uint32_t *n_out);

/* ---- multi-GPU: top-k exchange over peer memory, fused into the scan launch -------------------------
 * One process per GPU, shards = contiguous point-id ranges (id_base = first row).  Replaces, per query,
 * qz_scan_topk_dev + an all-gather of the k keys + qz_merge_topk_dev with ONE kernel launch per GPU: the scan's
 * last thread block stores its k best keys into every peer's buffer over NVLink, raises a flag, waits for the
 * peers' flags and merges.  Every rank must make the same sequence of calls (like a collective).  Any codec (PQ:
 * row pitch a multiple of 16 bytes), k <= 128, non-empty shards; a peer that never arrives makes the kernel trap after ~30 s.
 *   qz_peer_create   allocates this GPU's buffer; handle_out = QZ_PEER_HANDLE_BYTES bytes (a CUDA IPC handle)
 *                    that the caller ships to the other ranks
 *   qz_peer_connect  all_handles = [world][QZ_PEER_HANDLE_BYTES] in rank order; maps the peers' buffers
 *   qz_scan_topk_peer_dev  as qz_scan_topk_dev, results = the global top-k on every rank */
#define QZ_PEER_HANDLE_BYTES 64
typedef struct qz_peer qz_peer;
QZ_API qz_status qz_peer_create(qz_ctx *ctx, uint32_t world, uint32_t rank, uint8_t *handle_out, qz_peer **out);
QZ_API qz_status qz_peer_connect(qz_peer *peer, const uint8_t *all_handles);
QZ_API void qz_peer_destroy(qz_peer *peer);
QZ_API qz_status qz_scan_topk_peer_dev(qz_index *index, qz_peer *peer, const void *d_enc_query, uint32_t k,
                                       int32_t largest, uint64_t *d_keys, float *d_scores, uint32_t *d_ids);
/* host-buffer form (the multi-GPU qz_search_topk): f32 query in, the global top-k out on every rank, one call,
 * one synchronisation; *n_out = min(k, rows over all shards) */
QZ_API qz_status qz_search_topk_peer(qz_index *index, qz_peer *peer, const float *query, uint32_t k, int32_t largest,
                                     float *scores, uint32_t *ids, uint32_t *n_out);
/* Rooted form — the gather BASELINE.json's north_star names: only rank `root` waits for the other ranks' keys and
 * merges; its outputs hold the global top-k.  Every other rank stores its k keys into the root's buffer, raises its
 * flag and finishes its launch without waiting for anybody (it may run up to two steps ahead of the root; the root's
 * acknowledgement, written back over NVLink after each merge, is the flow control) — its outputs are NOT written
 * and the host form returns *n_out = 0 without synchronising.  Same call sequence on every rank, like a collective;
 * rooted and all-to-all steps may be mixed on one peer group. */
QZ_API qz_status qz_scan_topk_gather_dev(qz_index *index, qz_peer *peer, uint32_t root, const void *d_enc_query,
                                         uint32_t k, int32_t largest, uint64_t *d_keys, float *d_scores, uint32_t *d_ids);
QZ_API qz_status qz_search_topk_gather(qz_index *index, qz_peer *peer, uint32_t root, const float *query, uint32_t k,
                                       int32_t largest, float *scores, uint32_t *ids, uint32_t *n_out);

#ifdef __cplusplus
}
#endif
#endif /* QZ_CUDA_H */
