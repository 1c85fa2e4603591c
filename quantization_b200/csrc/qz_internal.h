// qz_internal.h — host-side structures behind the opaque handles of include/qz_cuda.h
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <mutex>
#include <string>
#include <unordered_set>

#include "../../include/qz_cuda.h"

struct qz_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int sm_count = 0;
    size_t smem_optin = 0;
    std::string err;
    uint64_t launches = 0;
    std::mutex mu;  // serialises host entry points that share the scratch below
    std::atomic<int> refs{1};  // the context handle + every live index; freed when the last one goes

    // device scratch
    uint64_t *d_partial = nullptr;  // [max(sm_count,256)][QZ_MAX_K] per-CTA top-k lists
    uint64_t *d_merge_tmp = nullptr;
    uint32_t *d_done = nullptr;     // CTA arrival counter of the scan's in-launch merge (zero between launches)
    uint64_t *d_keys = nullptr;     // [QZ_MAX_K]
    float *d_scores_k = nullptr;    // [QZ_MAX_K]
    uint32_t *d_ids_k = nullptr;    // [QZ_MAX_K]
    uint8_t *d_query = nullptr;     // staged encoded query / f32 query
    size_t d_query_bytes = 0;
    uint8_t *d_io = nullptr;        // generic staging (id lists, score arrays)
    size_t d_io_bytes = 0;
    // pinned host staging
    uint8_t *h_pin = nullptr;
    size_t h_pin_bytes = 0;
    // optional in-stream timing of the dominant scan kernel (qz_ctx_set_timing)
    bool timing = false;
    static constexpr int kTimingRing = 256;
    cudaEvent_t ev_start[kTimingRing] = {}, ev_stop[kTimingRing] = {};
    uint64_t ev_head = 0, ev_tail = 0;  // [tail, head) pairs are recorded and not yet read
    double timed_ms = 0.0;
    // Chained scans (row_scan.cu): a fused top-k scan launched right behind another one of this context may start
    // while the previous one's last CTA is still merging (programmatic dependent launch).  Eligible when nothing
    // else can have been enqueued between the two: the context owns its stream, or the caller said so
    // (qz_ctx_set_scan_chaining), and no other kernel of this library was launched in between.
    bool chain_ok = false;      // caller's promise for a stream the context does not own
    uint64_t chain_tail = 0;    // value of `launches` right after the last chainable scan launch, 0 = none
    uint64_t scan_seq = 0;      // fused scans launched with a tile counter (picks one of the two counters)
    bool query_dep_next = false;  // set by the host-facing search entry points right after their query-encode launch:
                                  // the next row scan is launched as its programmatic dependent (RowScanParams::query_dep)
    bool pdl_next = false;      // launch_one: give the next launch the programmatic-serialization attribute
    uint64_t *d_stamps = nullptr;  // qz_ctx_set_phase_stamps: [sm_count][QZ_STAMP_WORDS], or NULL
    uint64_t timed_n = 0;
    // One-time kernel attribute setup.  cudaFuncSetAttribute(MaxDynamicSharedMemorySize) applies to the CURRENT
    // device only, so "done" is a property of the context (= device), never of the process: a host that creates
    // contexts on several devices needs the attribute on each of them (qz::ensure_smem_attr).
    std::mutex attr_mu;
    std::unordered_set<const void *> attr_done;
    int batch_cl_pick = 0;                     // u8_batch.cu: cluster size chosen for this device, 0 = not yet
    int batch_cl_grid[3] = {0, 0, 0};          // resident CTAs for cluster size 1 / 2
};

struct qz_index {
    qz_ctx *ctx = nullptr;
    int codec = 0;
    uint64_t count = 0;
    uint32_t dim = 0;
    uint32_t row_bytes = 0;
    int distance = 0, invert = 0;
    uint32_t id_base = 0;
    uint8_t *d_storage = nullptr;
    size_t alloc_bytes = 0;
    // u8 (Metadata, encoded_vectors_u8.rs:24-31)
    uint32_t actual_dim = 0;
    float alpha = 0, offset = 0, multiplier = 0;
    // batched tensor-core path (u8_batch.cu): SoA mirror + per-batch scratch, built lazily
    uint8_t *d_codes_soa = nullptr;  // [count][actual_dim]
    float *d_voff_soa = nullptr;     // [count]
    int32_t *d_vq_soa = nullptr;     // [count] floor(vector offset / multiplier) (integer filter of the batched path)
    bool soa_valid = false;          // the mirror matches the storage and the Metadata scalars
    bool vq_ok = false;              // d_vq_soa is usable (finite non-zero multiplier, offsets in range)
    float voff_absmax = 0.0f;        // max |vector offset| (bound used by the batched fast filter)
    uint8_t *d_batch = nullptr;      // scratch arena for one batch call
    size_t batch_bytes = 0;
    // binary
    int store = 16;
    // pq (Metadata, encoded_vectors_pq.rs:39-44)
    uint32_t n_chunks = 0;
    float *d_centroids = nullptr;    // [256][dim]
    uint32_t *d_bounds = nullptr;    // [n_chunks+1]
    uint32_t *h_bounds = nullptr;
    float *d_lut_perm = nullptr;     // scratch: bank-permuted LUT of the current query
    float *d_pq_partial = nullptr;   // scratch: [n_slices][count] partial sums
    uint32_t pq_slices = 0;
};

namespace qz {

qz_status fail(qz_ctx *ctx, qz_status st, const std::string &msg);
void ctx_release(qz_ctx *ctx);  // drop one reference (qz_ctx_destroy, qz_index_destroy)
qz_status cuda_fail(qz_ctx *ctx, cudaError_t e, const char *what);
qz_status ensure_query(qz_ctx *ctx, size_t bytes);
qz_status ensure_io(qz_ctx *ctx, size_t bytes);
qz_status ensure_pin(qz_ctx *ctx, size_t bytes);
// opt the kernel `fn` into `bytes` of dynamic shared memory on this context's device, once per context
qz_status ensure_smem_attr(qz_ctx *ctx, const void *fn, int bytes);
// bracket the dominant kernel of a scan with events when timing is on
void timing_begin(qz_ctx *ctx);
void timing_end(qz_ctx *ctx);

#define QZ_CUDA(ctx, call)                                          \
    do {                                                            \
        cudaError_t e__ = (call);                                   \
        if (e__ != cudaSuccess) return qz::cuda_fail(ctx, e__, #call); \
    } while (0)

#define QZ_TRY(call)                   \
    do {                               \
        qz_status s__ = (call);        \
        if (s__ != QZ_OK) return s__;  \
    } while (0)

// Experiment switches (tile sizes, ring depth, schedules ...) exist only in builds made with -DQZ_TUNING
// (QZ_TUNING=1 python -m quantization_b200.build, used by the sweeps under tools/); a production build ignores
// the environment entirely.
#ifdef QZ_TUNING
#include <stdlib.h>
static inline uint32_t tune_u32(const char *name, uint32_t dflt) {
    const char *e = getenv(name);
    return e ? (uint32_t)atoi(e) : dflt;
}
static inline bool tune_flag(const char *name) { return getenv(name) != nullptr; }
#else
static inline uint32_t tune_u32(const char *, uint32_t dflt) { return dflt; }
static inline bool tune_flag(const char *) { return false; }
#endif

static inline uint32_t pow2_ceil(uint32_t v) {
    uint32_t p = 1;
    while (p < v) p <<= 1;
    return p;
}
static inline uint32_t topk_pad(uint32_t k) { return pow2_ceil(k < 128 ? 128 : k); }

// scans (row_scan.cu)
struct ScanFinal {  // where the scan kernel's in-launch final merge writes (each pointer may be NULL)
    uint64_t *d_keys;
    float *d_scores;
    uint32_t *d_ids;
    bool *fused;  // out: the kernel did the final merge
    const void *px = nullptr;  // optional qz::PeerExchange (row_scan.cuh): cross-GPU exchange in the same launch
};
qz_status launch_row_scan(qz_index *ix, const void *d_query, float *d_scores, uint32_t k, int largest,
                          uint32_t *n_lists, uint32_t *Kp_out, const ScanFinal *fin = nullptr);
// merge (topk.cu): lists at d_in -> final keys/scores/ids
qz_status launch_merge(qz_ctx *ctx, const uint64_t *d_in, uint32_t n_lists, uint32_t in_len, uint32_t in_stride,
                       uint32_t k, int largest, uint64_t *d_keys, float *d_scores, uint32_t *d_ids);

// point_kernels.cu
qz_status launch_u8_encode_query(qz_index *ix, const float *d_queries, uint32_t nq, void *d_out);
qz_status launch_gather(qz_index *ix, const void *d_query, const uint32_t *d_rows_a, const uint32_t *d_rows_b,
                        uint64_t n, float *d_out);
qz_status launch_scores_topk(qz_ctx *ctx, const float *d_parts, uint32_t n_parts, uint64_t part_stride, uint64_t n,
                             float *d_sum_out, uint32_t k, int largest, uint32_t id_base, uint32_t *n_lists,
                             uint32_t *Kp_out);

}  // namespace qz
