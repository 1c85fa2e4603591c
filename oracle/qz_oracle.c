/*
 * qz_oracle.c — CPU restatement of qdrant/quantization's encoded-vector hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under quantization_b200/ may import, link or
 * execute this file; it is the checker for tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py.
 *
 * Parity status: the integer kernels (dot, L1, xor-popcount) are pinned against the
 * reference's own C scorers compiled from /root/reference/quantization/cpp/{sse,avx2}.c
 * into oracle/_ref/libsimd_utils_ref.so (see oracle/Makefile, tests/test_oracle_vs_ref.py).
 * The Rust-only arithmetic (encode, headers, LUT, k-means) has no golden vectors in the
 * reference (its tests are tolerance properties on StdRng data) and no
 * rustc exists here, so for those functions parity is "restated line by line, pinned by
 * the reference's own test tolerances" (tests/test_reference_properties.py).
 *
 * All f32 arithmetic is written as separate IEEE operations in the reference's order;
 * build with -ffp-contract=off -fno-fast-math (oracle/Makefile does).
 *
 * Citations are path:line relative to /root/reference/quantization/.
 */
#include <math.h>
#include <float.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define QZO_DOT 0
#define QZO_L1 1
#define QZO_L2 2

#define QZO_EXPORT __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------
 * Deterministic synthetic data: SplitMix64 -> f32 in [0,1) by (x >> 40) * 2^-24.
 * Value i of stream `seed` depends only on (seed, i) so any slice can be generated anywhere.
 * ---------------------------------------------------------------------------------------- */
static inline uint64_t splitmix64_at(uint64_t seed, uint64_t i) {
    uint64_t z = seed + (i + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

QZO_EXPORT void qzo_fill_uniform(uint64_t seed, uint64_t first, uint64_t n, float lo, float hi, float *out) {
    for (uint64_t i = 0; i < n; i++) {
        float u = (float)(splitmix64_at(seed, first + i) >> 40) * (1.0f / 16777216.0f);
        out[i] = lo + (hi - lo) * u;
    }
}

/* +-1 data as in tests/test_binary.rs:14-25 and demos/benches/binary.rs:10-21 */
QZO_EXPORT void qzo_fill_signs(uint64_t seed, uint64_t first, uint64_t n, float *out) {
    for (uint64_t i = 0; i < n; i++) out[i] = (splitmix64_at(seed, first + i) >> 63) ? 1.0f : -1.0f;
}

/* ------------------------------------------------------------------------------------------
 * Ordering contract shared by oracle and GPU: a (score,id) pair becomes a u64
 * key, larger key = better.  High word: order-preserving map of the f32 bit pattern
 * (complemented when smaller scores are better); low word: 0xFFFFFFFF - id, so that ties on
 * score go to the lower point id.  Key 0 is the "empty" sentinel (ids <= 0xFFFFFFFE).
 * ---------------------------------------------------------------------------------------- */
static inline uint32_t f32_order_bits(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
static inline float f32_from_order_bits(uint32_t o) {
    uint32_t u = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
    float f;
    memcpy(&f, &u, 4);
    return f;
}
QZO_EXPORT uint64_t qzo_make_key(float score, uint32_t id, int largest) {
    uint32_t o = f32_order_bits(score);
    if (!largest) o = ~o;
    return ((uint64_t)o << 32) | (uint64_t)(0xFFFFFFFFu - id);
}
QZO_EXPORT void qzo_split_key(uint64_t key, int largest, float *score, uint32_t *id) {
    uint32_t o = (uint32_t)(key >> 32);
    if (!largest) o = ~o;
    *score = f32_from_order_bits(o);
    *id = 0xFFFFFFFFu - (uint32_t)key;
}

static int cmp_u64_desc(const void *a, const void *b) {
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? 1 : (x > y ? -1 : 0);
}

/* size-k selection over a score array; replaces the caller-side BinaryHeap of
 * demos/src/ann_benchmark_data.rs:150-167 with the deterministic order above.
 * Returns number of results written (min(k,n)), best first. */
QZO_EXPORT uint32_t qzo_topk(const float *scores, uint64_t n, uint32_t k, int largest, uint32_t id_base,
                             float *out_scores, uint32_t *out_ids) {
    if (k == 0 || n == 0) return 0;
    uint32_t cap = k;
    uint64_t *heap = (uint64_t *)malloc(sizeof(uint64_t) * cap); /* min-heap on key */
    uint32_t sz = 0;
    for (uint64_t i = 0; i < n; i++) {
        uint64_t key = qzo_make_key(scores[i], id_base + (uint32_t)i, largest);
        if (sz < cap) {
            uint32_t c = sz++;
            heap[c] = key;
            while (c > 0) {
                uint32_t p = (c - 1) / 2;
                if (heap[p] <= heap[c]) break;
                uint64_t t = heap[p]; heap[p] = heap[c]; heap[c] = t;
                c = p;
            }
        } else if (key > heap[0]) {
            heap[0] = key;
            uint32_t p = 0;
            for (;;) {
                uint32_t l = 2 * p + 1, r = l + 1, m = p;
                if (l < sz && heap[l] < heap[m]) m = l;
                if (r < sz && heap[r] < heap[m]) m = r;
                if (m == p) break;
                uint64_t t = heap[p]; heap[p] = heap[m]; heap[m] = t;
                p = m;
            }
        }
    }
    qsort(heap, sz, sizeof(uint64_t), cmp_u64_desc);
    for (uint32_t i = 0; i < sz; i++) qzo_split_key(heap[i], largest, &out_scores[i], &out_ids[i]);
    free(heap);
    return sz;
}

/* merge of per-shard top-k lists (keys), the step the GPU does after the NCCL all-gather */
QZO_EXPORT uint32_t qzo_merge_topk(const float *scores, const uint32_t *ids, uint64_t n, uint32_t k, int largest,
                                   float *out_scores, uint32_t *out_ids) {
    uint64_t *keys = (uint64_t *)malloc(sizeof(uint64_t) * (n ? n : 1));
    for (uint64_t i = 0; i < n; i++) keys[i] = qzo_make_key(scores[i], ids[i], largest);
    qsort(keys, n, sizeof(uint64_t), cmp_u64_desc);
    uint32_t m = (uint32_t)(n < k ? n : k);
    for (uint32_t i = 0; i < m; i++) qzo_split_key(keys[i], largest, &out_scores[i], &out_ids[i]);
    free(keys);
    return m;
}

/* ==========================================================================================
 * u8 scalar quantizer — src/encoded_vectors_u8.rs
 * ======================================================================================== */

/* encoded_vectors_u8.rs:257-259 (ALIGNMENT = 16, :12) */
QZO_EXPORT uint32_t qzo_u8_actual_dim(uint32_t dim) { return dim + (16 - dim % 16) % 16; }

/* quantile.rs:5-19 — fold from (f32::MAX, f32::MIN) with strict < / > (NaN never wins) */
QZO_EXPORT void qzo_minmax(const float *data, uint64_t n, float *mn, float *mx) {
    float lo = FLT_MAX, hi = -FLT_MAX;
    for (uint64_t i = 0; i < n; i++) {
        float v = data[i];
        if (v < lo) lo = v;
        if (v > hi) hi = v;
    }
    *mn = lo;
    *mx = hi;
}

/* encoded_vectors_u8.rs:228-232 */
QZO_EXPORT void qzo_u8_alpha_offset(float mn, float mx, float *alpha, float *offset) {
    *alpha = (mx - mn) / 127.0f;
    *offset = mn;
}

/* encoded_vectors_u8.rs:234-237 — Rust `clamp` keeps NaN, `as u8` truncates toward zero,
 * saturates and maps NaN to 0. */
static inline uint8_t f32_to_u8(float v, float alpha, float offset) {
    float t = (v - offset) / alpha;
    if (t != t) return 0;
    if (t < 0.0f) t = 0.0f;
    if (t > 127.0f) t = 127.0f;
    return (uint8_t)t;
}
QZO_EXPORT uint8_t qzo_u8_code(float v, float alpha, float offset) { return f32_to_u8(v, alpha, offset); }

static int cmp_f32(const void *a, const void *b) {
    float x = *(const float *)a, y = *(const float *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}

/* quantile.rs:21-71 applied to an already gathered sample (`values` = the flattened sampled
 * vectors, `slice_size` = number of sampled vectors).  The sample *ids* come from a randomly
 * keyed Permutor in the reference (quantile.rs:32-34), so they are an explicit input here.
 * Two select_nth_unstable_by calls keep sorted positions [cut+1, len-cut-1]; a full sort
 * gives the same two values.  `values` is reordered in place.  Returns 1 if an interval is
 * produced. */
QZO_EXPORT int qzo_quantile_interval(float *values, uint64_t len, uint64_t count, uint64_t slice_size, float quantile,
                                     float *lo, float *hi) {
    if (count < 127 || quantile >= 1.0f) return 0; /* quantile.rs:27-29 */
    if (len < 4) return 0;                          /* :48-51 */
    uint64_t a = (len - 1) / 2;
    float fb = (float)slice_size * (1.0f - quantile) / 2.0f; /* :55 — vector count, not value count */
    uint64_t b = fb <= 0.0f ? 0 : (uint64_t)fb;
    uint64_t cut = a < b ? a : b;
    if (cut < 1) cut = 1; /* :57 */
    qsort(values, len, sizeof(float), cmp_f32);
    /* left = sorted[0 .. len-cut) ; then right part of select_nth(cut) = sorted(cut .. len-cut) exclusive */
    uint64_t first = cut + 1, last = len - cut - 1; /* inclusive positions that remain */
    if (last < first || last - first + 1 < 2) return 0; /* :63-65 */
    *lo = values[first];
    *hi = values[last];
    return 1;
}

/* encoded_vectors_u8.rs:119-128 */
QZO_EXPORT float qzo_u8_multiplier(float alpha, int distance, int invert) {
    float m = distance == QZO_DOT ? alpha * alpha : (distance == QZO_L1 ? alpha : -2.0f * alpha * alpha);
    return invert ? -m : m;
}

/* one row: encoded_vectors_u8.rs:78-117.  out = [f32 header][actual_dim codes] */
static void u8_encode_row(const float *v, uint32_t dim, uint32_t actual_dim, int distance, int invert, float alpha,
                          float offset, uint8_t *out) {
    memset(out, 0, 4);
    uint8_t *codes = out + 4;
    for (uint32_t i = 0; i < dim; i++) codes[i] = f32_to_u8(v[i], alpha, offset);
    float placeholder = distance == QZO_DOT ? 0.0f : offset; /* :86-89 */
    uint8_t pad = f32_to_u8(placeholder, alpha, offset);
    for (uint32_t i = dim; i < actual_dim; i++) codes[i] = pad;
    float vo;
    if (distance == QZO_DOT) { /* :95-98; the iterator also visits the 4 zero header bytes */
        float s = 0.0f;
        for (uint32_t i = 0; i < actual_dim + 4; i++) s += (float)out[i];
        vo = (float)actual_dim * offset * offset + s * alpha * offset;
    } else if (distance == QZO_L1) {
        vo = 0.0f;
    } else { /* :100-108 */
        float s = 0.0f;
        for (uint32_t i = 0; i < actual_dim + 4; i++) s += (float)out[i] * (float)out[i];
        vo = (float)actual_dim * offset * offset + s * alpha * alpha;
    }
    if (invert) vo = -vo;
    memcpy(out, &vo, 4); /* native endian, :115-116 */
}

QZO_EXPORT void qzo_u8_encode(const float *data, uint64_t count, uint32_t dim, int distance, int invert, float alpha,
                              float offset, uint8_t *storage) {
    uint32_t ad = qzo_u8_actual_dim(dim);
    for (uint64_t r = 0; r < count; r++)
        u8_encode_row(data + r * dim, dim, ad, distance, invert, alpha, offset, storage + r * (uint64_t)(ad + 4));
}

/* encoded_vectors_u8.rs:290-329.  out = [f32 query offset][actual_dim codes] (same shape as a row) */
QZO_EXPORT void qzo_u8_encode_query(const float *q, uint32_t dim, int distance, int invert, float alpha, float offset,
                                    uint8_t *out) {
    uint32_t ad = qzo_u8_actual_dim(dim);
    uint8_t *codes = out + 4;
    for (uint32_t i = 0; i < dim; i++) codes[i] = f32_to_u8(q[i], alpha, offset);
    float placeholder = distance == QZO_DOT ? 0.0f : offset;
    uint8_t pad = f32_to_u8(placeholder, alpha, offset);
    for (uint32_t i = dim; i < ad; i++) codes[i] = pad;
    float qo;
    if (distance == QZO_DOT) {
        float s = 0.0f;
        for (uint32_t i = 0; i < ad; i++) s += (float)codes[i];
        qo = s * alpha * offset;
    } else if (distance == QZO_L1) {
        qo = 0.0f;
    } else {
        float s = 0.0f;
        for (uint32_t i = 0; i < ad; i++) s += (float)codes[i] * (float)codes[i];
        qo = s * alpha * alpha;
    }
    if (invert) qo = -qo;
    memcpy(out, &qo, 4);
}

/* encoded_vectors_u8.rs:456-464 / :466-474 — the scalar integer kernels, returned as f32 the
 * way score_point_simple casts them (:158) */
QZO_EXPORT float qzo_score_dot(const uint8_t *q, const uint8_t *v, uint32_t dim) {
    int32_t s = 0;
    for (uint32_t i = 0; i < dim; i++) s += (int32_t)q[i] * (int32_t)v[i];
    return (float)s;
}
QZO_EXPORT float qzo_score_l1(const uint8_t *q, const uint8_t *v, uint32_t dim) {
    int32_t s = 0;
    for (uint32_t i = 0; i < dim; i++) {
        int32_t d = (int32_t)q[i] - (int32_t)v[i];
        s += d < 0 ? -d : d;
    }
    return (float)s;
}

typedef float (*qzo_u8_kernel)(const uint8_t *, const uint8_t *, uint32_t);

/* encoded_vectors_u8.rs:331-384 for rows [first, first+n): multiplier*S + q.offset + v.offset,
 * evaluated left to right (:347).  `kernel` is the integer scorer: NULL = the scalar one above,
 * otherwise e.g. impl_score_dot_avx from oracle/_ref (cpp/avx2.c:31-69). */
QZO_EXPORT void qzo_u8_scan(const uint8_t *storage, uint64_t first, uint64_t n, uint32_t actual_dim, int distance,
                            float multiplier, const uint8_t *query, qzo_u8_kernel kernel, float *out) {
    if (!kernel) kernel = distance == QZO_L1 ? qzo_score_l1 : qzo_score_dot;
    float qo;
    memcpy(&qo, query, 4);
    const uint8_t *qc = query + 4;
    uint64_t stride = (uint64_t)actual_dim + 4;
    for (uint64_t i = 0; i < n; i++) {
        const uint8_t *row = storage + (first + i) * stride;
        float vo;
        memcpy(&vo, row, 4);
        float s = kernel(qc, row + 4, actual_dim);
        float t = multiplier * s;
        t = t + qo;
        out[i] = t + vo;
    }
}

/* encoded_vectors_u8.rs:386-453 */
QZO_EXPORT void qzo_u8_score_pairs(const uint8_t *storage, uint32_t actual_dim, int distance, int invert,
                                   float multiplier, float offset, const uint32_t *ia, const uint32_t *ib, uint64_t n,
                                   qzo_u8_kernel kernel, float *out) {
    if (!kernel) kernel = distance == QZO_L1 ? qzo_score_l1 : qzo_score_dot;
    uint64_t stride = (uint64_t)actual_dim + 4;
    float diff = (float)actual_dim * offset * offset; /* :389 */
    if (invert) diff = -diff;
    for (uint64_t p = 0; p < n; p++) {
        const uint8_t *ri = storage + ia[p] * stride, *rj = storage + ib[p] * stride;
        float oi, oj;
        memcpy(&oi, ri, 4);
        memcpy(&oj, rj, 4);
        float off = oi + oj - diff; /* :395 */
        float s = kernel(ri + 4, rj + 4, actual_dim);
        out[p] = multiplier * s + off; /* :409 */
    }
}

/* threaded scan + top-k: how a caller with all host cores would drive score_point
 * (contiguous id ranges per thread, per-thread top-k merged) — the CPU baseline of SURVEY §8d */
typedef struct {
    const uint8_t *storage; uint64_t first, n; uint32_t actual_dim; int distance; float multiplier;
    const uint8_t *query; qzo_u8_kernel kernel; uint32_t k; int largest; uint32_t id_base;
    float *scores; uint32_t *ids; uint32_t got;
} u8_job;

static void *u8_job_run(void *p) {
    u8_job *j = (u8_job *)p;
    float *tmp = (float *)malloc(sizeof(float) * (j->n ? j->n : 1));
    qzo_u8_scan(j->storage, j->first, j->n, j->actual_dim, j->distance, j->multiplier, j->query, j->kernel, tmp);
    j->got = qzo_topk(tmp, j->n, j->k, j->largest, j->id_base + (uint32_t)j->first, j->scores, j->ids);
    free(tmp);
    return NULL;
}

QZO_EXPORT uint32_t qzo_u8_scan_topk_mt(const uint8_t *storage, uint64_t count, uint32_t actual_dim, int distance,
                                        float multiplier, const uint8_t *query, qzo_u8_kernel kernel, uint32_t k,
                                        int largest, uint32_t id_base, int n_threads, float *out_scores,
                                        uint32_t *out_ids) {
    if (n_threads < 1) n_threads = 1;
    u8_job *jobs = (u8_job *)calloc(n_threads, sizeof(u8_job));
    pthread_t *th = (pthread_t *)calloc(n_threads, sizeof(pthread_t));
    float *ps = (float *)malloc(sizeof(float) * (size_t)n_threads * k);
    uint32_t *pi = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n_threads * k);
    for (int t = 0; t < n_threads; t++) {
        uint64_t a = count * t / n_threads, b = count * (t + 1) / n_threads;
        jobs[t] = (u8_job){storage, a, b - a, actual_dim, distance, multiplier, query, kernel, k, largest, id_base,
                           ps + (size_t)t * k, pi + (size_t)t * k, 0};
        if (n_threads > 1) pthread_create(&th[t], NULL, u8_job_run, &jobs[t]);
        else u8_job_run(&jobs[t]);
    }
    uint64_t tot = 0;
    float *cs = (float *)malloc(sizeof(float) * (size_t)n_threads * k);
    uint32_t *ci = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n_threads * k);
    for (int t = 0; t < n_threads; t++) {
        if (n_threads > 1) pthread_join(th[t], NULL);
        memcpy(cs + tot, jobs[t].scores, sizeof(float) * jobs[t].got);
        memcpy(ci + tot, jobs[t].ids, sizeof(uint32_t) * jobs[t].got);
        tot += jobs[t].got;
    }
    uint32_t m = qzo_merge_topk(cs, ci, tot, k, largest, out_scores, out_ids);
    free(cs); free(ci); free(ps); free(pi); free(th); free(jobs);
    return m;
}

/* ==========================================================================================
 * binary quantizer — src/encoded_vectors_binary.rs
 * ======================================================================================== */

/* bytes per row. store = 1 (u8 words, :99-116) or 16 (u128 words, :152-159,:210-213) */
QZO_EXPORT uint32_t qzo_bin_row_bytes(uint32_t dim, int store) {
    if (store == 16) return ((dim + 127) / 128) * 16;
    uint32_t bytes = dim > 128 ? 16 : (dim > 64 ? 8 : (dim > 32 ? 4 : 1));
    uint32_t bits = 8 * bytes;
    return ((dim + bits - 1) / bits) * bytes;
}

/* encode_vector, :193-208 — bit i = (v[i] > 0.0) at byte i/8, bit i%8 (same bytes for both
 * store types on a little-endian host) */
QZO_EXPORT void qzo_bin_encode(const float *data, uint64_t count, uint32_t dim, int store, uint8_t *storage) {
    uint32_t rb = qzo_bin_row_bytes(dim, store);
    for (uint64_t r = 0; r < count; r++) {
        uint8_t *row = storage + r * (uint64_t)rb;
        memset(row, 0, rb);
        const float *v = data + r * (uint64_t)dim;
        for (uint32_t i = 0; i < dim; i++)
            if (v[i] > 0.0f) row[i >> 3] |= (uint8_t)(1u << (i & 7));
    }
}

/* scalar xor_popcnt, :92-96 / :145-149 */
QZO_EXPORT uint32_t qzo_bin_xor_popcnt(const uint8_t *a, const uint8_t *b, uint32_t bytes) {
    uint32_t r = 0;
    for (uint32_t i = 0; i < bytes; i++) r += (uint32_t)__builtin_popcount((unsigned)(a[i] ^ b[i]));
    return r;
}

typedef uint32_t (*qzo_bin_kernel)(const uint8_t *, const uint8_t *, uint32_t);

/* dispatch of u8::xor_popcnt (:48-71) / u128::xor_popcnt (:123-131) onto the reference's C
 * kernels (k128/k64/k32 from oracle/_ref, any may be NULL => scalar loop) */
static uint32_t bin_popcnt_dispatch(const uint8_t *a, const uint8_t *b, uint32_t bytes, int store, qzo_bin_kernel k128,
                                    qzo_bin_kernel k64, qzo_bin_kernel k32) {
    if (store == 16) return k128 ? k128(a, b, bytes / 16) : qzo_bin_xor_popcnt(a, b, bytes);
    if (bytes > 16 && k128) return k128(a, b, bytes / 16);
    if (bytes > 8 && bytes <= 16 && k64) return k64(a, b, bytes / 8);
    if (bytes > 4 && bytes <= 8 && k32) return k32(a, b, bytes / 4);
    return qzo_bin_xor_popcnt(a, b, bytes);
}

/* calculate_metric, :237-252 */
static inline float bin_metric(uint32_t xor_cnt, uint32_t dim, int distance, int invert) {
    float x = (float)xor_cnt;
    float z = (float)dim - x;
    int dot = distance == QZO_DOT;
    if ((dot && !invert) || (!dot && invert)) return z - x;
    return x - z;
}
QZO_EXPORT float qzo_bin_metric(uint32_t xor_cnt, uint32_t dim, int distance, int invert) {
    return bin_metric(xor_cnt, dim, distance, invert);
}

/* score_point over rows [first, first+n), :293-300 */
QZO_EXPORT void qzo_bin_scan(const uint8_t *storage, uint64_t first, uint64_t n, uint32_t dim, int store, int distance,
                             int invert, const uint8_t *query, qzo_bin_kernel k128, qzo_bin_kernel k64,
                             qzo_bin_kernel k32, float *out) {
    uint32_t rb = qzo_bin_row_bytes(dim, store);
    for (uint64_t i = 0; i < n; i++) {
        const uint8_t *row = storage + (first + i) * (uint64_t)rb;
        out[i] = bin_metric(bin_popcnt_dispatch(row, query, rb, store, k128, k64, k32), dim, distance, invert);
    }
}

/* score_internal, :302-314 */
QZO_EXPORT void qzo_bin_score_pairs(const uint8_t *storage, uint32_t dim, int store, int distance, int invert,
                                    const uint32_t *ia, const uint32_t *ib, uint64_t n, float *out) {
    uint32_t rb = qzo_bin_row_bytes(dim, store);
    for (uint64_t p = 0; p < n; p++)
        out[p] = bin_metric(qzo_bin_xor_popcnt(storage + ia[p] * (uint64_t)rb, storage + ib[p] * (uint64_t)rb, rb), dim,
                            distance, invert);
}

typedef struct {
    const uint8_t *storage; uint64_t first, n; uint32_t dim; int store, distance, invert; const uint8_t *query;
    qzo_bin_kernel k128, k64, k32; uint32_t k; int largest; uint32_t id_base; float *scores; uint32_t *ids; uint32_t got;
} bin_job;

static void *bin_job_run(void *p) {
    bin_job *j = (bin_job *)p;
    float *tmp = (float *)malloc(sizeof(float) * (j->n ? j->n : 1));
    qzo_bin_scan(j->storage, j->first, j->n, j->dim, j->store, j->distance, j->invert, j->query, j->k128, j->k64, j->k32,
                 tmp);
    j->got = qzo_topk(tmp, j->n, j->k, j->largest, j->id_base + (uint32_t)j->first, j->scores, j->ids);
    free(tmp);
    return NULL;
}

QZO_EXPORT uint32_t qzo_bin_scan_topk_mt(const uint8_t *storage, uint64_t count, uint32_t dim, int store, int distance,
                                         int invert, const uint8_t *query, qzo_bin_kernel k128, qzo_bin_kernel k64,
                                         qzo_bin_kernel k32, uint32_t k, int largest, uint32_t id_base, int n_threads,
                                         float *out_scores, uint32_t *out_ids) {
    if (n_threads < 1) n_threads = 1;
    bin_job *jobs = (bin_job *)calloc(n_threads, sizeof(bin_job));
    pthread_t *th = (pthread_t *)calloc(n_threads, sizeof(pthread_t));
    float *ps = (float *)malloc(sizeof(float) * (size_t)n_threads * k);
    uint32_t *pi = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n_threads * k);
    for (int t = 0; t < n_threads; t++) {
        uint64_t a = count * t / n_threads, b = count * (t + 1) / n_threads;
        jobs[t] = (bin_job){storage, a, b - a, dim, store, distance, invert, query, k128, k64, k32, k, largest, id_base,
                            ps + (size_t)t * k, pi + (size_t)t * k, 0};
        if (n_threads > 1) pthread_create(&th[t], NULL, bin_job_run, &jobs[t]);
        else bin_job_run(&jobs[t]);
    }
    uint64_t tot = 0;
    float *cs = (float *)malloc(sizeof(float) * (size_t)n_threads * k);
    uint32_t *ci = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n_threads * k);
    for (int t = 0; t < n_threads; t++) {
        if (n_threads > 1) pthread_join(th[t], NULL);
        memcpy(cs + tot, jobs[t].scores, sizeof(float) * jobs[t].got);
        memcpy(ci + tot, jobs[t].ids, sizeof(uint32_t) * jobs[t].got);
        tot += jobs[t].got;
    }
    uint32_t m = qzo_merge_topk(cs, ci, tot, k, largest, out_scores, out_ids);
    free(cs); free(ci); free(ps); free(pi); free(th); free(jobs);
    return m;
}

/* ==========================================================================================
 * product quantizer — src/encoded_vectors_pq.rs, src/kmeans.rs, src/encoded_vectors.rs
 * ======================================================================================== */

/* get_vector_division, encoded_vectors_pq.rs:116-121; returns n_chunks, bounds[c]..bounds[c+1] */
QZO_EXPORT uint32_t qzo_pq_division(uint32_t dim, uint32_t chunk_size, uint32_t *bounds) {
    uint32_t n = 0;
    for (uint32_t i = 0; i < dim; i += chunk_size) {
        if (bounds) bounds[n] = i;
        n++;
    }
    if (bounds) bounds[n] = dim;
    return n;
}

/* DistanceType::distance, encoded_vectors.rs:37-45 — sequential f32 sum from 0.0 */
static inline float metric_distance(int distance, const float *a, const float *b, uint32_t n) {
    float s = 0.0f;
    if (distance == QZO_DOT) {
        for (uint32_t i = 0; i < n; i++) s += a[i] * b[i];
    } else if (distance == QZO_L1) {
        for (uint32_t i = 0; i < n; i++) s += fabsf(a[i] - b[i]);
    } else {
        for (uint32_t i = 0; i < n; i++) { float d = a[i] - b[i]; s += d * d; }
    }
    return s;
}

/* argmin over 256 centroids of the sequential squared distance, strict <, from f32::MAX:
 * encoded_vectors_pq.rs:244-264 and kmeans.rs:145-165.  centroids laid out [n_centroids][cdim] */
static inline uint32_t nearest_centroid(const float *sub, const float *centroids, uint32_t n_centroids, uint32_t cdim,
                                        uint32_t cstride, uint32_t coff) {
    float best = FLT_MAX;
    uint32_t bi = 0;
    for (uint32_t c = 0; c < n_centroids; c++) {
        const float *cen = centroids + (uint64_t)c * cstride + coff;
        float d = 0.0f;
        for (uint32_t t = 0; t < cdim; t++) { float x = sub[t] - cen[t]; d += x * x; }
        if (d < best) { best = d; bi = c; }
    }
    return bi;
}

/* encode_vector for all rows, encoded_vectors_pq.rs:237-265; centroids = [256][dim] */
QZO_EXPORT void qzo_pq_assign(const float *data, uint64_t count, uint32_t dim, const uint32_t *bounds,
                              uint32_t n_chunks, const float *centroids, uint32_t n_centroids, uint8_t *codes) {
    for (uint64_t r = 0; r < count; r++)
        for (uint32_t ch = 0; ch < n_chunks; ch++) {
            uint32_t a = bounds[ch], b = bounds[ch + 1];
            codes[r * n_chunks + ch] =
                (uint8_t)nearest_centroid(data + r * dim + a, centroids, n_centroids, b - a, dim, a);
        }
}

/* encode_query, encoded_vectors_pq.rs:525-547: lut[ch*256 + c] */
QZO_EXPORT void qzo_pq_lut(const float *q, uint32_t dim, const uint32_t *bounds, uint32_t n_chunks,
                           const float *centroids, uint32_t n_centroids, int distance, int invert, float *lut) {
    for (uint32_t ch = 0; ch < n_chunks; ch++) {
        uint32_t a = bounds[ch], b = bounds[ch + 1];
        for (uint32_t c = 0; c < n_centroids; c++) {
            float d = metric_distance(distance, q + a, centroids + (uint64_t)c * dim + a, b - a);
            lut[(uint64_t)ch * n_centroids + c] = invert ? -d : d;
        }
    }
}

/* score_point_sse, encoded_vectors_pq.rs:405-440: four interleaved lanes, (l0+l2)+(l1+l3), tail */
static inline float pq_score_sse_order(const uint8_t *codes, uint32_t n_chunks, const float *lut, uint32_t nc) {
    float l0 = 0.0f, l1 = 0.0f, l2 = 0.0f, l3 = 0.0f;
    uint32_t full = n_chunks / 4;
    const float *p = lut;
    for (uint32_t t = 0; t < full; t++) {
        l0 += p[codes[0]];
        l1 += p[nc + codes[1]];
        l2 += p[2 * nc + codes[2]];
        l3 += p[3 * nc + codes[3]];
        codes += 4;
        p += 4 * (uint64_t)nc;
    }
    float s0 = l0 + l2, s1 = l1 + l3; /* _mm_add_ps(sum128, movehl) :430 */
    float sum = s0 + s1;              /* _mm_add_ss(sum64, shuffle 0x55) :431 */
    for (uint32_t t = 0; t < n_chunks % 4; t++) {
        sum += p[codes[0]];
        codes++;
        p += nc;
    }
    return sum;
}
/* score_point_simple, :476-494 */
static inline float pq_score_simple(const uint8_t *codes, uint32_t n_chunks, const float *lut, uint32_t nc) {
    float s = 0.0f;
    for (uint32_t ch = 0; ch < n_chunks; ch++) s += lut[(uint64_t)ch * nc + codes[ch]];
    return s;
}

/* order: 0 = SSE order (what x86 runs, :549-553), 1 = scalar order */
QZO_EXPORT void qzo_pq_scan(const uint8_t *codes, uint64_t first, uint64_t n, uint32_t n_chunks, const float *lut,
                            uint32_t n_centroids, int order, float *out) {
    for (uint64_t i = 0; i < n; i++) {
        const uint8_t *row = codes + (first + i) * (uint64_t)n_chunks;
        out[i] = order == 0 ? pq_score_sse_order(row, n_chunks, lut, n_centroids)
                            : pq_score_simple(row, n_chunks, lut, n_centroids);
    }
}

/* score_internal, :566-593 */
QZO_EXPORT void qzo_pq_score_pairs(const uint8_t *codes, uint32_t dim, const uint32_t *bounds, uint32_t n_chunks,
                                   const float *centroids, int distance, int invert, const uint32_t *ia,
                                   const uint32_t *ib, uint64_t n, float *out) {
    for (uint64_t p = 0; p < n; p++) {
        const uint8_t *ci = codes + ia[p] * (uint64_t)n_chunks, *cj = codes + ib[p] * (uint64_t)n_chunks;
        float s = 0.0f;
        for (uint32_t ch = 0; ch < n_chunks; ch++) {
            uint32_t a = bounds[ch], b = bounds[ch + 1];
            s += metric_distance(distance, centroids + (uint64_t)ci[ch] * dim + a, centroids + (uint64_t)cj[ch] * dim + a,
                                 b - a);
        }
        out[p] = invert ? -s : s;
    }
}

typedef struct {
    const uint8_t *codes; uint64_t first, n; uint32_t n_chunks; const float *lut; uint32_t nc; int order; uint32_t k;
    int largest; uint32_t id_base; float *scores; uint32_t *ids; uint32_t got;
} pq_job;
static void *pq_job_run(void *p) {
    pq_job *j = (pq_job *)p;
    float *tmp = (float *)malloc(sizeof(float) * (j->n ? j->n : 1));
    qzo_pq_scan(j->codes, j->first, j->n, j->n_chunks, j->lut, j->nc, j->order, tmp);
    j->got = qzo_topk(tmp, j->n, j->k, j->largest, j->id_base + (uint32_t)j->first, j->scores, j->ids);
    free(tmp);
    return NULL;
}
QZO_EXPORT uint32_t qzo_pq_scan_topk_mt(const uint8_t *codes, uint64_t count, uint32_t n_chunks, const float *lut,
                                        uint32_t n_centroids, int order, uint32_t k, int largest, uint32_t id_base,
                                        int n_threads, float *out_scores, uint32_t *out_ids) {
    if (n_threads < 1) n_threads = 1;
    pq_job *jobs = (pq_job *)calloc(n_threads, sizeof(pq_job));
    pthread_t *th = (pthread_t *)calloc(n_threads, sizeof(pthread_t));
    float *ps = (float *)malloc(sizeof(float) * (size_t)n_threads * k);
    uint32_t *pi = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n_threads * k);
    for (int t = 0; t < n_threads; t++) {
        uint64_t a = count * t / n_threads, b = count * (t + 1) / n_threads;
        jobs[t] = (pq_job){codes, a, b - a, n_chunks, lut, n_centroids, order, k, largest, id_base,
                           ps + (size_t)t * k, pi + (size_t)t * k, 0};
        if (n_threads > 1) pthread_create(&th[t], NULL, pq_job_run, &jobs[t]);
        else pq_job_run(&jobs[t]);
    }
    uint64_t tot = 0;
    float *cs = (float *)malloc(sizeof(float) * (size_t)n_threads * k);
    uint32_t *ci = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n_threads * k);
    for (int t = 0; t < n_threads; t++) {
        if (n_threads > 1) pthread_join(th[t], NULL);
        memcpy(cs + tot, jobs[t].scores, sizeof(float) * jobs[t].got);
        memcpy(ci + tot, jobs[t].ids, sizeof(uint32_t) * jobs[t].got);
        tot += jobs[t].got;
    }
    uint32_t m = qzo_merge_topk(cs, ci, tot, k, largest, out_scores, out_ids);
    free(cs); free(ci); free(ps); free(pi); free(th); free(jobs);
    return m;
}

/* ---- k-means, kmeans.rs:7-167 -------------------------------------------------------------
 * data = [n][dim] sample sub-vectors of one chunk.  Deviations forced by the reference being
 * non-deterministic here: (1) an empty cluster is reseeded from thread_rng (kmeans.rs:111-118);
 * we draw the point id from SplitMix64(rng_seed) instead — statistical parity only; (2) f64
 * partial sums are merged in thread order over `max_threads` contiguous ranges (kmeans.rs:65-108),
 * reproduced exactly for a given max_threads.  The redundant final update_indexes (kmeans.rs:45)
 * has no effect on the returned centroids and is skipped.  Returns iterations run. */
QZO_EXPORT uint32_t qzo_kmeans(const float *data, uint64_t n, uint32_t dim, uint32_t n_centroids, uint32_t max_iter,
                               uint32_t max_threads, float accuracy, uint64_t rng_seed, float *centroids,
                               uint32_t *indexes_out) {
    if (max_threads < 1) max_threads = 1;
    memcpy(centroids, data, sizeof(float) * (size_t)n_centroids * dim); /* kmeans.rs:25 */
    uint32_t *idx = (uint32_t *)calloc(n ? n : 1, sizeof(uint32_t));
    size_t cd = (size_t)n_centroids * dim;
    double *acc = (double *)malloc(sizeof(double) * cd);
    double *tacc = (double *)malloc(sizeof(double) * cd);
    uint64_t *cnt = (uint64_t *)malloc(sizeof(uint64_t) * n_centroids);
    uint64_t rng_i = 0;
    uint32_t it = 0;
    for (; it < max_iter; it++) {
        for (uint64_t i = 0; i < n; i++) /* update_indexes, kmeans.rs:138-167 */
            idx[i] = nearest_centroid(data + i * dim, centroids, n_centroids, dim, dim, 0);
        /* update_centroids, kmeans.rs:49-136 */
        memset(acc, 0, sizeof(double) * cd);
        memset(cnt, 0, sizeof(uint64_t) * n_centroids);
        uint64_t chunk = n / max_threads;
        for (uint32_t t = 0; t < max_threads; t++) {
            uint64_t a = chunk * t, b = (t + 1 == max_threads) ? n : chunk * (t + 1);
            memset(tacc, 0, sizeof(double) * cd);
            for (uint64_t i = a; i < b; i++) {
                uint32_t c = idx[i];
                cnt[c]++;
                for (uint32_t d = 0; d < dim; d++) tacc[(size_t)c * dim + d] += (double)data[i * dim + d];
            }
            for (size_t j = 0; j < cd; j++) acc[j] += tacc[j];
        }
        for (uint32_t c = 0; c < n_centroids; c++) {
            if (cnt[c] == 0) {
                uint64_t pick = splitmix64_at(rng_seed, rng_i++) % n;
                for (uint32_t d = 0; d < dim; d++) acc[(size_t)c * dim + d] = (double)data[pick * dim + d];
            } else {
                double k = (double)cnt[c];
                for (uint32_t d = 0; d < dim; d++) acc[(size_t)c * dim + d] /= k;
            }
        }
        float diff = 0.0f;
        for (size_t j = 0; j < cd; j++) { /* kmeans.rs:125-134 */
            float nv = (float)acc[j];
            diff += fabsf(centroids[j] - nv);
            centroids[j] = nv;
        }
        if (diff < accuracy) { it++; break; }
    }
    if (indexes_out) {
        for (uint64_t i = 0; i < n; i++)
            indexes_out[i] = nearest_centroid(data + i * dim, centroids, n_centroids, dim, dim, 0);
    }
    free(idx); free(acc); free(tacc); free(cnt);
    return it;
}

/* find_centroids, encoded_vectors_pq.rs:278-342, on an explicit sorted sample id list (the
 * reference draws it from a randomly keyed Permutor, :300-307).  centroids out = [256][dim]. */
QZO_EXPORT void qzo_pq_find_centroids(const float *data, uint64_t count, uint32_t dim, const uint32_t *bounds,
                                      uint32_t n_chunks, const uint64_t *sample_ids, uint64_t n_sample,
                                      uint32_t n_centroids, uint32_t max_iter, uint32_t max_threads, float accuracy,
                                      uint64_t rng_seed, float *centroids) {
    if (count <= n_centroids) { /* :290-297 */
        memset(centroids, 0, sizeof(float) * (size_t)n_centroids * dim);
        memcpy(centroids, data, sizeof(float) * (size_t)count * dim);
        return;
    }
    for (uint32_t ch = 0; ch < n_chunks; ch++) {
        uint32_t a = bounds[ch], len = bounds[ch + 1] - bounds[ch];
        float *sub = (float *)malloc(sizeof(float) * (size_t)n_sample * len);
        for (uint64_t s = 0; s < n_sample; s++)
            memcpy(sub + s * len, data + sample_ids[s] * dim + a, sizeof(float) * len);
        float *cen = (float *)malloc(sizeof(float) * (size_t)n_centroids * len);
        qzo_kmeans(sub, n_sample, len, n_centroids, max_iter, max_threads, accuracy, rng_seed + ch, cen, NULL);
        for (uint32_t c = 0; c < n_centroids; c++) /* :336-338 */
            memcpy(centroids + (size_t)c * dim + a, cen + (size_t)c * len, sizeof(float) * len);
        free(sub);
        free(cen);
    }
}
