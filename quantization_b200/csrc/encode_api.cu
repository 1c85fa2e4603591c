// encode_api.cu — C-ABI batch encode entry points (u8 and binary).
//   qz_u8_encode[_dev]   EncodedVectorsU8::encode    (encoded_vectors_u8.rs:34-140, quantile.rs)
//   qz_bin_encode[_dev]  EncodedVectorsBin::encode   (encoded_vectors_binary.rs:165-208)
#include <string.h>

#include <algorithm>
#include <numeric>
#include <vector>

#include "qz_internal.h"

namespace qz {
qz_status bin_encode_rows(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, uint32_t row_bytes,
                          uint8_t *d_out);
qz_status dev_minmax(qz_ctx *ctx, const float *d_data, uint64_t n, uint32_t *d_mm, bool init, uint64_t base);
qz_status read_minmax(qz_ctx *ctx, const uint32_t *d_mm, float *mn, float *mx, uint64_t *zero_idx);
void fix_zero_sign(float *mn, float *mx, float first_zero);

// min / max of a device-resident corpus with the reference's first-seen-zero rule
static qz_status minmax_dev_resident(qz_ctx *ctx, const float *d_data, uint64_t n, uint32_t *d_mm, float *mn, float *mx) {
    uint64_t zi = ~0ull;
    QZ_TRY(dev_minmax(ctx, d_data, n, d_mm, true, 0));
    QZ_TRY(read_minmax(ctx, d_mm, mn, mx, &zi));
    if ((*mn == 0.0f || *mx == 0.0f) && zi != ~0ull) {
        float z = 0.0f;
        QZ_CUDA(ctx, cudaMemcpyAsync(&z, d_data + zi, sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        fix_zero_sign(mn, mx, z);
    }
    return QZ_OK;
}
qz_status dev_select2(qz_ctx *ctx, const float *d_vals, uint64_t n, uint64_t r0, uint64_t r1, float *v0, float *v1,
                      const uint64_t *d_ids, uint32_t dim);
qz_status launch_u8_encode(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, int distance, int invert,
                           float alpha, float offset, uint8_t *d_out);
qz_status launch_gather_rows(qz_ctx *ctx, const float *d_data, uint32_t dim, const uint64_t *d_ids, uint64_t n_ids,
                             float *d_out);

constexpr uint64_t kQuantileSample = 100000;  // QUANTILE_SAMPLE_SIZE, quantile.rs:3
constexpr uint64_t kChunkBytes = 256ull << 20;  // host<->device staging granularity

static inline bool stopped(const volatile int32_t *stop) { return stop && *stop != 0; }

static uint64_t gcd64(uint64_t a, uint64_t b) {
    while (b) {
        uint64_t t = a % b;
        a = b;
        b = t;
    }
    return a;
}

static inline uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// `n` distinct ids in [0,count), ascending: the caller's list, or (the reference draws them from a randomly keyed
// permutation, quantile.rs:32-34 / pq.rs:300-302 — permutation_iterator is a Feistel network too) the first n
// outputs of a 4-round Feistel permutation of [0, 2^bits) keyed by `seed`, cycle-walked into [0, count).
// quantization_b200/sharded.py::sample_ids is the same function (sharded encode needs the ids on every rank).
std::vector<uint64_t> pick_sample_ids(uint64_t count, uint64_t n, const uint64_t *ids, uint64_t n_ids, uint64_t seed) {
    std::vector<uint64_t> out;
    if (ids) {
        out.assign(ids, ids + n_ids);
    } else if (n >= count) {
        out.resize(count);
        std::iota(out.begin(), out.end(), 0ull);
        return out;
    } else {
        uint32_t bits = 2;
        while (bits < 64 && (1ull << bits) < count) bits += 2;  // even, so the halves are equal
        const uint32_t half = bits / 2;
        const uint64_t mask = (1ull << half) - 1;
        uint64_t keys[4];
        for (int r = 0; r < 4; r++) keys[r] = splitmix64(seed * 4 + r);
        auto perm = [&](uint64_t x) {
            do {
                uint64_t l = x >> half, r = x & mask;
                for (int k = 0; k < 4; k++) {
                    const uint64_t f = splitmix64(r ^ keys[k]) & mask;
                    const uint64_t t = l ^ f;
                    l = r;
                    r = t;
                }
                x = (l << half) | r;
            } while (x >= count);
            return x;
        };
        out.resize(n);
        if (count <= 256 * n) {  // mark a bitmap and read it back: the ids come out ascending without a sort
            std::vector<uint64_t> bm((count + 63) / 64, 0ull);
            for (uint64_t i = 0; i < n; i++) {
                const uint64_t x = perm(i);
                bm[x >> 6] |= 1ull << (x & 63);
            }
            size_t k = 0;
            for (size_t w = 0; w < bm.size(); w++)
                for (uint64_t m = bm[w]; m; m &= m - 1) out[k++] = (uint64_t)w * 64 + (uint64_t)__builtin_ctzll(m);
            return out;
        }
        for (uint64_t i = 0; i < n; i++) out[i] = perm(i);
    }
    std::sort(out.begin(), out.end());
    return out;
}

// alpha/offset/multiplier from (min,max): u8.rs:228-232, :119-128
static void u8_params(float mn, float mx, int distance, int invert, float *alpha, float *offset, float *multiplier) {
    const float a = (mx - mn) / 127.0f;
    float m;
    if (distance == QZ_DOT) m = a * a;
    else if (distance == QZ_L1) m = a;
    else {
        m = -2.0f * a;
        m = m * a;
    }
    *alpha = a;
    *offset = mn;
    *multiplier = invert ? -m : m;
}

// quantile.rs:53-70 on `len` sampled values resident on the device
// (d_ids != nullptr: the sample is rows d_ids[0 .. len / dim) of d_vals, read in place)
static qz_status quantile_interval(qz_ctx *ctx, const float *d_vals, uint64_t len, uint64_t slice_size, float q,
                                   bool *found, float *lo, float *hi, const uint64_t *d_ids = nullptr, uint32_t dim = 0) {
    *found = false;
    if (len < 4) return QZ_OK;
    const float fb = (float)slice_size * (1.0f - q) / 2.0f;  // vector count, not value count (quantile.rs:55)
    const uint64_t b = fb <= 0.0f ? 0 : (uint64_t)fb;
    uint64_t cut = std::min((len - 1) / 2, b);
    cut = std::max<uint64_t>(cut, 1);
    if (len < 2 * cut + 3) return QZ_OK;  // fewer than 2 values remain (quantile.rs:63-65)
    QZ_TRY(dev_select2(ctx, d_vals, len, cut + 1, len - cut - 1, lo, hi, d_ids, dim));
    *found = true;
    return QZ_OK;
}

}  // namespace qz

using namespace qz;

extern "C" {

qz_status qz_u8_encode_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, int32_t distance,
                           int32_t invert, const float *quantile, const uint64_t *sample_ids, uint64_t n_sample,
                           uint64_t seed, const volatile int32_t *stop, uint8_t *d_out, float *alpha, float *offset,
                           float *multiplier) {
    if (!ctx || !alpha || !offset || !multiplier) return fail(ctx, QZ_ARGUMENTS, "qz_u8_encode: NULL argument");
    if (distance < QZ_DOT || distance > QZ_L2 || dim == 0) return fail(ctx, QZ_ARGUMENTS, "qz_u8_encode: bad parameters");
    *alpha = *offset = *multiplier = 0.0f;
    if (count == 0) return QZ_OK;  // u8.rs:43-54
    if (!d_data || !d_out) return fail(ctx, QZ_ARGUMENTS, "qz_u8_encode: NULL data");
    std::lock_guard<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    if (stopped(stop)) return QZ_STOPPED;
    QZ_TRY(ensure_io(ctx, 256));
    uint32_t *d_mm = reinterpret_cast<uint32_t *>(ctx->d_io);
    float mn, mx;
    QZ_TRY(minmax_dev_resident(ctx, d_data, count * (uint64_t)dim, d_mm, &mn, &mx));
    if (quantile && count >= 127 && *quantile < 1.0f) {  // quantile.rs:27-29
        const uint64_t slice = std::min<uint64_t>(count, kQuantileSample);
        bool found = false;
        float lo = 0, hi = 0;
        if (!sample_ids && slice == count) {
            QZ_TRY(quantile_interval(ctx, d_data, count * (uint64_t)dim, slice, *quantile, &found, &lo, &hi));
        } else {
            std::vector<uint64_t> ids = pick_sample_ids(count, slice, sample_ids, n_sample, seed);
            for (uint64_t id : ids)
                if (id >= count) return fail(ctx, QZ_ARGUMENTS, "qz_u8_encode: sample id out of range");
            uint64_t *d_ids = nullptr;
            cudaError_t e = cudaMalloc((void **)&d_ids, ids.size() * sizeof(uint64_t));
            if (e == cudaSuccess)
                e = cudaMemcpyAsync(d_ids, ids.data(), ids.size() * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream);
            qz_status st = e == cudaSuccess ? QZ_OK : cuda_fail(ctx, e, "qz_u8_encode: sample alloc");
            if (st == QZ_OK)
                st = quantile_interval(ctx, d_data, ids.size() * (uint64_t)dim, ids.size(), *quantile, &found, &lo, &hi, d_ids, dim);
            cudaStreamSynchronize(ctx->stream);
            cudaFree(d_ids);
            if (st != QZ_OK) return st;
        }
        if (found) {
            mn = lo;
            mx = hi;
        }
    }
    u8_params(mn, mx, distance, invert, alpha, offset, multiplier);
    // encode in row chunks so the stop flag is polled between launches (u8.rs:74-76)
    const uint32_t row_bytes = qz_u8_row_bytes(dim);
    const uint64_t chunk = std::max<uint64_t>(1, kChunkBytes / ((uint64_t)dim * 4));
    for (uint64_t r = 0; r < count; r += chunk) {
        if (stopped(stop)) {
            cudaStreamSynchronize(ctx->stream);
            return QZ_STOPPED;
        }
        const uint64_t n = std::min(chunk, count - r);
        QZ_TRY(launch_u8_encode(ctx, d_data + r * dim, n, dim, distance, invert, *alpha, *offset, d_out + r * row_bytes));
    }
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return QZ_OK;
}

qz_status qz_u8_minmax_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, float *min_out,
                           float *max_out) {
    if (!ctx || !min_out || !max_out) return fail(ctx, QZ_ARGUMENTS, "qz_u8_minmax_dev: NULL argument");
    if (count && (!d_data || dim == 0)) return fail(ctx, QZ_ARGUMENTS, "qz_u8_minmax_dev: NULL data");
    std::lock_guard<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    QZ_TRY(ensure_io(ctx, 256));
    uint32_t *d_mm = reinterpret_cast<uint32_t *>(ctx->d_io);
    return minmax_dev_resident(ctx, d_data, count * (uint64_t)dim, d_mm, min_out, max_out);
}

qz_status qz_u8_quantile_dev(qz_ctx *ctx, const float *d_sample, uint64_t n_rows, uint32_t dim, uint64_t slice_size,
                             float quantile, int32_t *found, float *lo, float *hi) {
    if (!ctx || !found || !lo || !hi) return fail(ctx, QZ_ARGUMENTS, "qz_u8_quantile_dev: NULL argument");
    *found = 0;
    *lo = *hi = 0.0f;
    if (n_rows == 0 || dim == 0) return QZ_OK;
    if (!d_sample) return fail(ctx, QZ_ARGUMENTS, "qz_u8_quantile_dev: NULL sample");
    std::lock_guard<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    bool f = false;
    QZ_TRY(quantile_interval(ctx, d_sample, n_rows * (uint64_t)dim, slice_size, quantile, &f, lo, hi));
    *found = f ? 1 : 0;
    return QZ_OK;
}

qz_status qz_u8_encode_range_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, int32_t distance,
                                 int32_t invert, float min_value, float max_value, const volatile int32_t *stop,
                                 uint8_t *d_out, float *alpha, float *offset, float *multiplier) {
    if (!ctx || !alpha || !offset || !multiplier) return fail(ctx, QZ_ARGUMENTS, "qz_u8_encode_range_dev: NULL argument");
    if (distance < QZ_DOT || distance > QZ_L2 || dim == 0) return fail(ctx, QZ_ARGUMENTS, "qz_u8_encode_range_dev: bad parameters");
    if (count && (!d_data || !d_out)) return fail(ctx, QZ_ARGUMENTS, "qz_u8_encode_range_dev: NULL data");
    std::lock_guard<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    u8_params(min_value, max_value, distance, invert, alpha, offset, multiplier);
    const uint32_t row_bytes = qz_u8_row_bytes(dim);
    const uint64_t chunk = std::max<uint64_t>(1, kChunkBytes / ((uint64_t)dim * 4));
    for (uint64_t r = 0; r < count; r += chunk) {
        if (stopped(stop)) {
            cudaStreamSynchronize(ctx->stream);
            return QZ_STOPPED;
        }
        const uint64_t n = std::min(chunk, count - r);
        QZ_TRY(launch_u8_encode(ctx, d_data + r * dim, n, dim, distance, invert, *alpha, *offset, d_out + r * row_bytes));
    }
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return QZ_OK;
}

qz_status qz_u8_encode(qz_ctx *ctx, const float *data, uint64_t count, uint32_t dim, int32_t distance, int32_t invert,
                       const float *quantile, const uint64_t *sample_ids, uint64_t n_sample, uint64_t seed,
                       const volatile int32_t *stop, uint8_t *out_storage, float *alpha, float *offset,
                       float *multiplier) {
    if (!ctx || !alpha || !offset || !multiplier) return fail(ctx, QZ_ARGUMENTS, "qz_u8_encode: NULL argument");
    if (distance < QZ_DOT || distance > QZ_L2 || dim == 0) return fail(ctx, QZ_ARGUMENTS, "qz_u8_encode: bad parameters");
    *alpha = *offset = *multiplier = 0.0f;
    if (count == 0) return QZ_OK;
    if (!data || !out_storage) return fail(ctx, QZ_ARGUMENTS, "qz_u8_encode: NULL data");
    std::unique_lock<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint32_t row_bytes = qz_u8_row_bytes(dim);
    const uint64_t chunk = std::max<uint64_t>(1, kChunkBytes / ((uint64_t)dim * 4));
    const uint64_t crow = std::min(chunk, count);
    float *d_chunk = nullptr;
    uint8_t *d_enc = nullptr;
    uint32_t *d_mm = nullptr;
    float *d_sample = nullptr;
    qz_status st = QZ_OK;
    cudaError_t e = cudaMalloc((void **)&d_chunk, crow * (size_t)dim * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_enc, crow * (size_t)row_bytes + 16);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_mm, 256);
    if (e != cudaSuccess) st = cuda_fail(ctx, e, "qz_u8_encode: staging alloc");
    float mn = 0, mx = 0;
    // pass 1: min / max (u8.rs:57)
    for (uint64_t r = 0; r < count && st == QZ_OK; r += crow) {
        if (stopped(stop)) st = QZ_STOPPED;
        const uint64_t n = std::min(crow, count - r);
        if (st == QZ_OK) {
            e = cudaMemcpyAsync(d_chunk, data + r * dim, n * (size_t)dim * 4, cudaMemcpyHostToDevice, ctx->stream);
            if (e != cudaSuccess) st = cuda_fail(ctx, e, "qz_u8_encode: H2D");
        }
        if (st == QZ_OK) st = dev_minmax(ctx, d_chunk, n * (uint64_t)dim, d_mm, r == 0, r * (uint64_t)dim);
        if (st == QZ_OK) {
            e = cudaStreamSynchronize(ctx->stream);
            if (e != cudaSuccess) st = cuda_fail(ctx, e, "qz_u8_encode: minmax");
        }
    }
    uint64_t zero_idx = ~0ull;
    if (st == QZ_OK) st = read_minmax(ctx, d_mm, &mn, &mx, &zero_idx);
    if (st == QZ_OK && zero_idx != ~0ull) fix_zero_sign(&mn, &mx, data[zero_idx]);  // quantile.rs:5-19: first zero seen wins
    // pass 2: quantile cut on a sample (u8.rs:58-71)
    if (st == QZ_OK && quantile && count >= 127 && *quantile < 1.0f) {
        const uint64_t slice = std::min<uint64_t>(count, kQuantileSample);
        std::vector<uint64_t> ids = pick_sample_ids(count, slice, sample_ids, n_sample, seed);
        for (uint64_t id : ids)
            if (id >= count) st = fail(ctx, QZ_ARGUMENTS, "qz_u8_encode: sample id out of range");
        if (st == QZ_OK) {
            e = cudaMalloc((void **)&d_sample, ids.size() * (size_t)dim * 4);
            if (e != cudaSuccess) st = cuda_fail(ctx, e, "qz_u8_encode: sample alloc");
        }
        // copy maximal runs of consecutive ids
        for (size_t i = 0; i < ids.size() && st == QZ_OK;) {
            size_t j = i + 1;
            while (j < ids.size() && ids[j] == ids[j - 1] + 1) j++;
            e = cudaMemcpyAsync(d_sample + i * (size_t)dim, data + ids[i] * dim, (j - i) * (size_t)dim * 4,
                                cudaMemcpyHostToDevice, ctx->stream);
            if (e != cudaSuccess) st = cuda_fail(ctx, e, "qz_u8_encode: sample H2D");
            i = j;
        }
        bool found = false;
        float lo = 0, hi = 0;
        if (st == QZ_OK) st = quantile_interval(ctx, d_sample, ids.size() * (uint64_t)dim, ids.size(), *quantile, &found, &lo, &hi);
        if (st == QZ_OK && found) {
            mn = lo;
            mx = hi;
        }
    }
    if (st == QZ_OK) u8_params(mn, mx, distance, invert, alpha, offset, multiplier);
    // pass 3: codes + per-vector header (u8.rs:73-118)
    for (uint64_t r = 0; r < count && st == QZ_OK; r += crow) {
        if (stopped(stop)) {
            st = QZ_STOPPED;
            break;
        }
        const uint64_t n = std::min(crow, count - r);
        e = cudaMemcpyAsync(d_chunk, data + r * dim, n * (size_t)dim * 4, cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) st = cuda_fail(ctx, e, "qz_u8_encode: H2D");
        if (st == QZ_OK) st = launch_u8_encode(ctx, d_chunk, n, dim, distance, invert, *alpha, *offset, d_enc);
        if (st == QZ_OK) {
            e = cudaMemcpyAsync(out_storage + r * row_bytes, d_enc, n * (size_t)row_bytes, cudaMemcpyDeviceToHost, ctx->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
            if (e != cudaSuccess) st = cuda_fail(ctx, e, "qz_u8_encode: D2H");
        }
    }
    cudaStreamSynchronize(ctx->stream);
    cudaFree(d_chunk);
    cudaFree(d_enc);
    cudaFree(d_mm);
    cudaFree(d_sample);
    return st;
}

qz_status qz_bin_encode_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, int32_t store_bytes,
                            const volatile int32_t *stop, uint8_t *d_out) {
    if (!ctx) return QZ_ARGUMENTS;
    if (store_bytes != 1 && store_bytes != 16) return fail(ctx, QZ_ARGUMENTS, "qz_bin_encode: store_bytes must be 1 or 16");
    const uint32_t rb = qz_bin_row_bytes(dim, store_bytes);
    if (count == 0 || rb == 0) return stopped(stop) ? QZ_STOPPED : QZ_OK;
    if (!d_data || !d_out) return fail(ctx, QZ_ARGUMENTS, "qz_bin_encode: NULL data");
    std::lock_guard<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint64_t chunk = std::max<uint64_t>(1, kChunkBytes / ((uint64_t)dim * 4));
    for (uint64_t r = 0; r < count; r += chunk) {
        if (stopped(stop)) {
            cudaStreamSynchronize(ctx->stream);
            return QZ_STOPPED;
        }
        const uint64_t n = std::min(chunk, count - r);
        QZ_TRY(bin_encode_rows(ctx, d_data + r * dim, n, dim, rb, d_out + r * rb));
    }
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return QZ_OK;
}

qz_status qz_bin_encode(qz_ctx *ctx, const float *data, uint64_t count, uint32_t dim, int32_t store_bytes,
                        const volatile int32_t *stop, uint8_t *out_storage) {
    if (!ctx) return QZ_ARGUMENTS;
    if (store_bytes != 1 && store_bytes != 16) return fail(ctx, QZ_ARGUMENTS, "qz_bin_encode: store_bytes must be 1 or 16");
    const uint32_t rb = qz_bin_row_bytes(dim, store_bytes);
    if (count == 0 || rb == 0) return stopped(stop) ? QZ_STOPPED : QZ_OK;
    if (!data || !out_storage) return fail(ctx, QZ_ARGUMENTS, "qz_bin_encode: NULL data");
    std::unique_lock<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint64_t chunk = std::max<uint64_t>(1, kChunkBytes / ((uint64_t)dim * 4));
    const uint64_t crow = std::min(chunk, count);
    float *d_chunk = nullptr;
    uint8_t *d_enc = nullptr;
    qz_status st = QZ_OK;
    cudaError_t e = cudaMalloc((void **)&d_chunk, crow * (size_t)dim * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_enc, crow * (size_t)rb + 16);
    if (e != cudaSuccess) st = cuda_fail(ctx, e, "qz_bin_encode: staging alloc");
    for (uint64_t r = 0; r < count && st == QZ_OK; r += crow) {
        if (stopped(stop)) {
            st = QZ_STOPPED;
            break;
        }
        const uint64_t n = std::min(crow, count - r);
        e = cudaMemcpyAsync(d_chunk, data + r * dim, n * (size_t)dim * 4, cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) st = cuda_fail(ctx, e, "qz_bin_encode: H2D");
        if (st == QZ_OK) st = bin_encode_rows(ctx, d_chunk, n, dim, rb, d_enc);
        if (st == QZ_OK) {
            e = cudaMemcpyAsync(out_storage + r * rb, d_enc, n * (size_t)rb, cudaMemcpyDeviceToHost, ctx->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
            if (e != cudaSuccess) st = cuda_fail(ctx, e, "qz_bin_encode: D2H");
        }
    }
    cudaStreamSynchronize(ctx->stream);
    cudaFree(d_chunk);
    cudaFree(d_enc);
    return st;
}

}  // extern "C"
