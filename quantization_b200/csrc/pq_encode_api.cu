// pq_encode_api.cu — C-ABI of the PQ batch-encode step.
//   qz_pq_train[_dev]   find_centroids + kmeans   (encoded_vectors_pq.rs:278-342, kmeans.rs:7-47)
//   qz_pq_assign[_dev]  encode_storage / encode_vector (encoded_vectors_pq.rs:136-265)
#include <string.h>

#include <algorithm>
#include <vector>

#include "qz_internal.h"

namespace qz {
std::vector<uint64_t> pick_sample_ids(uint64_t count, uint64_t n, const uint64_t *ids, uint64_t n_ids, uint64_t seed);
qz_status launch_gather_rows(qz_ctx *ctx, const float *d_data, uint32_t dim, const uint64_t *d_ids, uint64_t n_ids,
                             float *d_out);
qz_status launch_pq_assign(qz_ctx *ctx, const float *d_data, uint64_t n, uint32_t dim, const float *d_ct,
                           const uint32_t *d_bounds, const std::vector<uint32_t> &bounds, const uint8_t *d_done,
                           uint8_t *d_codes, uint8_t *d_idx);
qz_status launch_kmeans_update(qz_ctx *ctx, const float *d_sample, uint64_t n, uint32_t dim, float *d_ct,
                               const uint32_t *d_bounds, const std::vector<uint32_t> &bounds, const uint8_t *d_idx,
                               uint8_t *d_done, uint64_t seed, uint32_t iter, float accuracy);

constexpr uint64_t kKmeansSample = 10000;  // KMEANS_SAMPLE_SIZE, pq.rs:22
constexpr uint32_t kNCentroids = 256;      // CENTROIDS_COUNT, pq.rs:25
constexpr uint64_t kPqChunkBytes = 256ull << 20;

static inline bool stopped(const volatile int32_t *stop) { return stop && *stop != 0; }

// get_vector_division, pq.rs:116-121
static std::vector<uint32_t> division(uint32_t dim, uint32_t chunk_size) {
    std::vector<uint32_t> b;
    for (uint32_t i = 0; i < dim; i += chunk_size) b.push_back(i);
    b.push_back(dim);
    return b;
}

// [256][dim] row-major  <->  chunk-major ct[bounds[ch]*256 + c*len + t]
static void to_chunk_major(const float *cen, uint32_t dim, const std::vector<uint32_t> &b, std::vector<float> &ct) {
    ct.resize((size_t)dim * kNCentroids);
    for (size_t ch = 0; ch + 1 < b.size(); ch++) {
        const uint32_t a = b[ch], len = b[ch + 1] - a;
        for (uint32_t c = 0; c < kNCentroids; c++)
            memcpy(&ct[(size_t)a * kNCentroids + (size_t)c * len], cen + (size_t)c * dim + a, len * sizeof(float));
    }
}
static void from_chunk_major(const std::vector<float> &ct, uint32_t dim, const std::vector<uint32_t> &b, float *cen) {
    for (size_t ch = 0; ch + 1 < b.size(); ch++) {
        const uint32_t a = b[ch], len = b[ch + 1] - a;
        for (uint32_t c = 0; c < kNCentroids; c++)
            memcpy(cen + (size_t)c * dim + a, &ct[(size_t)a * kNCentroids + (size_t)c * len], len * sizeof(float));
    }
}

struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 16); }
    template <class T>
    T *as() { return reinterpret_cast<T *>(p); }
};

// k-means on a device-resident sample [n][dim]; result [256][dim] on the host
static qz_status train_on_sample(qz_ctx *ctx, const float *d_sample, uint64_t n, uint32_t dim,
                                 const std::vector<uint32_t> &bounds, uint64_t seed, uint32_t max_iter, float accuracy,
                                 const volatile int32_t *stop, float *centroids, uint32_t *iterations) {
    const uint32_t n_chunks = (uint32_t)bounds.size() - 1;
    // centroids_0 = the first 256 sample points (kmeans.rs:25)
    std::vector<float> init((size_t)kNCentroids * dim), ct;
    QZ_CUDA(ctx, cudaMemcpyAsync(init.data(), d_sample, init.size() * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    to_chunk_major(init.data(), dim, bounds, ct);
    DevBuf d_ct, d_bounds, d_idx, d_done;
    cudaError_t e = d_ct.alloc(ct.size() * sizeof(float));
    if (e == cudaSuccess) e = d_bounds.alloc(bounds.size() * sizeof(uint32_t));
    if (e == cudaSuccess) e = d_idx.alloc((size_t)n_chunks * n);
    if (e == cudaSuccess) e = d_done.alloc(n_chunks);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_ct.p, ct.data(), ct.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(d_bounds.p, bounds.data(), bounds.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_done.p, 0, n_chunks, ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, e, "qz_pq_train: alloc");
    std::vector<uint8_t> done(n_chunks, 0);
    std::vector<uint32_t> iters(n_chunks, 0);
    for (uint32_t it = 0; it < max_iter; it++) {
        if (stopped(stop)) {  // kmeans.rs:29-31
            cudaStreamSynchronize(ctx->stream);
            return QZ_STOPPED;
        }
        QZ_TRY(launch_pq_assign(ctx, d_sample, n, dim, d_ct.as<float>(), d_bounds.as<uint32_t>(), bounds,
                                d_done.as<uint8_t>(), nullptr, d_idx.as<uint8_t>()));
        QZ_TRY(launch_kmeans_update(ctx, d_sample, n, dim, d_ct.as<float>(), d_bounds.as<uint32_t>(), bounds,
                                    d_idx.as<uint8_t>(), d_done.as<uint8_t>(), seed, it, accuracy));
        for (uint32_t ch = 0; ch < n_chunks; ch++)
            if (!done[ch]) iters[ch] = it + 1;
        QZ_CUDA(ctx, cudaMemcpyAsync(done.data(), d_done.p, n_chunks, cudaMemcpyDeviceToHost, ctx->stream));
        QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (std::all_of(done.begin(), done.end(), [](uint8_t d) { return d != 0; })) break;
    }
    QZ_CUDA(ctx, cudaMemcpyAsync(ct.data(), d_ct.p, ct.size() * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    from_chunk_major(ct, dim, bounds, centroids);
    if (iterations) memcpy(iterations, iters.data(), n_chunks * sizeof(uint32_t));
    return QZ_OK;
}

static qz_status train_common(qz_ctx *ctx, const float *data, bool on_device, uint64_t count, uint32_t dim,
                              uint32_t chunk_size, const uint64_t *sample_ids, uint64_t n_sample, uint64_t seed,
                              uint32_t max_iter, float accuracy, const volatile int32_t *stop, float *centroids,
                              uint32_t *iterations) {
    if (!ctx || !centroids) return fail(ctx, QZ_ARGUMENTS, "qz_pq_train: NULL argument");
    if (chunk_size == 0 || dim == 0) return fail(ctx, QZ_ARGUMENTS, "qz_pq_train: dim and chunk_size must be > 0");
    if (count && !data) return fail(ctx, QZ_ARGUMENTS, "qz_pq_train: NULL data");
    std::lock_guard<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const std::vector<uint32_t> bounds = division(dim, chunk_size);
    if (iterations) memset(iterations, 0, (bounds.size() - 1) * sizeof(uint32_t));
    if (count <= kNCentroids) {  // pq.rs:290-297: centroids are the points, the rest zeros
        memset(centroids, 0, (size_t)kNCentroids * dim * sizeof(float));
        if (count) {
            QZ_CUDA(ctx, cudaMemcpyAsync(centroids, data, count * (size_t)dim * sizeof(float),
                                         on_device ? cudaMemcpyDeviceToHost : cudaMemcpyHostToHost, ctx->stream));
            QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        }
        return QZ_OK;
    }
    const uint64_t ns = std::min<uint64_t>(kKmeansSample, count);  // pq.rs:286
    std::vector<uint64_t> ids = pick_sample_ids(count, ns, sample_ids, n_sample, seed);
    if (ids.size() < kNCentroids) return fail(ctx, QZ_ARGUMENTS, "qz_pq_train: need at least 256 sample ids");
    for (uint64_t id : ids)
        if (id >= count) return fail(ctx, QZ_ARGUMENTS, "qz_pq_train: sample id out of range");
    if (stopped(stop)) return QZ_STOPPED;  // pq.rs:303-305
    DevBuf d_sample, d_ids;
    cudaError_t e = d_sample.alloc(ids.size() * (size_t)dim * sizeof(float));
    if (e != cudaSuccess) return cuda_fail(ctx, e, "qz_pq_train: sample alloc");
    if (on_device) {
        e = d_ids.alloc(ids.size() * sizeof(uint64_t));
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(d_ids.p, ids.data(), ids.size() * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) return cuda_fail(ctx, e, "qz_pq_train: ids");
        QZ_TRY(launch_gather_rows(ctx, data, dim, d_ids.as<uint64_t>(), ids.size(), d_sample.as<float>()));
    } else {
        for (size_t i = 0; i < ids.size();) {  // copy maximal runs of consecutive ids
            size_t j = i + 1;
            while (j < ids.size() && ids[j] == ids[j - 1] + 1) j++;
            e = cudaMemcpyAsync(d_sample.as<float>() + i * (size_t)dim, data + ids[i] * dim, (j - i) * (size_t)dim * 4,
                                cudaMemcpyHostToDevice, ctx->stream);
            if (e != cudaSuccess) return cuda_fail(ctx, e, "qz_pq_train: sample H2D");
            i = j;
        }
    }
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return train_on_sample(ctx, d_sample.as<float>(), ids.size(), dim, bounds, seed, max_iter, accuracy, stop, centroids,
                           iterations);
}

}  // namespace qz

using namespace qz;

extern "C" {

qz_status qz_pq_train(qz_ctx *ctx, const float *data, uint64_t count, uint32_t dim, uint32_t chunk_size,
                      const uint64_t *sample_ids, uint64_t n_sample, uint64_t seed, uint32_t max_iterations,
                      float accuracy, const volatile int32_t *stop, float *centroids, uint32_t *iterations) {
    return train_common(ctx, data, false, count, dim, chunk_size, sample_ids, n_sample, seed, max_iterations, accuracy,
                        stop, centroids, iterations);
}

qz_status qz_pq_train_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, uint32_t chunk_size,
                          const uint64_t *sample_ids, uint64_t n_sample, uint64_t seed, uint32_t max_iterations,
                          float accuracy, const volatile int32_t *stop, float *centroids, uint32_t *iterations) {
    return train_common(ctx, d_data, true, count, dim, chunk_size, sample_ids, n_sample, seed, max_iterations, accuracy,
                        stop, centroids, iterations);
}

static qz_status assign_common(qz_ctx *ctx, const float *data, bool on_device, uint64_t count, uint32_t dim,
                               uint32_t chunk_size, const float *centroids, const volatile int32_t *stop,
                               uint8_t *codes) {
    if (!ctx) return QZ_ARGUMENTS;
    if (chunk_size == 0 || dim == 0) return fail(ctx, QZ_ARGUMENTS, "qz_pq_assign: dim and chunk_size must be > 0");
    if (count == 0) return stopped(stop) ? QZ_STOPPED : QZ_OK;
    if (!data || !centroids || !codes) return fail(ctx, QZ_ARGUMENTS, "qz_pq_assign: NULL argument");
    std::lock_guard<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    const std::vector<uint32_t> bounds = division(dim, chunk_size);
    const uint32_t n_chunks = (uint32_t)bounds.size() - 1;
    std::vector<float> ct;
    to_chunk_major(centroids, dim, bounds, ct);
    DevBuf d_ct, d_bounds, d_chunk, d_codes;
    cudaError_t e = d_ct.alloc(ct.size() * sizeof(float));
    if (e == cudaSuccess) e = d_bounds.alloc(bounds.size() * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_ct.p, ct.data(), ct.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(d_bounds.p, bounds.data(), bounds.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, e, "qz_pq_assign: alloc");
    const uint64_t crow = std::min<uint64_t>(count, std::max<uint64_t>(1, kPqChunkBytes / ((uint64_t)dim * 4)));
    if (!on_device) {
        e = d_chunk.alloc(crow * (size_t)dim * 4);
        if (e == cudaSuccess) e = d_codes.alloc(crow * (size_t)n_chunks);
        if (e != cudaSuccess) return cuda_fail(ctx, e, "qz_pq_assign: staging alloc");
    }
    for (uint64_t r = 0; r < count; r += crow) {
        if (stopped(stop)) {  // pq.rs:198-200
            cudaStreamSynchronize(ctx->stream);
            return QZ_STOPPED;
        }
        const uint64_t n = std::min(crow, count - r);
        const float *src = data + r * dim;
        uint8_t *dst = codes + r * n_chunks;
        if (!on_device) {
            QZ_CUDA(ctx, cudaMemcpyAsync(d_chunk.p, src, n * (size_t)dim * 4, cudaMemcpyHostToDevice, ctx->stream));
            src = d_chunk.as<float>();
            dst = d_codes.as<uint8_t>();
        }
        QZ_TRY(launch_pq_assign(ctx, src, n, dim, d_ct.as<float>(), d_bounds.as<uint32_t>(), bounds, nullptr, dst, nullptr));
        if (!on_device) {
            QZ_CUDA(ctx, cudaMemcpyAsync(codes + r * n_chunks, d_codes.p, n * (size_t)n_chunks, cudaMemcpyDeviceToHost, ctx->stream));
            QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        }
    }
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return stopped(stop) ? QZ_STOPPED : QZ_OK;  // pq.rs:95-106 polls once more at the end
}

qz_status qz_pq_assign(qz_ctx *ctx, const float *data, uint64_t count, uint32_t dim, uint32_t chunk_size,
                       const float *centroids, const volatile int32_t *stop, uint8_t *codes) {
    return assign_common(ctx, data, false, count, dim, chunk_size, centroids, stop, codes);
}

qz_status qz_pq_assign_dev(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, uint32_t chunk_size,
                           const float *centroids, const volatile int32_t *stop, uint8_t *d_codes) {
    return assign_common(ctx, d_data, true, count, dim, chunk_size, centroids, stop, d_codes);
}

}  // extern "C"
