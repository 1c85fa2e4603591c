// pq.cu — product-quantizer scoring (EncodedVectorsPQ, quantization/src/encoded_vectors_pq.rs)
//   pq_lut_kernel          encode_query: lut[ch*256+c] = +-distance(q[range], centroid c[range])   (:525-547)
//   pq_lut_permute_kernel  re-lays one query's LUT so that every lane of a warp gathers from its
//                          own shared-memory bank
//   pq_scan_kernel         score_point for every row: sum of LUT gathers                             (:405-494)
//   pq_points_kernel       score_point on an id list in the exact x86 (SSE) summation order       (:405-440)
//   pq_pairs_kernel        score_internal                                                            (:566-593)
//
// Centroids live on the device "chunk-major": ct[bounds[ch]*256 + c*len + t] = centroids[c][bounds[ch]+t],
// so one chunk's 256 sub-centroids are contiguous.
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "qz_internal.h"
#include "common.cuh"

namespace qz {

bool pq_scan2_launch(qz_ctx *ctx, const uint8_t *d_codes, uint64_t count, uint32_t n_chunks, uint32_t n_slices,
                     const float *d_perm, float *d_scores, uint32_t k, int largest, uint32_t id_base, uint32_t *n_lists,
                     const ScanFinal *fin, cudaError_t *err);

constexpr int kPqCentroids = 256;     // CENTROIDS_COUNT, pq.rs:25
constexpr int kSliceChunks = 128;     // chunks per LUT slice: 32 lanes x 4 code bytes
constexpr int kSliceFloats = kSliceChunks * kPqCentroids;  // 32768 floats = 128 KiB
constexpr int kPqWarps = 16;

// DistanceType::distance, encoded_vectors.rs:37-45: sequential f32 sum from 0.0, no FMA
__device__ __forceinline__ float metric_term(int distance, float a, float b) {
    if (distance == kDot) return __fmul_rn(a, b);
    const float d = __fsub_rn(a, b);
    return distance == kL1 ? fabsf(d) : __fmul_rn(d, d);
}

__global__ void __launch_bounds__(256) pq_lut_kernel(const float *queries, uint32_t dim, const uint32_t *bounds,
                                                     uint32_t n_chunks, const float *ct, int distance, int invert,
                                                     float *lut_out) {
    const uint32_t ch = blockIdx.x, c = threadIdx.x;
    const float *q = queries + (size_t)blockIdx.y * dim;
    float *lut = lut_out + (size_t)blockIdx.y * n_chunks * kPqCentroids;
    const uint32_t a = bounds[ch], len = bounds[ch + 1] - a;
    const float *cen = ct + (size_t)a * kPqCentroids + (size_t)c * len;
    float s = 0.0f;
    for (uint32_t t = 0; t < len; t++) s = __fadd_rn(s, metric_term(distance, q[a + t], cen[t]));
    lut[(size_t)ch * kPqCentroids + c] = invert ? -s : s;
}

// perm[slice][(b >> 1)*16384 + c*64 + (b & 1)*32 + w] = lut[(slice*128 + 4*w + b)*256 + c]   (0 beyond n_chunks):
// lane w's four gathers (b = 0..3) fall in bank w whatever the codes, and a code's byte offset (c * 256) leaves the
// low byte free for the lane's own offset, so one byte permute forms the whole variable part of the address
__global__ void __launch_bounds__(256) pq_lut_permute_kernel(const float *lut, uint32_t n_chunks, float *perm) {
    const uint32_t slice = blockIdx.y;
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < (uint32_t)kSliceFloats;
         idx += gridDim.x * blockDim.x) {
        // read-coalesced over c; the scattered side is the (cheap, L2-resident) store
        const uint32_t cl = idx / kPqCentroids, c = idx % kPqCentroids;
        const uint32_t ch = slice * kSliceChunks + cl;
        const float v = ch < n_chunks ? lut[(size_t)ch * kPqCentroids + c] : 0.0f;
        const uint32_t w = cl >> 2, b = cl & 3;
        perm[(size_t)slice * kSliceFloats + (b >> 1) * 16384 + c * 64 + (b & 1) * 32 + w] = v;
    }
}

struct PqScanParams {
    const uint8_t *codes;
    uint64_t count;
    uint32_t n_chunks;
    const float *perm;  // [n_slices][kSliceFloats]
    float *partial;     // [n_slices][count]
    int aligned;        // rows are whole, aligned 32-bit words
};

// grid = (row CTAs, slices).  Each CTA keeps its 128 KiB LUT slice in shared memory; a warp scores
// 16 rows per step: lane w owns code word w of the slice (4 chunks), so its four gathers
// (layout of pq_lut_permute_kernel) always fall in bank w: conflict free for any codes.
__global__ void __launch_bounds__(kPqWarps * 32, 1) pq_scan_kernel(PqScanParams p) {
    extern __shared__ __align__(128) float slut[];
    __shared__ __align__(8) uint64_t bar;
    const uint32_t slice = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar, kSliceFloats * 4);
        const char *src = reinterpret_cast<const char *>(p.perm + (size_t)slice * kSliceFloats);
        for (int i = 0; i < 4; i++)  // 4 x 32 KiB bulk copies
            bulk_g2s(reinterpret_cast<char *>(slut) + i * 32768, src + i * 32768, 32768, &bar);
    }
    mbar_wait(&bar, 0);

    const uint32_t ch0 = slice * kSliceChunks + 4 * lane;  // first chunk of this lane's code word
    const bool upper16 = lane & 16, upper8 = lane & 8, upper4 = lane & 4, upper2 = lane & 2;
    const float *my = slut + lane;
    float *out = p.partial + (size_t)slice * p.count;
    const uint64_t n_batches = (p.count + 15) / 16;
    const uint64_t stride = (uint64_t)gridDim.x * kPqWarps;
    for (uint64_t bt = (uint64_t)blockIdx.x * kPqWarps + warp; bt < n_batches; bt += stride) {
        const uint64_t row0 = bt * 16;
        uint32_t cw[16];
#pragma unroll
        for (int r = 0; r < 16; r++) {
            uint64_t row = row0 + r;
            if (row >= p.count) row = p.count - 1;
            const uint8_t *rp = p.codes + row * p.n_chunks;
            if (p.aligned) {
                cw[r] = ch0 < p.n_chunks ? __ldg(reinterpret_cast<const uint32_t *>(rp) + (ch0 >> 2)) : 0u;
            } else {
                uint32_t v = 0;
#pragma unroll
                for (int b = 0; b < 4; b++)
                    if (ch0 + b < p.n_chunks) v |= (uint32_t)__ldg(rp + ch0 + b) << (8 * b);
                cw[r] = v;
            }
        }
        float acc[16];
#pragma unroll
        for (int r = 0; r < 16; r++) {
            // chunks past n_chunks gather the zero padding of the permuted LUT
            float s = my[(cw[r] & 0xFFu) * 64];
            s = __fadd_rn(s, my[((cw[r] >> 8) & 0xFFu) * 64 + 32]);
            s = __fadd_rn(s, my[16384 + ((cw[r] >> 16) & 0xFFu) * 64]);
            s = __fadd_rn(s, my[16384 + (cw[r] >> 24) * 64 + 32]);
            acc[r] = s;
        }
#pragma unroll
        for (int i = 0; i < 8; i++) {
            float keep = upper16 ? acc[i + 8] : acc[i], send = upper16 ? acc[i] : acc[i + 8];
            acc[i] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, 16));
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {
            float keep = upper8 ? acc[i + 4] : acc[i], send = upper8 ? acc[i] : acc[i + 4];
            acc[i] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, 8));
        }
#pragma unroll
        for (int i = 0; i < 2; i++) {
            float keep = upper4 ? acc[i + 2] : acc[i], send = upper4 ? acc[i] : acc[i + 2];
            acc[i] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, 4));
        }
        {
            float keep = upper2 ? acc[1] : acc[0], send = upper2 ? acc[0] : acc[1];
            acc[0] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, 2));
        }
        const float total = __fadd_rn(acc[0], __shfl_xor_sync(0xffffffffu, acc[0], 1));
        const uint64_t row = row0 + (lane >> 1);
        if (!(lane & 1) && row < p.count) out[row] = total;
    }
}

// exact x86 order (score_point_sse, pq.rs:405-440): four interleaved lanes, (l0+l2)+(l1+l3), tail
__global__ void __launch_bounds__(128) pq_points_kernel(const uint8_t *codes, uint32_t n_chunks, const float *lut,
                                                        const uint32_t *rows, uint64_t n, float *out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t *c = codes + (uint64_t)rows[i] * n_chunks;
    float l0 = 0.0f, l1 = 0.0f, l2 = 0.0f, l3 = 0.0f;
    const float *p = lut;
    const uint32_t full = n_chunks / 4;
    for (uint32_t t = 0; t < full; t++) {
        l0 = __fadd_rn(l0, p[c[0]]);
        l1 = __fadd_rn(l1, p[256 + c[1]]);
        l2 = __fadd_rn(l2, p[512 + c[2]]);
        l3 = __fadd_rn(l3, p[768 + c[3]]);
        c += 4;
        p += 1024;
    }
    float sum = __fadd_rn(__fadd_rn(l0, l2), __fadd_rn(l1, l3));
    for (uint32_t t = 0; t < n_chunks % 4; t++) {
        sum = __fadd_rn(sum, p[c[0]]);
        c++;
        p += 256;
    }
    out[i] = sum;
}

__global__ void __launch_bounds__(128) pq_pairs_kernel(const uint8_t *codes, uint32_t n_chunks, const uint32_t *bounds,
                                                       const float *ct, int distance, int invert, const uint32_t *ra,
                                                       const uint32_t *rb, uint64_t n, float *out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t *ci = codes + (uint64_t)ra[i] * n_chunks, *cj = codes + (uint64_t)rb[i] * n_chunks;
    float s = 0.0f;
    for (uint32_t ch = 0; ch < n_chunks; ch++) {
        const uint32_t a = bounds[ch], len = bounds[ch + 1] - a;
        const float *x = ct + (size_t)a * kPqCentroids + (size_t)ci[ch] * len;
        const float *y = ct + (size_t)a * kPqCentroids + (size_t)cj[ch] * len;
        float d = 0.0f;
        for (uint32_t t = 0; t < len; t++) d = __fadd_rn(d, metric_term(distance, x[t], y[t]));
        s = __fadd_rn(s, d);
    }
    out[i] = invert ? -s : s;
}

// ---------------------------------------------------------------------------------------------
static qz_status check_launch(qz_ctx *ctx, const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(ctx, e, what);
    ctx->launches++;
    return QZ_OK;
}

qz_status pq_encode_query(qz_index *ix, const float *d_queries, uint32_t nq, void *d_out) {
    if (nq == 0 || ix->n_chunks == 0) return QZ_OK;
    dim3 grid(ix->n_chunks, nq);
    pq_lut_kernel<<<grid, 256, 0, ix->ctx->stream>>>(d_queries, ix->dim, ix->d_bounds, ix->n_chunks, ix->d_centroids,
                                                     ix->distance, ix->invert, (float *)d_out);
    return check_launch(ix->ctx, "pq_lut_kernel");
}

qz_status pq_scan(qz_index *ix, const void *d_lut, float *d_scores, uint32_t k, int largest, uint32_t *n_lists,
                  uint32_t *Kp_out, const ScanFinal *fin) {
    if (fin) *fin->fused = false;
    qz_ctx *ctx = ix->ctx;
    QZ_TRY(ensure_smem_attr(ctx, (const void *)pq_scan_kernel, kSliceFloats * 4));
    const uint32_t slices = ix->pq_slices;
    if (slices == 0) {  // dim 0: every score is the empty sum
        QZ_TRY(ensure_io(ctx, ix->count * sizeof(float)));
        QZ_CUDA(ctx, cudaMemsetAsync(ctx->d_io, 0, ix->count * sizeof(float), ctx->stream));
        return launch_scores_topk(ctx, (const float *)ctx->d_io, 1, 0, ix->count, d_scores, k, largest, ix->id_base,
                                  n_lists, Kp_out);
    }
    dim3 pg(32, slices);  // 4 LUT entries per thread (the kernel is latency-bound: 8 blocks per slice took 5 us)
    pq_lut_permute_kernel<<<pg, 256, 0, ctx->stream>>>((const float *)d_lut, ix->n_chunks, ix->d_lut_perm);
    QZ_TRY(check_launch(ctx, "pq_lut_permute_kernel"));
    // streaming kernel (pq_scan2.cu): k <= 128 is reduced inside the kernel; otherwise the final scores go to
    // partial[0] and the generic top-k pass runs over that one part
    const bool env_v1 = tune_flag("QZ_PQ_V1");
    if (!env_v1) {
        const bool fused = k > 0 && k <= 128;
        float *sc = fused ? d_scores : (d_scores ? d_scores : ix->d_pq_partial);
        cudaError_t e2 = cudaSuccess;
        uint32_t lists = 0;
        if (pq_scan2_launch(ctx, ix->d_storage, ix->count, ix->n_chunks, slices, ix->d_lut_perm, sc, fused ? k : 0, largest,
                            ix->id_base, &lists, fused ? fin : nullptr, &e2)) {
            if (e2 != cudaSuccess) return cuda_fail(ctx, e2, "pq_scan2_kernel");
            ctx->launches++;
            if (fused) {
                if (n_lists) *n_lists = lists;
                if (Kp_out) *Kp_out = 128;
                return QZ_OK;
            }
            if (!k) return QZ_OK;
            return launch_scores_topk(ctx, sc, 1, ix->count, ix->count, nullptr, k, largest, ix->id_base, n_lists, Kp_out);
        }
    }
    PqScanParams p{ix->d_storage, ix->count, ix->n_chunks, ix->d_lut_perm, ix->d_pq_partial,
                   (ix->n_chunks % 4 == 0) ? 1 : 0};
    uint32_t gx = (uint32_t)ctx->sm_count / slices;
    if (gx == 0) gx = 1;
    const uint64_t n_batches = (ix->count + 15) / 16;
    const uint64_t need = (n_batches + kPqWarps - 1) / kPqWarps;
    if (gx > need) gx = (uint32_t)need;
    dim3 grid(gx, slices);
    timing_begin(ctx);
    pq_scan_kernel<<<grid, kPqWarps * 32, kSliceFloats * 4, ctx->stream>>>(p);
    timing_end(ctx);
    QZ_TRY(check_launch(ctx, "pq_scan_kernel"));
    // ordered sum of the slice partials (+ fused top-k lists)
    return launch_scores_topk(ctx, ix->d_pq_partial, slices, ix->count, ix->count, d_scores, k, largest, ix->id_base,
                              n_lists, Kp_out);
}

qz_status pq_score_points(qz_index *ix, const void *d_lut, const uint32_t *d_rows, uint64_t n, float *d_out) {
    pq_points_kernel<<<(uint32_t)((n + 127) / 128), 128, 0, ix->ctx->stream>>>(ix->d_storage, ix->n_chunks,
                                                                              (const float *)d_lut, d_rows, n, d_out);
    return check_launch(ix->ctx, "pq_points_kernel");
}

qz_status pq_score_pairs(qz_index *ix, const uint32_t *d_a, const uint32_t *d_b, uint64_t n, float *d_out) {
    pq_pairs_kernel<<<(uint32_t)((n + 127) / 128), 128, 0, ix->ctx->stream>>>(
        ix->d_storage, ix->n_chunks, ix->d_bounds, ix->d_centroids, ix->distance, ix->invert, d_a, d_b, n, d_out);
    return check_launch(ix->ctx, "pq_pairs_kernel");
}

void pq_index_free(qz_index *ix) {
    cudaFree(ix->d_centroids);
    cudaFree(ix->d_bounds);
    cudaFree(ix->d_lut_perm);
    cudaFree(ix->d_pq_partial);
    delete[] ix->h_bounds;
    ix->d_centroids = nullptr;
    ix->d_bounds = nullptr;
    ix->d_lut_perm = nullptr;
    ix->d_pq_partial = nullptr;
    ix->h_bounds = nullptr;
}

qz_status index_alloc_storage(qz_index *ix, const uint8_t *host_storage);

}  // namespace qz

using namespace qz;

extern "C" {

qz_status qz_pq_index_create(qz_ctx *ctx, const uint8_t *codes, uint64_t count, uint32_t dim, uint32_t chunk_size,
                             const float *centroids, int32_t distance, int32_t invert, uint32_t id_base,
                             qz_index **out) {
    if (!ctx || !out) return fail(ctx, QZ_ARGUMENTS, "qz_pq_index_create: NULL argument");
    *out = nullptr;
    if (distance < QZ_DOT || distance > QZ_L2) return fail(ctx, QZ_ARGUMENTS, "qz_pq_index_create: bad distance");
    if (chunk_size == 0) return fail(ctx, QZ_ARGUMENTS, "qz_pq_index_create: chunk_size is 0");
    if (dim && !centroids) return fail(ctx, QZ_ARGUMENTS, "qz_pq_index_create: NULL centroids");
    if (count + id_base > 0xFFFFFFFEull) return fail(ctx, QZ_ARGUMENTS, "qz_pq_index_create: point ids exceed u32");
    std::lock_guard<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    qz_index *ix = new qz_index();
    ctx->refs.fetch_add(1);
    ix->ctx = ctx;
    ix->codec = QZ_CODEC_PQ;
    ix->count = count;
    ix->dim = dim;
    ix->distance = distance;
    ix->invert = invert != 0;
    ix->id_base = id_base;
    // get_vector_division, pq.rs:116-121
    const uint32_t n_chunks = (dim + chunk_size - 1) / chunk_size;
    ix->n_chunks = n_chunks;
    ix->row_bytes = n_chunks;
    ix->h_bounds = new uint32_t[n_chunks + 1];
    for (uint32_t c = 0; c <= n_chunks; c++) ix->h_bounds[c] = (c * chunk_size < dim) ? c * chunk_size : dim;
    ix->pq_slices = (n_chunks + kSliceChunks - 1) / kSliceChunks;
    // chunk-major centroid table
    std::vector<float> ct((size_t)dim * kPqCentroids);
    for (uint32_t ch = 0; ch < n_chunks; ch++) {
        const uint32_t a = ix->h_bounds[ch], len = ix->h_bounds[ch + 1] - a;
        for (uint32_t c = 0; c < (uint32_t)kPqCentroids; c++)
            memcpy(&ct[(size_t)a * kPqCentroids + (size_t)c * len], centroids + (size_t)c * dim + a, len * sizeof(float));
    }
    cudaError_t e = cudaMalloc((void **)&ix->d_centroids, ct.size() * sizeof(float) + 16);
    if (e == cudaSuccess) e = cudaMalloc((void **)&ix->d_bounds, (n_chunks + 1) * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc((void **)&ix->d_lut_perm, (size_t)(ix->pq_slices ? ix->pq_slices : 1) * kSliceFloats * 4);
    if (e == cudaSuccess)
        e = cudaMalloc((void **)&ix->d_pq_partial, ((size_t)(ix->pq_slices ? ix->pq_slices : 1) * count + 4) * sizeof(float));
    if (e == cudaSuccess && !ct.empty())
        e = cudaMemcpyAsync(ix->d_centroids, ct.data(), ct.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(ix->d_bounds, ix->h_bounds, (n_chunks + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice,
                            ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    qz_status st = e == cudaSuccess ? index_alloc_storage(ix, codes) : cuda_fail(ctx, e, "qz_pq_index_create");
    if (st != QZ_OK) {
        cudaFree(ix->d_storage);
        pq_index_free(ix);
        ctx->refs.fetch_sub(1);
        delete ix;
        return st;
    }
    *out = ix;
    return QZ_OK;
}

}  // extern "C"
