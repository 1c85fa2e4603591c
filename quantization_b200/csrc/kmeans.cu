// kmeans.cu — PQ training and assignment.
//   pq_assign_kernel      encode_vector (encoded_vectors_pq.rs:237-265) == update_indexes (kmeans.rs:138-167):
//                         per chunk argmin_c sum_t (a_t - b_t)^2, sequential f32, strict <, lowest c wins
//   kmeans_update_kernel  update_centroids (kmeans.rs:49-136) for every chunk at once: f64 sums in point
//                         order (== the reference with max_threads = 1), mean -> f32, empty cluster <-
//                         pseudo-random sample point, converged iff sum |delta| < accuracy
// All 'n_chunks' k-means problems of find_centroids (pq.rs:310-339) run side by side instead of serially.
#include <string.h>

#include <algorithm>
#include <vector>

#include "qz_internal.h"
#include "common.cuh"

namespace qz {

constexpr int kCentroids = 256;

struct AssignParams {
    const float *data;  // [n][dim]
    uint64_t n;
    uint32_t dim;
    const float *ct;         // chunk-major centroids
    const uint32_t *bounds;  // [n_chunks + 1]
    uint32_t n_chunks;
    uint32_t chunks_per_cta;
    const uint8_t *done;  // optional [n_chunks]: skip converged chunks
    uint8_t *codes;       // corpus layout [n][n_chunks]   (exactly one of codes / idx is set)
    uint8_t *idx;         // training layout [n_chunks][n]
};

// one thread = one point; the CTA walks a range of chunks, staging each chunk's 256 sub-centroids in smem
template <int LEN>
__global__ void __launch_bounds__(256) pq_assign_kernel(AssignParams p) {
    extern __shared__ float cen[];  // [256][len]
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < p.n;
    const float *row = p.data + (live ? i : 0) * p.dim;
    const uint32_t ch0 = blockIdx.y * p.chunks_per_cta;
    const uint32_t ch1 = min(ch0 + p.chunks_per_cta, p.n_chunks);
    for (uint32_t ch = ch0; ch < ch1; ch++) {
        if (p.done && p.done[ch]) continue;
        const uint32_t a = p.bounds[ch], len = LEN ? LEN : p.bounds[ch + 1] - a;
        __syncthreads();
        for (uint32_t j = threadIdx.x; j < kCentroids * len; j += blockDim.x) cen[j] = p.ct[(size_t)a * kCentroids + j];
        __syncthreads();
        if (!live) continue;
        float best = 3.402823466e+38f;  // f32::MAX
        uint32_t bi = 0;
        if (LEN) {
            float v[LEN ? LEN : 1];
#pragma unroll
            for (int t = 0; t < LEN; t++) v[t] = row[a + t];
#pragma unroll 4
            for (uint32_t c = 0; c < (uint32_t)kCentroids; c++) {
                float d = 0.0f;
#pragma unroll
                for (int t = 0; t < LEN; t++) {
                    const float x = __fsub_rn(v[t], cen[c * LEN + t]);
                    d = __fadd_rn(d, __fmul_rn(x, x));
                }
                if (d < best) {
                    best = d;
                    bi = c;
                }
            }
        } else {
            for (uint32_t c = 0; c < (uint32_t)kCentroids; c++) {
                float d = 0.0f;
                for (uint32_t t = 0; t < len; t++) {
                    const float x = __fsub_rn(row[a + t], cen[c * len + t]);
                    d = __fadd_rn(d, __fmul_rn(x, x));
                }
                if (d < best) {
                    best = d;
                    bi = c;
                }
            }
        }
        if (p.codes) p.codes[i * p.n_chunks + ch] = (uint8_t)bi;
        else p.idx[(size_t)ch * p.n + i] = (uint8_t)bi;
    }
}

// Second generation, for uniform chunk lengths 1/2/4/8 and dim % 8 == 0 (every BASELINE shape).  Same arithmetic, less
// overhead per (point, centroid) pair: a thread owns TWO points (one LDS of a sub-centroid serves both), reads its
// rows in whole 32-byte sectors (G = 8 / LEN chunks per load, the next group's loads issued before this group's
// arithmetic), drops the `0.0 + x*x` of the first term (x*x is never -0, so the sum is bit-identical), and writes
// the G codes of a group as one word.
template <int LEN>
__global__ void __launch_bounds__(256) pq_assign2_kernel(AssignParams p) {
    constexpr int G = 8 / LEN;
    __shared__ __align__(16) float cen[8 * kCentroids];  // [G][256][LEN]
    const uint64_t i0 = (uint64_t)blockIdx.x * 512 + threadIdx.x, i1 = i0 + 256;
    const bool live0 = i0 < p.n, live1 = i1 < p.n;
    const float *row0 = p.data + (live0 ? i0 : 0) * p.dim;
    const float *row1 = p.data + (live1 ? i1 : 0) * p.dim;
    const uint32_t n_groups = p.n_chunks / G;
    const uint32_t g0 = blockIdx.y * p.chunks_per_cta, g1 = min(g0 + p.chunks_per_cta, n_groups);  // in groups
    float va[8], vb[8], na[8], nb[8];
    auto load8 = [](const float *src, float *dst) {
        const float4 x = *reinterpret_cast<const float4 *>(src), y = *reinterpret_cast<const float4 *>(src + 4);
        dst[0] = x.x; dst[1] = x.y; dst[2] = x.z; dst[3] = x.w;
        dst[4] = y.x; dst[5] = y.y; dst[6] = y.z; dst[7] = y.w;
    };
    if (g0 < g1) {
        load8(row0 + 8 * g0, na);
        load8(row1 + 8 * g0, nb);
    }
    for (uint32_t g = g0; g < g1; g++) {
#pragma unroll
        for (int t = 0; t < 8; t++) {
            va[t] = na[t];
            vb[t] = nb[t];
        }
        bool any = p.done == nullptr;
        if (!any)
            for (int k = 0; k < G; k++) any |= p.done[g * G + k] == 0;
        if (g + 1 < g1) {
            load8(row0 + 8 * (g + 1), na);
            load8(row1 + 8 * (g + 1), nb);
        }
        if (!any) continue;
        __syncthreads();
        {
            const float4 *src = reinterpret_cast<const float4 *>(p.ct + (size_t)g * 8 * kCentroids);
            float4 *dst = reinterpret_cast<float4 *>(cen);
            dst[threadIdx.x] = src[threadIdx.x];
            dst[threadIdx.x + 256] = src[threadIdx.x + 256];
        }
        __syncthreads();
        uint32_t code0[G], code1[G];
#pragma unroll
        for (int k = 0; k < G; k++) {
            code0[k] = code1[k] = 0;
            if (p.done && p.done[g * G + k]) continue;
            float best0 = 3.402823466e+38f, best1 = 3.402823466e+38f;  // f32::MAX
            uint32_t b0 = 0, b1 = 0;
            const float *cc = cen + k * kCentroids * LEN;
#pragma unroll 4
            for (uint32_t c = 0; c < (uint32_t)kCentroids; c++) {
                float d0, d1;
#pragma unroll
                for (int t = 0; t < LEN; t++) {
                    const float ce = cc[c * LEN + t];
                    const float x0 = __fsub_rn(va[k * LEN + t], ce), x1 = __fsub_rn(vb[k * LEN + t], ce);
                    const float s0 = __fmul_rn(x0, x0), s1 = __fmul_rn(x1, x1);
                    d0 = t == 0 ? s0 : __fadd_rn(d0, s0);
                    d1 = t == 0 ? s1 : __fadd_rn(d1, s1);
                }
                if (d0 < best0) {
                    best0 = d0;
                    b0 = c;
                }
                if (d1 < best1) {
                    best1 = d1;
                    b1 = c;
                }
            }
            code0[k] = b0;
            code1[k] = b1;
        }
        if (p.codes) {
            uint8_t *o0 = p.codes + i0 * p.n_chunks + (size_t)g * G, *o1 = p.codes + i1 * p.n_chunks + (size_t)g * G;
            if (G == 4 && p.done == nullptr && (p.n_chunks & 3) == 0 && ((uintptr_t)p.codes & 3) == 0) {
                const uint32_t w0 = code0[0] | (code0[G > 1 ? 1 : 0] << 8) | (code0[G > 2 ? 2 : 0] << 16) | (code0[G > 3 ? 3 : 0] << 24);
                const uint32_t w1 = code1[0] | (code1[G > 1 ? 1 : 0] << 8) | (code1[G > 2 ? 2 : 0] << 16) | (code1[G > 3 ? 3 : 0] << 24);
                if (live0) *reinterpret_cast<uint32_t *>(o0) = w0;
                if (live1) *reinterpret_cast<uint32_t *>(o1) = w1;
            } else {
#pragma unroll
                for (int k = 0; k < G; k++) {
                    if (p.done && p.done[g * G + k]) continue;
                    if (live0) o0[k] = (uint8_t)code0[k];
                    if (live1) o1[k] = (uint8_t)code1[k];
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < G; k++) {
                if (p.done && p.done[g * G + k]) continue;
                uint8_t *o = p.idx + (size_t)(g * G + k) * p.n;
                if (live0) o[i0] = (uint8_t)code0[k];
                if (live1) o[i1] = (uint8_t)code1[k];
            }
        }
    }
}

__device__ __forceinline__ uint64_t splitmix_dev(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

struct UpdateParams {
    const float *sample;  // [n][dim]
    uint64_t n;
    uint32_t dim;
    float *ct;
    const uint32_t *bounds;
    const uint8_t *idx;  // [n_chunks][n]
    uint8_t *done;       // [n_chunks] in/out
    uint64_t seed;
    uint32_t iter;
    float accuracy;
};

// CTA = one chunk, thread c = centroid c.  Members are visited in point order, so every f64 sum is
// the reference's sequential sum.
__global__ void __launch_bounds__(256) kmeans_update_kernel(UpdateParams p) {
    extern __shared__ float newc[];  // [256][len]
    __shared__ uint8_t tile[1024];
    const uint32_t ch = blockIdx.x, c = threadIdx.x;
    if (p.done[ch]) return;
    const uint32_t a = p.bounds[ch], len = p.bounds[ch + 1] - a;
    const uint8_t *idx = p.idx + (size_t)ch * p.n;
    uint64_t cnt = 0;
    for (uint32_t t0 = 0; t0 < len; t0 += 8) {
        double acc[8];
#pragma unroll
        for (int t = 0; t < 8; t++) acc[t] = 0.0;
        cnt = 0;
        for (uint64_t base = 0; base < p.n; base += 1024) {
            __syncthreads();
            for (uint32_t j = threadIdx.x; j < 1024 && base + j < p.n; j += blockDim.x) tile[j] = idx[base + j];
            __syncthreads();
            const uint32_t m = (uint32_t)min((uint64_t)1024, p.n - base);
            for (uint32_t j = 0; j < m; j++) {
                if (tile[j] == c) {
                    cnt++;
                    const float *v = p.sample + (base + j) * p.dim + a + t0;
#pragma unroll
                    for (int t = 0; t < 8; t++)
                        if (t0 + t < len) acc[t] += (double)v[t];
                }
            }
        }
        // mean in f64 -> f32 (kmeans.rs:119-121,129); empty cluster <- a sample point (kmeans.rs:111-118)
        const uint64_t pick = splitmix_dev(p.seed ^ splitmix_dev(((uint64_t)p.iter << 40) ^ ((uint64_t)ch << 8) ^ c)) % p.n;
#pragma unroll
        for (int t = 0; t < 8; t++) {
            if (t0 + t < len) {
                float nv;
                if (cnt == 0) nv = p.sample[pick * p.dim + a + t0 + t];
                else nv = (float)(acc[t] / (double)cnt);
                newc[c * len + t0 + t] = nv;
            }
        }
    }
    __syncthreads();
    float *cen = p.ct + (size_t)a * kCentroids;
    if (threadIdx.x == 0) {
        float diff = 0.0f;  // sequential f32 sum over the flattened [c][t] order (kmeans.rs:125-134)
        for (uint32_t j = 0; j < kCentroids * len; j++) diff = __fadd_rn(diff, fabsf(__fsub_rn(cen[j], newc[j])));
        if (diff < p.accuracy) p.done[ch] = 1;
    }
    __syncthreads();
    for (uint32_t j = threadIdx.x; j < kCentroids * len; j += blockDim.x) cen[j] = newc[j];
}

static qz_status launched(qz_ctx *ctx, const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(ctx, e, what);
    ctx->launches++;
    return QZ_OK;
}

static uint32_t max_chunk_len(const std::vector<uint32_t> &b) {
    uint32_t m = 0;
    for (size_t i = 0; i + 1 < b.size(); i++) m = std::max(m, b[i + 1] - b[i]);
    return m;
}
static bool uniform_len(const std::vector<uint32_t> &b, uint32_t len) {
    for (size_t i = 0; i + 1 < b.size(); i++)
        if (b[i + 1] - b[i] != len) return false;
    return true;
}

qz_status launch_pq_assign(qz_ctx *ctx, const float *d_data, uint64_t n, uint32_t dim, const float *d_ct,
                           const uint32_t *d_bounds, const std::vector<uint32_t> &bounds, const uint8_t *d_done,
                           uint8_t *d_codes, uint8_t *d_idx) {
    const uint32_t n_chunks = (uint32_t)bounds.size() - 1;
    if (n == 0 || n_chunks == 0) return QZ_OK;
    const uint32_t mlen = max_chunk_len(bounds);
    const size_t smem = (size_t)kCentroids * mlen * sizeof(float);
    if (smem > 200 * 1024) return fail(ctx, QZ_ARGUMENTS, "pq: chunk_size too large (256*chunk_size floats must fit shared memory)");
    AssignParams p{d_data, n, dim, d_ct, d_bounds, n_chunks, 0, d_done, d_codes, d_idx};
    for (uint32_t len : {1u, 2u, 4u, 8u}) {
        if (!uniform_len(bounds, len) || dim % 8 != 0 || ((uintptr_t)d_data & 15) != 0 || ((uintptr_t)d_ct & 15) != 0) continue;
        const uint32_t n_groups = n_chunks / (8 / len);  // dim % 8 == 0 => whole groups
        const uint32_t gx2 = (uint32_t)((n + 511) / 512);
        uint32_t gy2 = std::max<uint32_t>(1, std::min<uint32_t>(n_groups, (uint32_t)(8ull * ctx->sm_count / std::max<uint32_t>(gx2, 1)) + 1));
        p.chunks_per_cta = (n_groups + gy2 - 1) / gy2;  // counted in groups for this kernel
        gy2 = (n_groups + p.chunks_per_cta - 1) / p.chunks_per_cta;
        dim3 grid2(gx2, gy2);
        if (len == 1) pq_assign2_kernel<1><<<grid2, 256, 0, ctx->stream>>>(p);
        else if (len == 2) pq_assign2_kernel<2><<<grid2, 256, 0, ctx->stream>>>(p);
        else if (len == 4) pq_assign2_kernel<4><<<grid2, 256, 0, ctx->stream>>>(p);
        else pq_assign2_kernel<8><<<grid2, 256, 0, ctx->stream>>>(p);
        return launched(ctx, "pq_assign2_kernel");
    }
    const uint32_t gx = (uint32_t)((n + 255) / 256);
    // enough CTAs to fill the machine, but long chunk runs per CTA so a point's row stays in L1
    uint32_t gy = std::max<uint32_t>(1, std::min<uint32_t>(n_chunks, (uint32_t)(4ull * ctx->sm_count / std::max<uint32_t>(gx, 1)) + 1));
    p.chunks_per_cta = (n_chunks + gy - 1) / gy;
    gy = (n_chunks + p.chunks_per_cta - 1) / p.chunks_per_cta;
    dim3 grid(gx, gy);
#define QZ_ASSIGN(L)                                                                                            \
    do {                                                                                                        \
        if (smem > 48 * 1024)                                                                                   \
            cudaFuncSetAttribute(pq_assign_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);  \
        pq_assign_kernel<L><<<grid, 256, smem, ctx->stream>>>(p);                                               \
    } while (0)
    if (uniform_len(bounds, 1)) QZ_ASSIGN(1);
    else if (uniform_len(bounds, 2)) QZ_ASSIGN(2);
    else if (uniform_len(bounds, 4)) QZ_ASSIGN(4);
    else if (uniform_len(bounds, 8)) QZ_ASSIGN(8);
    else QZ_ASSIGN(0);
#undef QZ_ASSIGN
    return launched(ctx, "pq_assign_kernel");
}

qz_status launch_kmeans_update(qz_ctx *ctx, const float *d_sample, uint64_t n, uint32_t dim, float *d_ct,
                               const uint32_t *d_bounds, const std::vector<uint32_t> &bounds, const uint8_t *d_idx,
                               uint8_t *d_done, uint64_t seed, uint32_t iter, float accuracy) {
    const uint32_t n_chunks = (uint32_t)bounds.size() - 1;
    const size_t smem = (size_t)kCentroids * max_chunk_len(bounds) * sizeof(float);
    UpdateParams p{d_sample, n, dim, d_ct, d_bounds, d_idx, d_done, seed, iter, accuracy};
    if (smem > 48 * 1024) cudaFuncSetAttribute(kmeans_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kmeans_update_kernel<<<n_chunks, 256, smem, ctx->stream>>>(p);
    return launched(ctx, "kmeans_update_kernel");
}

}  // namespace qz
