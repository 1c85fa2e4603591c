// encode.cu — batch encode kernels.
//   bin_encode_kernel     EncodedVectorsBin::encode_vector          (encoded_vectors_binary.rs:193-208)
//   minmax_kernel         find_min_max_from_iter                    (quantile.rs:5-19)
//   u8_encode_kernel      EncodedVectorsU8::encode hot loop         (encoded_vectors_u8.rs:73-118)
//   radix-select kernels  find_quantile_interval's two order statistics (quantile.rs:53-70)
#include <float.h>
#include <string.h>

#include <vector>

#include "qz_internal.h"
#include "common.cuh"

namespace qz {

// ---------------------------------------------------------------------------------------------
// binary: bit i = (v[i] > 0) at byte i/8, bit i%8.  One warp ballot = one 32-bit word.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bin_encode_kernel(const float *data, uint64_t count, uint32_t dim,
                                                         uint32_t row_bytes, uint8_t *out) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t words = (row_bytes + 3) / 4;
    const uint64_t total = count * words;
    const uint64_t nwarps = (uint64_t)gridDim.x * (blockDim.x >> 5);
    for (uint64_t it = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); it < total; it += nwarps) {
        const uint64_t row = it / words;
        const uint32_t w = (uint32_t)(it - row * words);
        const uint32_t i = w * 32 + lane;
        const float v = i < dim ? data[row * dim + i] : 0.0f;
        const uint32_t bits = __ballot_sync(0xffffffffu, v > 0.0f);
        uint8_t *dst = out + row * row_bytes + (uint64_t)w * 4;
        if ((row_bytes & 3) == 0) {
            if (lane == 0) *reinterpret_cast<uint32_t *>(dst) = bits;
        } else if (lane < 4 && w * 4 + lane < row_bytes) {
            dst[lane] = (uint8_t)(bits >> (8 * lane));
        }
    }
}

// Fast path for rows of whole float4s and whole 32-bit output words (every BASELINE shape): one warp per row,
// lane l reads the float4 at floats [128 j + 4 l, +4) of 128-float segment j (all loads of a 4-segment batch
// are issued before the first use), turns it into a nibble, and three xor-shuffles gather the eight nibbles of
// a 32-bit word; lanes 0, 8, 16, 24 store words 4 j + l / 8.  No per-element index arithmetic (the generic
// kernel above spends its time in a 64-bit division per loaded float).
__global__ void __launch_bounds__(256) bin_encode_v4_kernel(const float *data, uint64_t count, uint32_t dim,
                                                            uint32_t row_bytes, uint8_t *out) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t nwarps = (uint64_t)gridDim.x * (blockDim.x >> 5);
    const uint32_t segs = (row_bytes + 15) / 16;
    const uint32_t sh = 4 * (lane & 7);
    for (uint64_t row = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < count; row += nwarps) {
        const float *v = data + row * dim;
        uint8_t *o = out + row * row_bytes;
        for (uint32_t j0 = 0; j0 < segs; j0 += 4) {
            float4 f[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const uint32_t i = (j0 + u) * 128 + 4 * lane;
                f[u] = i < dim ? __ldcs(reinterpret_cast<const float4 *>(v + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                uint32_t w = ((f[u].x > 0.0f ? 1u : 0u) | (f[u].y > 0.0f ? 2u : 0u) | (f[u].z > 0.0f ? 4u : 0u) |
                              (f[u].w > 0.0f ? 8u : 0u)) << sh;
                w |= __shfl_xor_sync(0xffffffffu, w, 1);
                w |= __shfl_xor_sync(0xffffffffu, w, 2);
                w |= __shfl_xor_sync(0xffffffffu, w, 4);
                const uint32_t byte = (j0 + u) * 16 + (lane >> 3) * 4;
                if ((lane & 7) == 0 && byte < row_bytes) *reinterpret_cast<uint32_t *>(o + byte) = w;
            }
        }
    }
}

qz_status bin_encode_rows(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, uint32_t row_bytes,
                          uint8_t *d_out) {
    if (count == 0 || row_bytes == 0) return QZ_OK;
    const bool v4 = dim % 4 == 0 && row_bytes % 4 == 0 && ((uintptr_t)d_data & 15) == 0 && ((uintptr_t)d_out & 3) == 0;
    if (v4) {
        uint64_t blocks = (count + 7) / 8;
        const uint64_t cap = (uint64_t)ctx->sm_count * 8;
        if (blocks > cap) blocks = cap;
        bin_encode_v4_kernel<<<(uint32_t)blocks, 256, 0, ctx->stream>>>(d_data, count, dim, row_bytes, d_out);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail(ctx, e, "bin_encode_v4_kernel");
        ctx->launches++;
        return QZ_OK;
    }
    const uint64_t total = count * ((row_bytes + 3) / 4);
    uint64_t blocks = (total + 7) / 8;
    const uint64_t cap = (uint64_t)ctx->sm_count * 32;
    if (blocks > cap) blocks = cap;
    bin_encode_kernel<<<(uint32_t)blocks, 256, 0, ctx->stream>>>(d_data, count, dim, row_bytes, d_out);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(ctx, e, "bin_encode_kernel");
    ctx->launches++;
    return QZ_OK;
}


// ---------------------------------------------------------------------------------------------
// global min / max with the reference's strict-compare fold (NaN never wins), quantile.rs:5-19.
// mm[0] = ordered bits of min (init order(FLT_MAX)), mm[1] = ordered bits of max (init order(-FLT_MAX))
// ---------------------------------------------------------------------------------------------
// mm[0] / mm[1]: min / max as order bits; mm[2..3]: (u64) flat index of the FIRST element equal to zero (either sign),
// ~0 when there is none.  The reference folds with strict < and > (quantile.rs:5-19), so when the minimum (maximum)
// is a zero and zeros of both signs occur, it keeps the zero it saw first; order bits alone would always say -0.0
// (+0.0).  The host resolves the sign from that element (fix_zero_sign).  `base` = flat index of data[0].
__global__ void __launch_bounds__(256) minmax_kernel(const float *data, uint64_t n, uint64_t base, uint32_t *mm) {
    float lo = FLT_MAX, hi = -FLT_MAX;
    unsigned long long zi = ~0ull;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t n4 = (((uintptr_t)data & 15) == 0) ? n / 4 : 0;
    const float4 *d4 = reinterpret_cast<const float4 *>(data);
    for (uint64_t i = tid; i < n4; i += stride) {
        const float4 v = d4[i];
        lo = fminf(lo, fminf(fminf(v.x, v.y), fminf(v.z, v.w)));
        hi = fmaxf(hi, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
        if (zi == ~0ull && (v.x == 0.0f || v.y == 0.0f || v.z == 0.0f || v.w == 0.0f))  // (indices only grow: first hit)
            zi = 4 * i + (v.x == 0.0f ? 0 : (v.y == 0.0f ? 1 : (v.z == 0.0f ? 2 : 3)));
    }
    for (uint64_t i = n4 * 4 + tid; i < n; i += stride) {
        lo = fminf(lo, data[i]);
        hi = fmaxf(hi, data[i]);
        if (data[i] == 0.0f && i < zi) zi = i;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        const unsigned long long z2 = __shfl_xor_sync(0xffffffffu, zi, o);
        zi = z2 < zi ? z2 : zi;
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&mm[0], f32_order_bits(__float_as_uint(lo)));
        atomicMax(&mm[1], f32_order_bits(__float_as_uint(hi)));
        if (zi != ~0ull) atomicMin(reinterpret_cast<unsigned long long *>(mm + 2), zi + base);
    }
}

// ---------------------------------------------------------------------------------------------
// u8 rows: [f32 vector_offset][actual_dim codes], encoded_vectors_u8.rs:78-117.  One warp per row;
// lane l builds code words l, l+32, ... (4 codes each).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t u8_code_of(float v, float alpha, float offset) {
    float t = __fdiv_rn(__fsub_rn(v, offset), alpha);  // IEEE division, u8.rs:235
    if (t != t) return 0u;                             // NaN as u8 == 0
    t = fminf(fmaxf(t, 0.0f), 127.0f);
    return (uint32_t)__float2int_rz(t);
}

// Same code, without the IEEE division on the common path: q' = (v - offset) * (1 / alpha) is within 2^-22 relative
// of the true quotient, so for q' in [-1, 128] it is off by < 4e-5; whenever q' is at least 1e-3 away from every
// integer, trunc(clamp(q')) == trunc(clamp(fl((v - offset) / alpha))).  Everything else — near an integer, NaN,
// infinities, |q'| >= 2^23 (rint(q') == q'), alpha == 0 — takes the exact division above.
__device__ __forceinline__ uint32_t u8_code_fast(float v, float alpha, float inv_alpha, float offset) {
    const float q = __fmul_rn(__fsub_rn(v, offset), inv_alpha);
    if (!(fabsf(__fsub_rn(q, rintf(q))) >= 1e-3f)) return u8_code_of(v, alpha, offset);
    return (uint32_t)__float2int_rz(fminf(fmaxf(q, 0.0f), 127.0f));
}

struct U8EncodeParams {
    const float *data;  // [count][dim]
    uint64_t count;
    uint32_t dim, actual_dim;
    int distance, invert;
    float alpha, offset;
    uint8_t *out;  // [count][actual_dim + 4]
};

__global__ void __launch_bounds__(256) u8_encode_kernel(U8EncodeParams p) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t nwarps = (uint64_t)gridDim.x * (blockDim.x >> 5);
    const uint32_t words = p.actual_dim / 4;
    const uint32_t pad = u8_code_of(p.distance == kDot ? 0.0f : p.offset, p.alpha, p.offset);  // u8.rs:84-93
    const bool vec4 = (p.dim % 4 == 0) && (((uintptr_t)p.data & 15) == 0);
    // a subnormal reciprocal has lost precision: NaN sends every element down the exact path
    const float rcp = __fdiv_rn(1.0f, p.alpha);
    const float inv_alpha = (fabsf(rcp) >= FLT_MIN && fabsf(rcp) <= FLT_MAX) ? rcp : __uint_as_float(0x7fc00000u);
    for (uint64_t row = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < p.count; row += nwarps) {
        const float *v = p.data + row * p.dim;
        uint32_t *o = reinterpret_cast<uint32_t *>(p.out + row * (uint64_t)(p.actual_dim + 4));
        uint32_t s1 = 0, s2 = 0;
        uint32_t w_first = lane;
        if (vec4) {
            // whole float4 words, four per lane in flight before the first use (the pad words, if any, and odd
            // shapes go through the loop below)
            const uint32_t full = p.dim / 4;  // words made of source floats only
            for (; w_first + 96 < full; w_first += 128) {
                float4 f[4];
#pragma unroll
                for (int u = 0; u < 4; u++) f[u] = __ldcs(reinterpret_cast<const float4 *>(v + 4 * (w_first + 32 * u)));
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const uint32_t c0 = u8_code_fast(f[u].x, p.alpha, inv_alpha, p.offset);
                    const uint32_t c1 = u8_code_fast(f[u].y, p.alpha, inv_alpha, p.offset);
                    const uint32_t c2 = u8_code_fast(f[u].z, p.alpha, inv_alpha, p.offset);
                    const uint32_t c3 = u8_code_fast(f[u].w, p.alpha, inv_alpha, p.offset);
                    o[1 + w_first + 32 * u] = c0 | (c1 << 8) | (c2 << 16) | (c3 << 24);
                    s1 += c0 + c1 + c2 + c3;
                    s2 += c0 * c0 + c1 * c1 + c2 * c2 + c3 * c3;
                }
            }
        }
        for (uint32_t w = w_first; w < words; w += 32) {
            const uint32_t i = 4 * w;
            uint32_t c0, c1, c2, c3;
            if (vec4 && i + 3 < p.dim) {
                const float4 f = __ldcs(reinterpret_cast<const float4 *>(v + i));
                c0 = u8_code_fast(f.x, p.alpha, inv_alpha, p.offset);
                c1 = u8_code_fast(f.y, p.alpha, inv_alpha, p.offset);
                c2 = u8_code_fast(f.z, p.alpha, inv_alpha, p.offset);
                c3 = u8_code_fast(f.w, p.alpha, inv_alpha, p.offset);
            } else {
                c0 = i < p.dim ? u8_code_of(v[i], p.alpha, p.offset) : pad;
                c1 = i + 1 < p.dim ? u8_code_of(v[i + 1], p.alpha, p.offset) : pad;
                c2 = i + 2 < p.dim ? u8_code_of(v[i + 2], p.alpha, p.offset) : pad;
                c3 = i + 3 < p.dim ? u8_code_of(v[i + 3], p.alpha, p.offset) : pad;
            }
            o[1 + w] = c0 | (c1 << 8) | (c2 << 16) | (c3 << 24);
            s1 += c0 + c1 + c2 + c3;
            s2 += c0 * c0 + c1 * c1 + c2 * c2 + c3 * c3;
        }
#pragma unroll
        for (int k = 16; k > 0; k >>= 1) {
            s1 += __shfl_xor_sync(0xffffffffu, s1, k);
            s2 += __shfl_xor_sync(0xffffffffu, s2, k);
        }
        __syncwarp();
        if (lane == 0) {
            // vector_offset, u8.rs:94-114.  The reference sums code values sequentially in f32; that
            // equals the exact integer while the total stays below 2^24, else replay the order.
            float vo = 0.0f;
            if (p.distance != kL1) {
                uint32_t tot = p.distance == kDot ? s1 : s2;
                float S = (float)tot;
                if (tot >= (1u << 24)) {
                    const uint8_t *codes = reinterpret_cast<const uint8_t *>(o + 1);
                    S = 0.0f;
                    for (uint32_t i = 0; i < p.actual_dim; i++) {
                        const float c = (float)codes[i];
                        S = __fadd_rn(S, p.distance == kDot ? c : __fmul_rn(c, c));
                    }
                }
                const float base = __fmul_rn(__fmul_rn((float)p.actual_dim, p.offset), p.offset);
                const float second = p.distance == kDot ? __fmul_rn(__fmul_rn(S, p.alpha), p.offset)
                                                        : __fmul_rn(__fmul_rn(S, p.alpha), p.alpha);
                vo = __fadd_rn(base, second);
            }
            if (p.invert) vo = -vo;
            o[0] = __float_as_uint(vo);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// radix select over ordered f32 bits: histogram of the next digit among elements whose high bits
// match `prefix` (two independent ranks in one pass), quantile.rs:58-61's two order statistics
// ---------------------------------------------------------------------------------------------
struct SelectPass {
    const float *data;
    uint64_t n;
    uint32_t shift, bits;      // digit = (key >> shift) & ((1<<bits)-1)
    uint32_t hi_shift;         // prefix compare: key >> hi_shift == prefix[r]  (hi_shift = 32 => no prefix)
    uint32_t prefix[2];
    uint32_t *hist;  // [2][2048]
    const uint64_t *ids;  // optional: rows of the sample inside `data`
    uint32_t dim;
};

// One element into the CTA histograms.  Lanes of a warp that hit the same digit are combined first (match.any) and
// one lane adds their count: on real data most keys share a few sign/exponent digits in the first pass, and
// same-address shared-memory atomics serialise.
__device__ __forceinline__ void select_hist_add(uint32_t (*h)[2048], const SelectPass &p, float x, uint32_t mask) {
    const uint32_t key = f32_order_bits(__float_as_uint(x));
    const uint32_t hi = p.hi_shift >= 32 ? 0u : key >> p.hi_shift;
    const uint32_t d = (key >> p.shift) & mask;
    const unsigned act = __activemask();
    const bool m0 = hi == p.prefix[0], m1 = hi == p.prefix[1];
    const unsigned b0 = __ballot_sync(act, m0), b1 = __ballot_sync(act, m1);
    if ((b0 | b1) == 0) return;  // passes 2 and 3: almost no key carries a selected prefix
    if (__popc(b0 | b1) <= 4) {  // (warp-uniform branches: match.any below is executed by all of `act` or none)
        if (m0) atomicAdd(&h[0][d], 1u);
        if (m1) atomicAdd(&h[1][d], 1u);
        return;
    }
    const unsigned peers = __match_any_sync(act, d);
    if ((threadIdx.x & 31) == (uint32_t)(__ffs(peers) - 1)) {
        const uint32_t c0 = __popc(peers & b0), c1 = __popc(peers & b1);
        if (c0) atomicAdd(&h[0][d], c0);
        if (c1) atomicAdd(&h[1][d], c1);
    }
}

// ids == nullptr: the n values at data[0 .. n).  Otherwise the sample is read in place, n / dim rows data[ids[r] * dim ..]
// (one CTA per row at a time) — no gathered copy of the sample is materialised.
__global__ void __launch_bounds__(256) select_hist_kernel(SelectPass p) {
    __shared__ uint32_t h[2][2048];
    for (uint32_t i = threadIdx.x; i < 4096; i += blockDim.x) (&h[0][0])[i] = 0;
    __syncthreads();
    const uint32_t mask = (1u << p.bits) - 1;
    if (p.ids == nullptr) {
        const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
        for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.n; i += stride)
            select_hist_add(h, p, p.data[i], mask);
    } else {
        const uint64_t n_rows = p.n / p.dim;
        for (uint64_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
            const float *src = p.data + p.ids[r] * p.dim;
            for (uint32_t j = threadIdx.x; j < p.dim; j += blockDim.x) select_hist_add(h, p, src[j], mask);
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < 4096; i += blockDim.x) {
        const uint32_t v = (&h[0][0])[i];
        if (v) atomicAdd(&p.hist[i], v);
    }
}

__global__ void __launch_bounds__(256) gather_rows_kernel(const float *data, uint32_t dim, const uint64_t *ids,
                                                          uint64_t n_ids, float *out) {
    const uint64_t total = n_ids * dim;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const uint64_t r = i / dim;
        out[i] = data[ids[r] * dim + (i - r * dim)];
    }
}

static qz_status launched(qz_ctx *ctx, const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(ctx, e, what);
    ctx->launches++;
    return QZ_OK;
}

static uint32_t grid_for(qz_ctx *ctx, uint64_t items, uint32_t per_block, uint32_t waves) {
    uint64_t b = (items + per_block - 1) / per_block;
    const uint64_t cap = (uint64_t)ctx->sm_count * waves;
    if (b > cap) b = cap;
    return b ? (uint32_t)b : 1u;
}

// min / max of a device array -> host floats.  `base` = flat index of d_data[0] within the whole corpus (the host
// encode streams the corpus through the device in pieces); *zero_idx = flat index of the first zero, ~0 if none.
qz_status dev_minmax(qz_ctx *ctx, const float *d_data, uint64_t n, uint32_t *d_mm, bool init, uint64_t base) {
    if (init) {
        uint32_t h[4] = {f32_order_bits(0x7F7FFFFFu), f32_order_bits(0xFF7FFFFFu), 0xFFFFFFFFu, 0xFFFFFFFFu};  // FLT_MAX, -FLT_MAX, ~0
        QZ_CUDA(ctx, cudaMemcpyAsync(d_mm, h, sizeof(h), cudaMemcpyHostToDevice, ctx->stream));
        QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // h is a stack temporary
    }
    if (n == 0) return QZ_OK;
    minmax_kernel<<<grid_for(ctx, n, 256 * 16, 8), 256, 0, ctx->stream>>>(d_data, n, base, d_mm);
    return launched(ctx, "minmax_kernel");
}

qz_status read_minmax(qz_ctx *ctx, const uint32_t *d_mm, float *mn, float *mx, uint64_t *zero_idx) {
    uint32_t h[4];
    QZ_CUDA(ctx, cudaMemcpyAsync(h, d_mm, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    uint32_t a = f32_from_order_bits(h[0]), b = f32_from_order_bits(h[1]);
    memcpy(mn, &a, 4);
    memcpy(mx, &b, 4);
    if (zero_idx) *zero_idx = ((uint64_t)h[3] << 32) | h[2];
    return QZ_OK;
}

// find_min_max_from_iter keeps the first zero it meets when the extreme is a zero (strict comparisons): give a zero
// minimum / maximum the sign of the corpus' first zero element
void fix_zero_sign(float *mn, float *mx, float first_zero) {
    if (*mn == 0.0f) *mn = first_zero;
    if (*mx == 0.0f) *mx = first_zero;
}

// the values at ascending ranks r0 and r1 of a device array (exact, 3 histogram passes)
qz_status dev_select2(qz_ctx *ctx, const float *d_vals, uint64_t n, uint64_t r0, uint64_t r1, float *v0, float *v1,
                      const uint64_t *d_ids, uint32_t dim) {
    uint32_t *d_hist = nullptr;
    QZ_CUDA(ctx, cudaMalloc((void **)&d_hist, 4096 * sizeof(uint32_t)));
    std::vector<uint32_t> h(4096);
    uint64_t rank[2] = {r0, r1};
    uint32_t prefix[2] = {0, 0};
    const uint32_t shifts[3] = {21, 10, 0}, bits[3] = {11, 11, 10};
    qz_status st = QZ_OK;
    for (int pass = 0; pass < 3 && st == QZ_OK; pass++) {
        cudaError_t e = cudaMemsetAsync(d_hist, 0, 4096 * sizeof(uint32_t), ctx->stream);
        if (e != cudaSuccess) { st = cuda_fail(ctx, e, "select memset"); break; }
        SelectPass p{d_vals, n, shifts[pass], bits[pass], pass == 0 ? 32u : shifts[pass - 1], {prefix[0], prefix[1]}, d_hist, d_ids, dim};
        select_hist_kernel<<<grid_for(ctx, n, 256 * 32, 4), 256, 0, ctx->stream>>>(p);
        st = launched(ctx, "select_hist_kernel");
        if (st != QZ_OK) break;
        e = cudaMemcpyAsync(h.data(), d_hist, 4096 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) { st = cuda_fail(ctx, e, "select readback"); break; }
        for (int r = 0; r < 2; r++) {
            uint64_t acc = 0;
            uint32_t d = 0;
            const uint32_t nb = 1u << bits[pass];
            for (; d < nb; d++) {
                if (acc + h[r * 2048 + d] > rank[r]) break;
                acc += h[r * 2048 + d];
            }
            if (d == nb) d = nb - 1;
            rank[r] -= acc;
            prefix[r] = (prefix[r] << bits[pass]) | d;
        }
    }
    cudaFree(d_hist);
    if (st != QZ_OK) return st;
    uint32_t a = f32_from_order_bits(prefix[0]), b = f32_from_order_bits(prefix[1]);
    memcpy(v0, &a, 4);
    memcpy(v1, &b, 4);
    return QZ_OK;
}

qz_status launch_u8_encode(qz_ctx *ctx, const float *d_data, uint64_t count, uint32_t dim, int distance, int invert,
                           float alpha, float offset, uint8_t *d_out) {
    if (count == 0) return QZ_OK;
    U8EncodeParams p{d_data, count, dim, qz_u8_actual_dim(dim), distance, invert, alpha, offset, d_out};
    u8_encode_kernel<<<grid_for(ctx, count, 8, 16), 256, 0, ctx->stream>>>(p);
    return launched(ctx, "u8_encode_kernel");
}

qz_status launch_gather_rows(qz_ctx *ctx, const float *d_data, uint32_t dim, const uint64_t *d_ids, uint64_t n_ids,
                             float *d_out) {
    if (n_ids == 0 || dim == 0) return QZ_OK;
    gather_rows_kernel<<<grid_for(ctx, n_ids * dim, 256 * 8, 8), 256, 0, ctx->stream>>>(d_data, dim, d_ids, n_ids, d_out);
    return launched(ctx, "gather_rows_kernel");
}

}  // namespace qz
