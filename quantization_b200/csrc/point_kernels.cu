// point_kernels.cu — small kernels around the scans:
//   u8_encode_query_kernel   EncodedVectorsU8::encode_query        (encoded_vectors_u8.rs:290-329)
//   gather_score_kernel      score_point / score_internal on id lists for u8 and binary
//                            (u8.rs:331-453, encoded_vectors_binary.rs:293-314); also the fallback
//                            full scan for row shapes the streaming kernel does not take
//   scores_topk_kernel       top-k lists from a score array (optionally the ordered sum of
//                            several partial arrays)
#include "qz_internal.h"
#include "row_scan.cuh"

namespace qz {

// f32_to_u8, encoded_vectors_u8.rs:234-237: ((v - offset) / alpha).clamp(0,127) as u8
__device__ __forceinline__ uint32_t u8_code(float v, float alpha, float offset) {
    float t = __fdiv_rn(__fsub_rn(v, offset), alpha);
    if (t != t) return 0u;
    t = fminf(fmaxf(t, 0.0f), 127.0f);
    return (uint32_t)__float2int_rz(t);
}

struct U8QueryParams {
    const float *queries;  // [nq][dim]
    uint8_t *out;          // [nq][actual_dim + 4]
    uint32_t dim, actual_dim;
    int distance, invert;
    float alpha, offset;
};

__global__ void __launch_bounds__(256) u8_encode_query_kernel(U8QueryParams p) {
    __shared__ unsigned long long s1, s2;
    // a scan launched behind this kernel as its programmatic dependent (qz_search_topk) may start setting up now; it
    // waits for this grid before it reads the encoded query.  No effect on ordinary launches.
    asm volatile("griddepcontrol.launch_dependents;");
    const float *q = p.queries + (size_t)blockIdx.x * p.dim;
    uint8_t *out = p.out + (size_t)blockIdx.x * (p.actual_dim + 4);
    if (threadIdx.x == 0) s1 = s2 = 0;
    __syncthreads();
    const uint32_t pad = u8_code(p.distance == kDot ? 0.0f : p.offset, p.alpha, p.offset);
    unsigned long long a1 = 0, a2 = 0;
    // The query may sit in pinned HOST memory (qz_search_topk passes its staging buffer): a load is then a PCIe round
    // trip of 1-2 us, so all loads of a pass are issued before the first code is stored (a load after a byte store
    // through `out` could not be hoisted: the two pointers may alias).  8 x 256 values per pass: one pass up to 2048-d.
    constexpr uint32_t kInFlight = 8;
    for (uint32_t base = 0; base < p.actual_dim; base += kInFlight * 256) {
        float v[kInFlight];
#pragma unroll
        for (uint32_t u = 0; u < kInFlight; u++) {
            const uint32_t i = base + u * 256 + threadIdx.x;
            v[u] = 0.0f;
            if (i < p.dim) asm volatile("ld.global.f32 %0, [%1];" : "=f"(v[u]) : "l"(q + i));
        }
#pragma unroll
        for (uint32_t u = 0; u < kInFlight; u++) {
            const uint32_t i = base + u * 256 + threadIdx.x;
            if (i < p.actual_dim) {
                const uint32_t c = i < p.dim ? u8_code(v[u], p.alpha, p.offset) : pad;
                out[4 + i] = (uint8_t)c;
                a1 += c;
                a2 += c * c;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {  // one 64-bit shared atomic per warp, not per thread (CAS loop)
        a1 += __shfl_xor_sync(0xffffffffu, a1, o);
        a2 += __shfl_xor_sync(0xffffffffu, a2, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s1, a1);
        atomicAdd(&s2, a2);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // integer sums equal the reference's sequential f32 sums while they stay below 2^24;
        // beyond that replay the f32 order (u8.rs:309,315)
        float S;
        if (p.distance == kDot) {
            S = (float)s1;
            if (s1 >= (1ull << 24)) {
                S = 0.0f;
                for (uint32_t i = 0; i < p.actual_dim; i++) S = __fadd_rn(S, (float)out[4 + i]);
            }
        } else {
            S = (float)s2;
            if (s2 >= (1ull << 24)) {
                S = 0.0f;
                for (uint32_t i = 0; i < p.actual_dim; i++) {
                    float c = (float)out[4 + i];
                    S = __fadd_rn(S, __fmul_rn(c, c));
                }
            }
        }
        float qo;
        if (p.distance == kDot) qo = __fmul_rn(__fmul_rn(S, p.alpha), p.offset);
        else if (p.distance == kL1) qo = 0.0f;
        else qo = __fmul_rn(__fmul_rn(S, p.alpha), p.alpha);
        if (p.invert) qo = -qo;
        *reinterpret_cast<float *>(out) = qo;
    }
}

qz_status launch_u8_encode_query(qz_index *ix, const float *d_queries, uint32_t nq, void *d_out) {
    if (nq == 0) return QZ_OK;
    U8QueryParams p{d_queries, (uint8_t *)d_out, ix->dim, ix->actual_dim, ix->distance, ix->invert, ix->alpha, ix->offset};
    u8_encode_query_kernel<<<nq, 256, 0, ix->ctx->stream>>>(p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(ix->ctx, e, "u8_encode_query_kernel");
    ix->ctx->launches++;
    return QZ_OK;
}

// ---------------------------------------------------------------------------------------------
// warp-per-item scoring of (operand a, row b):  a = the encoded query (rows_a == NULL) or a row
// ---------------------------------------------------------------------------------------------
struct GatherParams {
    const uint8_t *storage;
    uint32_t row_bytes, hdr_bytes;  // hdr_bytes = 4 for u8
    const uint8_t *query;           // blob (same shape as a row) or NULL for pairs
    const uint32_t *rows_a, *rows_b;  // rows_b NULL => identity (full scan)
    uint64_t n;
    int op;  // kOpDot / kOpL1 / kOpXor
    float multiplier, pair_diff;
    uint32_t dim;
    int bin_zx;
    float *out;
};

__global__ void __launch_bounds__(256) gather_score_kernel(GatherParams p) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warps = (uint64_t)gridDim.x * (blockDim.x >> 5);
    for (uint64_t it = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); it < p.n; it += warps) {
        const uint64_t rb = p.rows_b ? p.rows_b[it] : it;
        const uint8_t *b = p.storage + rb * p.row_bytes;
        const uint8_t *a = p.rows_a ? p.storage + (uint64_t)p.rows_a[it] * p.row_bytes : p.query;
        const uint32_t nbytes = p.row_bytes - p.hdr_bytes;
        uint32_t acc = 0;
        if ((p.row_bytes & 3) == 0 && (((uintptr_t)a) & 3) == 0) {
            const uint32_t *aw = reinterpret_cast<const uint32_t *>(a + p.hdr_bytes);
            const uint32_t *bw = reinterpret_cast<const uint32_t *>(b + p.hdr_bytes);
            for (uint32_t w = lane; w < nbytes / 4; w += 32) {
                uint32_t x = aw[w], y = bw[w];
                acc = p.op == kOpDot ? __dp4a(x, y, acc) : (p.op == kOpL1 ? acc + __vsadu4(x, y) : acc + __popc(x ^ y));
            }
        } else {
            for (uint32_t i = lane; i < nbytes; i += 32) {
                uint32_t x = a[p.hdr_bytes + i], y = b[p.hdr_bytes + i];
                acc += p.op == kOpDot ? x * y : (p.op == kOpL1 ? (x > y ? x - y : y - x) : (uint32_t)__popc(x ^ y));
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) {
            float score;
            if (p.op == kOpXor) {
                const float x = __uint2float_rn(acc);
                const float z = __fsub_rn(__uint2float_rn(p.dim), x);
                score = p.bin_zx ? __fsub_rn(z, x) : __fsub_rn(x, z);
            } else {
                const float oa = *reinterpret_cast<const float *>(a), ob = *reinterpret_cast<const float *>(b);
                const float ms = __fmul_rn(p.multiplier, __uint2float_rn(acc));
                if (p.rows_a) score = __fadd_rn(ms, __fsub_rn(__fadd_rn(oa, ob), p.pair_diff));  // u8.rs:395,409
                else score = __fadd_rn(__fadd_rn(ms, oa), ob);                                    // u8.rs:347
            }
            p.out[it] = score;
        }
    }
}

qz_status launch_gather(qz_index *ix, const void *d_query, const uint32_t *d_rows_a, const uint32_t *d_rows_b,
                        uint64_t n, float *d_out) {
    if (n == 0) return QZ_OK;
    qz_ctx *ctx = ix->ctx;
    GatherParams p{};
    p.storage = ix->d_storage;
    p.row_bytes = ix->row_bytes;
    p.hdr_bytes = ix->codec == QZ_CODEC_U8 ? 4 : 0;
    p.query = (const uint8_t *)d_query;
    p.rows_a = d_rows_a;
    p.rows_b = d_rows_b;
    p.n = n;
    p.op = ix->codec == QZ_CODEC_BIN ? kOpXor : (ix->distance == QZ_L1 ? kOpL1 : kOpDot);
    p.multiplier = ix->multiplier;
    // diff = actual_dim * offset * offset, negated when inverted (u8.rs:389-394); same rounding order
    float diff = (float)ix->actual_dim * ix->offset;
    diff = diff * ix->offset;
    p.pair_diff = ix->invert ? -diff : diff;
    p.dim = ix->dim;
    const bool dot = ix->distance == QZ_DOT;
    p.bin_zx = (dot && !ix->invert) || (!dot && ix->invert);
    p.out = d_out;
    uint64_t blocks = (n + 7) / 8;
    const uint64_t cap = (uint64_t)ctx->sm_count * 16;
    if (blocks > cap) blocks = cap;
    gather_score_kernel<<<(uint32_t)blocks, 256, 0, ctx->stream>>>(p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(ctx, e, "gather_score_kernel");
    ctx->launches++;
    return QZ_OK;
}

// ---------------------------------------------------------------------------------------------
// per-CTA top-k lists from score arrays: score[i] = ((parts[0][i] + parts[1][i]) + ...)
// ---------------------------------------------------------------------------------------------
struct ScoresTopkParams {
    const float *parts;  // [n_parts][part_stride]
    uint32_t n_parts;
    uint64_t part_stride, n;
    float *sum_out;  // optional: the summed score per row
    uint64_t *partial;
    uint32_t k, Kp, flip, id_base;
};

__global__ void __launch_bounds__(256) scores_topk_kernel(ScoresTopkParams p) {
    extern __shared__ __align__(16) uint64_t sm[];
    uint64_t *top = sm;
    uint64_t *cand_all = top + p.Kp;
    uint64_t *thr = cand_all + 8 * kCandPerWarp;
    int *lock = reinterpret_cast<int *>(thr + 1);
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint32_t i = threadIdx.x; i < p.Kp; i += blockDim.x) top[i] = 0;
    if (threadIdx.x == 0) {
        *thr = 0;
        *lock = 0;
    }
    __syncthreads();
    CtaTopK st{top, thr, lock, p.k, p.Kp};
    uint64_t *cand = cand_all + warp * kCandPerWarp;
    int cnt = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t n_round = (p.n + stride - 1) / stride * stride;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool valid = i < p.n;
        float s = 0.0f;
        if (valid) {
            s = p.parts[i];
            for (uint32_t j = 1; j < p.n_parts; j++) s = __fadd_rn(s, p.parts[j * p.part_stride + i]);
            if (p.sum_out) p.sum_out[i] = s;
        }
        if (p.partial) {
            const uint64_t key = make_key(s, p.id_base + (uint32_t)i, p.flip);
            const bool pass = valid && key > *thr;
            const uint32_t m = __ballot_sync(0xffffffffu, pass);
            if (m) {
                const int n = __popc(m);
                if (cnt + n > kCandPerWarp) {
                    cta_topk_flush(st, cand, cnt, lane);
                    cnt = 0;
                }
                if (pass) cand[cnt + __popc(m & ((1u << lane) - 1))] = key;
                cnt += n;
            }
        }
    }
    if (p.partial) {
        __syncwarp();
        if (cnt > 0) cta_topk_flush(st, cand, cnt, lane);
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < p.Kp; i += blockDim.x) p.partial[(size_t)blockIdx.x * p.Kp + i] = top[i];
    }
}

qz_status launch_scores_topk(qz_ctx *ctx, const float *d_parts, uint32_t n_parts, uint64_t part_stride, uint64_t n,
                             float *d_sum_out, uint32_t k, int largest, uint32_t id_base, uint32_t *n_lists,
                             uint32_t *Kp_out) {
    const uint32_t Kp = k ? topk_pad(k) : 128;
    ScoresTopkParams p{d_parts, n_parts, part_stride, n, d_sum_out, k ? ctx->d_partial : nullptr, k ? k : 1, Kp,
                       largest ? 0u : 0xFFFFFFFFu, id_base};
    uint64_t blocks = (n + 2047) / 2048;
    if (blocks > (uint64_t)ctx->sm_count) blocks = ctx->sm_count;
    if (blocks == 0) blocks = 1;
    const size_t smem = (size_t)Kp * 8 + 8 * kCandPerWarp * 8 + 16;
    scores_topk_kernel<<<(uint32_t)blocks, 256, smem, ctx->stream>>>(p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(ctx, e, "scores_topk_kernel");
    ctx->launches++;
    if (n_lists) *n_lists = (uint32_t)blocks;
    if (Kp_out) *Kp_out = Kp;
    return QZ_OK;
}

}  // namespace qz
