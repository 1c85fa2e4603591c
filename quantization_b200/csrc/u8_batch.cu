// u8_batch.cu — batched multi-query u8 search: tcgen05 int8 contraction + fused threshold filter,
// candidate compaction, and the C-ABI entry points qz_u8_search_batch[_dev].  See u8_batch.cuh.
#include <math.h>
#include <string.h>

#include <vector>

#include "qz_internal.h"
#include "u8_batch.cuh"

namespace qz {

qz_status launch_u8_encode_query(qz_index *ix, const float *d_queries, uint32_t nq, void *d_out);

// ---------------------------------------------------------------------------------------------
// AoS rows [f32 offset][codes] -> SoA codes[rows][K] + offsets[rows]  (mirror build, query split)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) u8_split_rows_kernel(const uint8_t *aos, uint64_t rows, uint32_t K, uint8_t *codes,
                                                            float *offs) {
    const uint32_t words = K / 4, rw = words + 1;
    const uint32_t *src = reinterpret_cast<const uint32_t *>(aos);
    uint32_t *dst = reinterpret_cast<uint32_t *>(codes);
    const uint64_t total = rows * rw;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r = i / rw;
        const uint32_t w = (uint32_t)(i - r * rw);
        const uint32_t v = src[i];
        if (w == 0) offs[r] = __uint_as_float(v);
        else dst[r * words + (w - 1)] = v;
    }
}

// ---------------------------------------------------------------------------------------------
// the contraction
// ---------------------------------------------------------------------------------------------
template <bool LARGEST>
__global__ void __launch_bounds__(kBatchThreads, 1)
    u8_batch_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_v,
                    const U8BatchParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t *stage_base = smem;                                             // [stages][A 16K | B 32K], 1024-aligned
    float *voff_s = reinterpret_cast<float *>(smem + (size_t)kBatchStages * kStageBytes);  // [2][256]
    uint64_t *bars = reinterpret_cast<uint64_t *>(voff_s + 2 * kBatchN);
    uint64_t *full = bars, *empty = bars + kBatchStages, *tfull = bars + 2 * kBatchStages, *tempty = tfull + 2;
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(tempty + 2);

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t n_tiles = p.n_qb * p.n_ct;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kBatchStages; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int b = 0; b < 2; b++) {
            mbar_init(&tfull[b], 1);
            mbar_init(&tempty[b], 4);
        }
        fence_mbar_init();
    }
    if (warp == 1) {  // TMEM: all 512 columns (two 256-column accumulator buffers)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_ptr)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ---------------- TMA producer ----------------
        if (lane == 0) {
            uint32_t stage = 0, round = 0;
            for (uint32_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                const uint32_t qb = t % p.n_qb, ct = t / p.n_qb;
                const int32_t q0 = (int32_t)(qb * kBatchM), r0 = (int32_t)(p.row_begin + ct * kBatchN);
                for (uint32_t kc = 0; kc < p.k_chunks; kc++) {
                    if (round > 0) mbar_wait_bounded(&empty[stage], (round - 1) & 1);
                    uint8_t *a = stage_base + (size_t)stage * kStageBytes, *b = a + kStageA;
                    mbar_arrive_expect_tx(&full[stage], kStageBytes);
                    tma_load_2d(a, &map_q, (int32_t)(kc * kBatchKC), q0, &full[stage]);
                    tma_load_2d(b, &map_v, (int32_t)(kc * kBatchKC), r0, &full[stage]);
                    if (++stage == kBatchStages) {
                        stage = 0;
                        round++;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer ----------------
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_u8(kBatchM, kBatchN);
            uint32_t stage = 0, round = 0, it = 0;
            for (uint32_t t = blockIdx.x; t < n_tiles; t += gridDim.x, it++) {
                const uint32_t buf = it & 1, use = it >> 1;  // use = how many times this buffer was filled before
                if (use > 0) mbar_wait_bounded(&tempty[buf], (use - 1) & 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * kBatchN;
                for (uint32_t kc = 0; kc < p.k_chunks; kc++) {
                    mbar_wait_bounded(&full[stage], round & 1);
                    tc_fence_after();
                    const uint8_t *a = stage_base + (size_t)stage * kStageBytes, *b = a + kStageA;
                    const uint64_t ad = umma_desc_sw128(a), bd = umma_desc_sw128(b);
#pragma unroll
                    for (uint32_t k4 = 0; k4 < kBatchKC / 32; k4++)  // K = 32 bytes per instruction: +32 B => +2 in the address field
                        umma_i8(d_tmem, ad + 2 * k4, bd + 2 * k4, idesc, (kc | k4) != 0);
                    umma_commit(&empty[stage]);  // arrives when the MMAs above have read this stage
                    if (++stage == kBatchStages) {
                        stage = 0;
                        round++;
                    }
                }
                umma_commit(&tfull[buf]);  // accumulator complete
            }
        }
    } else {
        // ---------------- epilogue: thread = TMEM lane = one query of the block ----------------
        const uint32_t quarter = warp & 3;  // warps 2,3,4,5 -> lane quarters 2,3,0,1
        const uint32_t m = quarter * 32 + lane;
        const uint32_t et = threadIdx.x - 64;  // 0..127 among the epilogue threads
        uint32_t it = 0;
        for (uint32_t t = blockIdx.x; t < n_tiles; t += gridDim.x, it++) {
            const uint32_t qb = t % p.n_qb, ct = t / p.n_qb;
            const uint32_t buf = it & 1, use = it >> 1;
            const uint32_t q = qb * kBatchM + m;
            const bool q_ok = q < p.nq;
            const uint32_t r0 = p.row_begin + ct * kBatchN;
            const uint32_t n_valid = p.row_end - r0 < (uint32_t)kBatchN ? p.row_end - r0 : (uint32_t)kBatchN;
            // stage this tile's vector offsets (shared by all queries)
            float *vo = voff_s + buf * kBatchN;
            for (uint32_t i = et; i < (uint32_t)kBatchN; i += 128) vo[i] = i < n_valid ? __ldg(p.voff + r0 + i) : 0.0f;
            const uint64_t thr_key = q_ok ? __ldcg(p.thr + q) : ~0ull;
            const float qo = q_ok ? __ldg(p.qoff + q) : 0.0f;
            // score part of the threshold key; key 0 decodes to NaN so that everything passes
            const float thr_f = __uint_as_float(f32_from_order_bits((uint32_t)(thr_key >> 32) ^ p.flip));
            named_bar_sync(1, 128);
            mbar_wait_bounded(&tfull[buf], use & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + buf * kBatchN + ((quarter * 32) << 16);
            for (uint32_t c = 0; c < (uint32_t)kBatchN / 32; c++) {
                uint32_t r[32];
                tmem_ld32(taddr + c * 32, r);
                if (c * 32 >= n_valid) continue;  // (uniform) nothing valid in this column group
                uint32_t pass = 0;
                float sc[32];
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    // multiplier * S + query.offset + vector_offset, left to right, no FMA (u8.rs:347)
                    const float s = __fadd_rn(__fadd_rn(__fmul_rn(p.multiplier, __int2float_rn((int)r[j])), qo), vo[c * 32 + j]);
                    sc[j] = s;
                    const bool ok = LARGEST ? !(s < thr_f) : !(s > thr_f);
                    pass |= ok ? (1u << j) : 0u;
                }
                if (!q_ok) pass = 0;
                const uint32_t left = n_valid - c * 32;
                if (left < 32) pass &= (1u << left) - 1;
                while (pass) {
                    const int j = __ffs(pass) - 1;
                    pass &= pass - 1;
                    float s = sc[0];
#pragma unroll
                    for (int jj = 1; jj < 32; jj++) s = (j == jj) ? sc[jj] : s;
                    const uint64_t key = make_key(s, p.id_base + r0 + c * 32 + j, p.flip);
                    if (key > thr_key) {
                        const uint32_t pos = atomicAdd(p.cnt + q, 1u);
                        if (pos < p.cap) p.cand[(size_t)q * p.cap + pos] = key;
                        else *p.overflow = 1u;
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[buf]);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
    }
}

// ---------------------------------------------------------------------------------------------
// Candidate compaction, one CTA per query: sort the staged keys (descending), keep the best k, publish the
// k-th as the query's new threshold; the last call also decodes the result rows.
// ---------------------------------------------------------------------------------------------
struct CompactParams {
    uint64_t *cand;
    uint32_t *cnt;
    uint64_t *thr;
    uint32_t cap, k, flip;
    int final_out;
    float *scores;  // [nq][k]
    uint32_t *ids;  // [nq][k]
};

__global__ void __launch_bounds__(512) u8_batch_compact_kernel(CompactParams p) {
    extern __shared__ __align__(16) uint64_t ks[];
    const uint32_t q = blockIdx.x, tid = threadIdx.x, T = blockDim.x;
    uint64_t *cq = p.cand + (size_t)q * p.cap;
    uint32_t n = p.cnt[q];
    if (n > p.cap) n = p.cap;
    if (!p.final_out && n <= p.k) return;  // nothing to drop yet; threshold stays
    uint32_t P = 32;
    while (P < n) P <<= 1;
    for (uint32_t i = tid; i < P; i += T) ks[i] = i < n ? cq[i] : 0ull;
    __syncthreads();
    for (uint32_t size = 2; size <= P; size <<= 1) {
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            for (uint32_t t = tid; t < (P >> 1); t += T) {
                const uint32_t lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                const bool desc = (lo & size) == 0 || size == P;
                const uint64_t x = ks[lo], y = ks[hi];
                if (desc ? (x < y) : (x > y)) {
                    ks[lo] = y;
                    ks[hi] = x;
                }
            }
            __syncthreads();
        }
    }
    const uint32_t keep = n < p.k ? n : p.k;
    for (uint32_t i = tid; i < keep; i += T) cq[i] = ks[i];
    if (tid == 0) {
        p.cnt[q] = keep;
        p.thr[q] = n >= p.k ? ks[p.k - 1] : 0ull;
    }
    if (p.final_out) {
        for (uint32_t i = tid; i < p.k; i += T) {
            float sc;
            uint32_t id;
            split_key(i < keep ? ks[i] : 0ull, p.flip, sc, id);
            p.scores[(size_t)q * p.k + i] = sc;
            p.ids[(size_t)q * p.k + i] = id;
        }
    }
}

}  // namespace qz

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
using namespace qz;

typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_tiled() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(p);
    }
    return fn;
}

// 2-D u8 tensor [rows][K] (K contiguous), box = 128 bytes x box_rows, 128-byte swizzle, OOB reads give 0
static bool make_map(CUtensorMap *map, const void *base, uint64_t rows, uint32_t K, uint32_t box_rows) {
    PFN_encodeTiled enc = get_encode_tiled();
    if (!enc) return false;
    cuuint64_t dims[2] = {K, rows};
    cuuint64_t strides[1] = {K};
    cuuint32_t box[2] = {(cuuint32_t)kBatchKC, box_rows};
    cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static qz_status batch_search(qz_index *ix, const float *d_queries, uint32_t nq, uint32_t k, int largest, float *d_scores,
                              uint32_t *d_ids) {
    qz_ctx *ctx = ix->ctx;
    const uint32_t K = ix->actual_dim;
    const uint64_t N = ix->count;
    // ---- SoA mirror of the corpus (once) ----
    if (!ix->d_codes_soa && N) {
        QZ_CUDA(ctx, cudaMalloc((void **)&ix->d_codes_soa, align_up((size_t)N * K, 256)));
        QZ_CUDA(ctx, cudaMalloc((void **)&ix->d_voff_soa, align_up((size_t)N * 4, 256)));
        u8_split_rows_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(ix->d_storage, N, K, ix->d_codes_soa, ix->d_voff_soa);
        QZ_CUDA(ctx, cudaGetLastError());
        ctx->launches++;
    }
    // ---- scratch arena ----
    const uint32_t cap = k <= 256 ? 4096u : 8192u;
    const size_t o_blob = 0, o_qcodes = o_blob + align_up((size_t)nq * (K + 4), 256);
    const size_t o_qoff = o_qcodes + align_up((size_t)nq * K, 256), o_thr = o_qoff + align_up((size_t)nq * 4, 256);
    const size_t o_cnt = o_thr + align_up((size_t)nq * 8, 256), o_ovf = o_cnt + align_up((size_t)nq * 4, 256);
    const size_t o_cand = o_ovf + 256, total = o_cand + (size_t)nq * cap * 8;
    if (ix->batch_bytes < total) {
        QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        cudaFree(ix->d_batch);
        ix->d_batch = nullptr;
        ix->batch_bytes = 0;
        QZ_CUDA(ctx, cudaMalloc((void **)&ix->d_batch, total));
        ix->batch_bytes = total;
    }
    uint8_t *a = ix->d_batch;
    uint8_t *d_blob = a + o_blob, *d_qcodes = a + o_qcodes;
    float *d_qoff = reinterpret_cast<float *>(a + o_qoff);
    uint64_t *d_thr = reinterpret_cast<uint64_t *>(a + o_thr), *d_cand = reinterpret_cast<uint64_t *>(a + o_cand);
    uint32_t *d_cnt = reinterpret_cast<uint32_t *>(a + o_cnt), *d_ovf = reinterpret_cast<uint32_t *>(a + o_ovf);
    // ---- encode + split the queries ----
    QZ_TRY(launch_u8_encode_query(ix, d_queries, nq, d_blob));
    u8_split_rows_kernel<<<(nq * (K / 4 + 1) + 255) / 256, 256, 0, ctx->stream>>>(d_blob, nq, K, d_qcodes, d_qoff);
    QZ_CUDA(ctx, cudaGetLastError());
    ctx->launches++;

    static bool attr_done = false;
    if (!attr_done) {
        QZ_CUDA(ctx, cudaFuncSetAttribute(u8_batch_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBatchSmem));
        QZ_CUDA(ctx, cudaFuncSetAttribute(u8_batch_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBatchSmem));
        QZ_CUDA(ctx, cudaFuncSetAttribute(u8_batch_compact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 8));
        attr_done = true;
    }
    CUtensorMap map_q, map_v;
    if (!make_map(&map_q, d_qcodes, nq, K, kBatchM) || (N && !make_map(&map_v, ix->d_codes_soa, N, K, kBatchN)))
        return fail(ctx, QZ_ENCODING, "cuTensorMapEncodeTiled failed");
    if (!N) memset(&map_v, 0, sizeof map_v);

    U8BatchParams p{};
    p.nq = nq;
    p.k_chunks = (K + kBatchKC - 1) / kBatchKC;
    p.n_qb = (nq + kBatchM - 1) / kBatchM;
    p.qoff = d_qoff;
    p.voff = ix->d_voff_soa;
    p.multiplier = ix->multiplier;
    p.id_base = ix->id_base;
    p.flip = largest ? 0u : 0xFFFFFFFFu;
    p.cap = cap;
    p.thr = d_thr;
    p.cand = d_cand;
    p.cnt = d_cnt;
    p.overflow = d_ovf;
    CompactParams cp{d_cand, d_cnt, d_thr, cap, k, p.flip, 0, d_scores, d_ids};

    // Chunk schedule.  The first chunk is small enough that every row fits the candidate buffer; after a
    // compaction at n rows the threshold is the k-th best of n, so (for rows in random order) a chunk of R rows
    // adds about R*k/n candidates: grow geometrically with an 8x margin.  `safe` = every chunk <= cap - k rows,
    // which can never overflow whatever the order; it is the fallback when the fast schedule overflowed.
    for (int safe = 0; safe < 2; safe++) {
        QZ_CUDA(ctx, cudaMemsetAsync(a + o_thr, 0, o_cand - o_thr, ctx->stream));  // thr, cnt, overflow
        uint64_t done = 0;
        const uint64_t first = (cap - (safe ? k : 0)) / kBatchN * kBatchN ? (cap - (safe ? k : 0)) / kBatchN * kBatchN : kBatchN;
        do {
            uint64_t R = first;
            if (!safe && done) {
                const uint64_t grow = done * cap / (8ull * k);
                R = grow > first ? grow / kBatchN * kBatchN : first;
            }
            if (safe && done) R = (cap - k) / kBatchN * kBatchN ? (cap - k) / kBatchN * kBatchN : kBatchN;
            if (done + R > N) R = N - done;
            const bool last = done + R >= N;
            if (R) {
                p.row_begin = (uint32_t)done;
                p.row_end = (uint32_t)(done + R);
                p.n_ct = (uint32_t)((R + kBatchN - 1) / kBatchN);
                const uint32_t tiles = p.n_qb * p.n_ct;
                const uint32_t grid = tiles < (uint32_t)ctx->sm_count ? tiles : (uint32_t)ctx->sm_count;
                timing_begin(ctx);
                if (largest) u8_batch_kernel<true><<<grid, kBatchThreads, kBatchSmem, ctx->stream>>>(map_q, map_v, p);
                else u8_batch_kernel<false><<<grid, kBatchThreads, kBatchSmem, ctx->stream>>>(map_q, map_v, p);
                timing_end(ctx);
                QZ_CUDA(ctx, cudaGetLastError());
                ctx->launches++;
            }
            cp.final_out = last;
            u8_batch_compact_kernel<<<nq, 512, (size_t)cap * 8, ctx->stream>>>(cp);
            QZ_CUDA(ctx, cudaGetLastError());
            ctx->launches++;
            done += R;
        } while (done < N);
        uint32_t ovf = 0;
        QZ_CUDA(ctx, cudaMemcpyAsync(&ovf, d_ovf, 4, cudaMemcpyDeviceToHost, ctx->stream));
        QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (!ovf) return QZ_OK;
    }
    return fail(ctx, QZ_ENCODING, "batched search: candidate buffer overflow in safe mode");
}

extern "C" {

qz_status qz_u8_search_batch_dev(qz_index *ix, const float *d_queries, uint32_t nq, uint32_t k, int32_t largest,
                                 float *d_scores, uint32_t *d_ids) {
    if (!ix || !ix->ctx) return QZ_ARGUMENTS;
    qz_ctx *ctx = ix->ctx;
    if (ix->codec != QZ_CODEC_U8) return fail(ctx, QZ_ARGUMENTS, "search_batch: u8 index required");
    if (ix->distance == QZ_L1)
        return fail(ctx, QZ_ARGUMENTS, "search_batch: L1 is not a contraction; use the single-query scan");
    if (k == 0 || k > QZ_MAX_K) return fail(ctx, QZ_ARGUMENTS, "search_batch: k must be in [1, QZ_MAX_K]");
    if (nq == 0) return QZ_OK;
    if (!d_queries || !d_scores || !d_ids) return fail(ctx, QZ_ARGUMENTS, "search_batch: NULL argument");
    if (ix->count >= 0xFFFFFF00ull) return fail(ctx, QZ_ARGUMENTS, "search_batch: too many rows");
    std::lock_guard<std::mutex> g(ctx->mu);
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    return batch_search(ix, d_queries, nq, k, largest, d_scores, d_ids);
}

qz_status qz_u8_search_batch(qz_index *ix, const float *queries, uint32_t nq, uint32_t k, int32_t largest, float *scores,
                             uint32_t *ids) {
    if (!ix || !ix->ctx) return QZ_ARGUMENTS;
    qz_ctx *ctx = ix->ctx;
    if (nq == 0) return QZ_OK;
    if (!queries || !scores || !ids) return fail(ctx, QZ_ARGUMENTS, "search_batch: NULL argument");
    if (k == 0 || k > QZ_MAX_K) return fail(ctx, QZ_ARGUMENTS, "search_batch: k must be in [1, QZ_MAX_K]");
    const size_t qb = (size_t)nq * ix->dim * 4, rb = (size_t)nq * k * 4;
    uint8_t *d = nullptr;
    {
        std::lock_guard<std::mutex> g(ctx->mu);
        QZ_CUDA(ctx, cudaSetDevice(ctx->device));
        QZ_CUDA(ctx, cudaMalloc((void **)&d, align_up(qb, 256) + 2 * align_up(rb, 256)));
        QZ_CUDA(ctx, cudaMemcpyAsync(d, queries, qb, cudaMemcpyHostToDevice, ctx->stream));
    }
    float *d_s = reinterpret_cast<float *>(d + align_up(qb, 256));
    uint32_t *d_i = reinterpret_cast<uint32_t *>(d + align_up(qb, 256) + align_up(rb, 256));
    qz_status st = qz_u8_search_batch_dev(ix, reinterpret_cast<const float *>(d), nq, k, largest, d_s, d_i);
    cudaError_t e = cudaSuccess;
    if (st == QZ_OK) {
        e = cudaMemcpyAsync(scores, d_s, rb, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(ids, d_i, rb, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    }
    cudaFree(d);
    if (st != QZ_OK) return st;
    if (e != cudaSuccess) return cuda_fail(ctx, e, "search_batch copy");
    return QZ_OK;
}

}  // extern "C"
