// u8_batch.cu — batched multi-query u8 search: tcgen05 int8 contraction + fused threshold filter,
// candidate compaction, and the C-ABI entry points qz_u8_search_batch[_dev].  See u8_batch.cuh.
#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "qz_internal.h"
#include "u8_batch.cuh"
#include "tensor_map.h"
#include "topk.cuh"

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

// max |x| over n floats (non-negative floats order like their bit patterns); NaN offsets poison the bound to +inf
__global__ void __launch_bounds__(256) absmax_kernel(const float *x, uint64_t n, uint32_t *out) {
    uint32_t m = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t b = __float_as_uint(x[i]) & 0x7FFFFFFFu;
        if (b > 0x7F800000u) b = 0x7F800000u;
        m = b > m ? b : m;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t v = __shfl_xor_sync(0xffffffffu, m, o);
        m = v > m ? v : m;
    }
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

// vq[i] = floor(voff[i] / multiplier) computed in double: the vector offset of row i in units of one step of the
// integer sum S (score = multiplier * S + query offset + vector offset).  |voff / multiplier| < 2^30 is checked by
// the host before this image is used; the double quotient is within 1e-6 of the real one, so
// vq - 1e-6 <= voff / multiplier <= vq + 1 + 1e-6.
__global__ void __launch_bounds__(256) u8_vq_kernel(const float *voff, uint64_t n, double multiplier, int32_t *vq) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        vq[i] = (int32_t)floor((double)voff[i] / multiplier);
}

// ---------------------------------------------------------------------------------------------
// the contraction
// ---------------------------------------------------------------------------------------------
// Drain one epilogue warp's candidate queue: every lane appends its share to the per-query buffers, so the whole
// warp pays one counter round trip per 32 candidates.
// (Arguments by value: taking the address of the kernel parameter block would move it to local memory, and with
// 227 KB of shared memory carved out there is practically no L1 to hold it.)
__device__ __noinline__ void queue_flush(uint32_t *cnt, uint64_t *cand, uint32_t cap, uint32_t *overflow,
                                         const uint64_t *qkeys, const uint32_t *qq, uint32_t n, uint32_t lane) {
    __syncwarp();
    // all counter round trips of a flush are in flight together (the rolled loop paid them one after another: up to
    // kQueueCap / 32 = 8 L2 atomics of ~1 us per flush, which paced the small early chunks where most of what a tile
    // produces is a candidate)
    constexpr int kRounds = kQueueCap / 32;
    uint32_t q[kRounds], pos[kRounds];
#pragma unroll
    for (int i = 0; i < kRounds; i++) {
        const uint32_t e = lane + 32 * i;
        q[i] = e < n ? qq[e] : 0u;
        pos[i] = e < n ? atomicAdd(cnt + q[i], 1u) : 0u;
    }
#pragma unroll
    for (int i = 0; i < kRounds; i++) {
        const uint32_t e = lane + 32 * i;
        if (e < n) {
            if (pos[i] < cap) cand[(size_t)q[i] * cap + pos[i]] = qkeys[e];
            else *overflow = 1u;
        }
    }
    __syncwarp();
}

// Filter modes.  The epilogue has to look at Q x N accumulators and keep ~k * ln(N) of them per query, so what it
// costs per accumulator decides whether the tensor pipe or the epilogue paces the kernel (v1: I2F + FFMA + FSETP +
// mask update per element = 4 issue slots, I2F at a quarter of the FP32 rate: the MMA warp waited for TMEM buffers,
// tensor pipe 62 % busy).
//   kFiltIntMax / kFiltIntMin   one VIADDMNMX per accumulator.  With M = multiplier, the exact test
//       fl(fl(fl(M*S) + qo) + vo_j) >= thr  (largest; <= for smallest) implies, in real arithmetic,
//       M*S + vo_j >= thr - qo - margin, i.e. for M > 0:  S + vo_j/M >= y := (thr - qo - margin) / M, and with the
//       integer image F_j = floor(vo_j / M) (u8_vq_kernel):  S + F_j >= floor(y) - G - 2, G bounding the f32
//       rounding of y.  Per query (lane) and tile the right-hand side is ONE integer A; the lane folds
//       max_j(S_j + F_j) over 8 columns at a time with VIADDMNMX and compares the maximum with A.  M < 0 or
//       smallest-first flip the inequality (min instead of max).  The filter is conservative (never drops a true
//       candidate).  Only when some lane of the warp has a chain maximum at or beyond its bound are the per-column
//       tests of that chain evaluated (8 x IADD + ISETP + mask bit); the flagged columns are then re-read from TMEM
//       one at a time (uniform address, rolled loop: no dynamically indexed registers, no local memory — with
//       227 KB of shared memory carved out the L1 is ~1 KB and every stack access is an L2 round trip) and
//       re-scored exactly.
//   kFiltFloat   the float form of the same test (one I2F + FFMA + FSETP + mask bit per element) for indexes whose
//       metadata cannot be mapped to integers: multiplier 0 / non-finite, |vo / M| >= 2^30, non-finite offsets.
constexpr int kFiltFloat = 0, kFiltIntMax = 1, kFiltIntMin = 2;

template <int MODE>
__device__ __forceinline__ int viaddmm(int a, int b, int c) {
    return MODE == kFiltIntMax ? __viaddmax_s32(a, b, c) : __viaddmin_s32(a, b, c);
}

// CL = CTAs per cluster.  The CL CTAs of a cluster work on CL consecutive query blocks against the SAME corpus
// tile: each loads its own A block and 1/CL of the B block, multicast to all of them, which divides the
// L2 -> SM traffic of B by CL (the kernel is L2-bandwidth bound without it).
template <int MODE, int CL>
__global__ void __launch_bounds__(kBatchThreads, 1)
    u8_batch_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_v,
                    const U8BatchParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);      // 1024-aligned, still "shared"
    uint8_t *stage_base = smem;                                                       // [stages][A 16K | B 32K]
    float *voff_s = reinterpret_cast<float *>(smem + (size_t)kBatchStages * kStageBytes);  // [kVoffBufs][256] f32
    int *vq_s = reinterpret_cast<int *>(voff_s + kVoffBufs * kBatchN);                     // [kVoffBufs][256] s32
    uint64_t *bars = reinterpret_cast<uint64_t *>(vq_s + kVoffBufs * kBatchN);
    uint64_t *full = bars, *empty = bars + kBatchStages, *tfull = bars + 2 * kBatchStages, *tempty = tfull + 2;
    uint64_t *vfull = tempty + 2, *vempty = vfull + kVoffBufs;
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(vempty + kVoffBufs);
    uint64_t *queue_base = reinterpret_cast<uint64_t *>(bars + 32);  // [kEpiWarps][kQueueCap] keys, then query ids

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // work items: (corpus tile, group of CL query blocks); cluster c takes items c, c + n_clusters, ...
    const uint32_t rank = CL > 1 ? cluster_ctarank() : 0u;
    const uint32_t cluster_id = blockIdx.x / CL, n_clusters = gridDim.x / CL;
    const uint32_t qgroups = p.n_qb / CL;  // n_qb is padded to a multiple of CL by the host
    const uint32_t n_items = qgroups * p.n_ct;
    constexpr uint16_t kMask = (uint16_t)((1u << CL) - 1);
    constexpr uint32_t kVBytes = MODE == kFiltFloat ? kBatchN * 4 : 2 * kBatchN * 4;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kBatchStages; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], CL);  // one tcgen05.commit arrival from every CTA of the cluster
        }
        for (int b = 0; b < 2; b++) {
            mbar_init(&tfull[b], 1);
            mbar_init(&tempty[b], kEpiWarps);
        }
        for (int b = 0; b < kVoffBufs; b++) {
            mbar_init(&vfull[b], 1);
            mbar_init(&vempty[b], kEpiWarps);
        }
        fence_mbar_init();
    }
    if (warp == 1) {  // TMEM: all 512 columns (two 256-column accumulator buffers)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_ptr)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (CL > 1) cluster_sync_all();  // every CTA's barriers exist before any remote arrival or multicast write
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    // The two single-thread roles run their loops with the WHOLE warp (uniform control flow, loop state in uniform
    // registers) and elect one lane only for the TMA / tcgen05 instructions themselves.  With `if (lane == 0)`
    // around the loops every UTCIMMA cost an ELECT + R2UR.BROADCAST + 64-bit descriptor add in a divergent thread,
    // ~17 instructions per MMA: the issuing thread itself paced the tensor pipe (it was never seen waiting).
    if (warp == 0) {
        // ---------------- TMA producer ----------------
        uint32_t stage = 0, round = 0, it = 0;
        uint32_t qg = cluster_id % qgroups, ct = cluster_id / qgroups;
        const uint32_t step_q = n_clusters % qgroups, step_c = n_clusters / qgroups;
        for (uint32_t t = cluster_id; t < n_items; t += n_clusters, it++) {
            const int32_t q0 = (int32_t)((qg * CL + rank) * kBatchM), r0 = (int32_t)(p.row_begin + ct * kBatchN);
            qg += step_q;
            ct += step_c;
            if (qg >= qgroups) {
                qg -= qgroups;
                ct++;
            }
            {   // this tile's 256 vector offsets and their integer images (1 KB each; the arrays are padded)
                // through their own small ring
                const uint32_t vb = it % kVoffBufs, vuse = it / kVoffBufs;
                if (vuse > 0) mbar_wait_bounded(&vempty[vb], (vuse - 1) & 1);
                if (elect_one()) {
                    mbar_arrive_expect_tx(&vfull[vb], kVBytes);
                    bulk_g2s(voff_s + vb * kBatchN, p.voff + r0, kBatchN * 4, &vfull[vb]);
                    if (MODE != kFiltFloat) bulk_g2s(vq_s + vb * kBatchN, p.vq + r0, kBatchN * 4, &vfull[vb]);
                }
                __syncwarp();
            }
            for (uint32_t kc = 0; kc < p.k_chunks; kc++) {
                if (round > 0) mbar_wait_bounded(&empty[stage], (round - 1) & 1);  // free in EVERY CTA of the cluster
                if (elect_one()) {
                    uint8_t *a = stage_base + (size_t)stage * kStageBytes, *b = a + kStageA;
                    mbar_arrive_expect_tx(&full[stage], kStageBytes);
                    tma_load_2d(a, &map_q, (int32_t)(kc * kBatchKC), q0, &full[stage]);
                    if (CL == 1) {
                        tma_load_2d(b, &map_v, (int32_t)(kc * kBatchKC), r0, &full[stage]);
                    } else {  // this CTA's 1/CL of the rows, delivered to all CL CTAs
                        constexpr uint32_t part = kBatchN / CL;
                        tma_load_2d_mc(b + (size_t)rank * part * kBatchKC, &map_v, (int32_t)(kc * kBatchKC),
                                       r0 + (int32_t)(rank * part), &full[stage], kMask);
                    }
                }
                __syncwarp();
                if (++stage == kBatchStages) {
                    stage = 0;
                    round++;
                }
            }
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer ----------------
        constexpr uint32_t idesc = umma_idesc_u8(kBatchM, kBatchN);
        uint32_t stage = 0, round = 0, it = 0;
        for (uint32_t t = cluster_id; t < n_items; t += n_clusters, it++) {
            const uint32_t buf = it & 1, use = it >> 1;  // use = how many times this buffer was filled before
            if (use > 0) mbar_wait_bounded(&tempty[buf], (use - 1) & 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + buf * kBatchN;
            for (uint32_t kc = 0; kc < p.k_chunks; kc++) {
                mbar_wait_bounded(&full[stage], round & 1);
                tc_fence_after();
                if (elect_one()) {  // (tcgen05.commit tracks the MMAs of the executing thread: same lane for both)
                    const uint8_t *a = stage_base + (size_t)stage * kStageBytes;
                    const uint32_t ad = umma_desc_sw128_lo(a), bd = umma_desc_sw128_lo(a + kStageA);
#pragma unroll
                    for (uint32_t k4 = 0; k4 < kBatchKC / 32; k4++)  // K = 32 bytes per instruction: +32 B => +2 in the address field
                        umma_i8_lo(d_tmem, ad + 2 * k4, bd + 2 * k4, kUmmaDescSw128Hi, idesc, (kc | k4) != 0);
                    // arrives (in every CTA of the cluster) when the MMAs above have read this stage
                    if (CL == 1) umma_commit(&empty[stage]);
                    else umma_commit_mc(&empty[stage], kMask);
                    if (kc + 1 == p.k_chunks) umma_commit(&tfull[buf]);  // accumulator complete
                }
                __syncwarp();
                if (++stage == kBatchStages) {
                    stage = 0;
                    round++;
                }
            }
        }
    } else {
        // ---------------- epilogue: thread = TMEM lane = one query; two warps per lane quarter ----------------
        // Warp e = warp - 2 reads TMEM lanes [32 * (warp % 4), +32) (the hardware ties a warp to that quarter) and
        // takes columns [128 * (e / 4), +128) of the accumulator buffer: four groups of 32, two register sets.
        // The MMA of tile t + 2 needs ALL epilogue warps to have released tile t's buffer, so the slowest warp paces
        // the tensor pipe.  The buffer is therefore released as soon as the last group has been LOADED (after two of
        // the four groups have been filtered), not when the tile is finished: everything after the loads works on
        // registers, including the re-scoring of flagged accumulators.
        // (Tried and measured slower: 16 epilogue warps x 2 groups — 96 registers per thread, spills, and five
        // warps per scheduler competing with the MMA-issuing warp: 3.10 ms per C5 step against 3.02.)
        const uint32_t quarter = warp & 3;
        const uint32_t slice = (warp - 2) >> 2;  // 0..1
        const uint32_t m = quarter * 32 + lane;
        const bool largest = p.flip == 0;
        const float kmax = fabsf(p.multiplier) * 16129.0f * (float)(p.k_chunks * kBatchKC);  // |multiplier * S| bound
        uint64_t *qkeys = queue_base + (size_t)(warp - 2) * kQueueCap;                       // this warp's candidate queue
        uint32_t *qq = reinterpret_cast<uint32_t *>(queue_base + kEpiWarps * kQueueCap) + (size_t)(warp - 2) * kQueueCap;
        uint32_t qcount = 0;
        uint32_t it = 0;
        // (query group, corpus tile) of item t without a division per tile
        const uint32_t step_q = n_clusters % qgroups, step_c = n_clusters / qgroups;
        uint32_t qg = cluster_id % qgroups, ct = cluster_id / qgroups;
        // per-query state of the NEXT tile is fetched while the current one is processed
        uint64_t thr_next = ~0ull;
        float qo_next = 0.0f;
        if (cluster_id < n_items) {
            const uint32_t q = (qg * CL + rank) * kBatchM + m;
            if (q < p.nq) {
                thr_next = __ldcg(p.thr + q);
                qo_next = __ldg(p.qoff + q);
            }
        }
        for (uint32_t t = cluster_id; t < n_items; t += n_clusters, it++) {
            const uint32_t buf = it & 1, use = it >> 1;
            const uint32_t q = (qg * CL + rank) * kBatchM + m;
            const bool q_ok = q < p.nq;
            const uint32_t r0 = p.row_begin + ct * kBatchN;
            const uint32_t n_valid = p.row_end - r0 < (uint32_t)kBatchN ? p.row_end - r0 : (uint32_t)kBatchN;
            const uint32_t vb = it % kVoffBufs, vuse = it / kVoffBufs;
            const float *vo = voff_s + vb * kBatchN;
            const uint64_t thr_key = thr_next;  // ~0 for a lane without a query: nothing beats it
            const float qo = qo_next;
            {   // next item of this cluster
                qg += step_q;
                ct += step_c;
                if (qg >= qgroups) {
                    qg -= qgroups;
                    ct++;
                }
                thr_next = ~0ull;
                qo_next = 0.0f;
                if (t + n_clusters < n_items) {
                    const uint32_t qn = (qg * CL + rank) * kBatchM + m;
                    if (qn < p.nq) {
                        thr_next = __ldcg(p.thr + qn);
                        qo_next = __ldg(p.qoff + qn);
                    }
                }
            }
            // score part of the threshold key; key 0 decodes to NaN
            const float thr_f = __uint_as_float(f32_from_order_bits((uint32_t)(thr_key >> 32) ^ p.flip));
            // Loosening of the threshold: far more than the rounding difference between real arithmetic and the
            // reference's three roundings (each <= 2^-24 of the magnitudes involved; the margin is 2^-20 of their sum).
            const float mag = kmax + fabsf(qo) + p.voff_absmax + fabsf(thr_f);
            const float margin = mag * 9.5367431640625e-7f;
            const float cthr = largest ? (thr_f - qo) - margin : (thr_f - qo) + margin;
            const float thr2 = cthr;  // float filter: multiplier * S + vo against this (NaN when there is no threshold yet)
            int A = 0;                // integer filter: S + vq against this
            if (MODE != kFiltFloat) {
                constexpr int kAlways = MODE == kFiltIntMax ? INT_MIN : INT_MAX, kNever = MODE == kFiltIntMax ? INT_MAX : INT_MIN;
                float y = cthr * p.rmult;
                const float G = mag * fabsf(p.rmult) * 4.76837158203125e-7f + 3.0f;  // f32 rounding of y, in steps of S
                if (!q_ok) A = kNever;
                else if (thr_key == 0ull || !(y == y) || !(G < 1.0e6f)) A = kAlways;
                else {
                    y = fminf(fmaxf(y, -1.6e9f), 1.6e9f);  // |S + vq| < 2^24 + 2^30: beyond +-1.6e9 nothing / everything passes
                    const int g = (int)G + 2;
                    A = MODE == kFiltIntMax ? __float2int_rd(y) - g : __float2int_rd(y) + g + 1;
                }
            }
            mbar_wait_bounded(&tfull[buf], use & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + buf * kBatchN + slice * 128 + ((quarter * 32) << 16);
            uint32_t rs[2][32];  // two register sets: groups 0 / 1, then 2 / 3
            tmem_ld32_nowait(taddr, rs[0]);
            tmem_ld32_nowait(taddr + 32, rs[1]);
            mbar_wait_bounded(&vfull[vb], vuse & 1);
#pragma unroll 1
            for (uint32_t hh = 0; hh < 2; hh++) {  // rolled: two filter bodies are the whole hot path (instruction cache)
              tmem_ld_wait(rs[0]);
              tmem_ld_wait(rs[1]);
#pragma unroll
              for (uint32_t ss = 0; ss < 2; ss++) {
                uint32_t(&rc)[32] = rs[ss];
                const uint32_t c = 2 * hh + ss;
                const uint32_t col0 = slice * 128 + c * 32;
                uint32_t mask = 0;  // bit j = column col0 + j may beat this lane's threshold
                const uint32_t left = col0 < n_valid ? n_valid - col0 : 0u;
#ifdef QZ_TUNING
                if (!(p.dbg & 1))
#endif
                if (left) {  // (uniform) something valid in this column group
                if (MODE == kFiltFloat) {
#pragma unroll
                    for (int j = 0; j < 32; j++) {
                        const float x = fmaf(p.multiplier, __int2float_rn((int)rc[j]), vo[col0 + j]);
                        const bool ok = largest ? !(x < thr2) : !(x > thr2);
                        mask |= ok ? (1u << j) : 0u;
                    }
                    if (!q_ok) mask = 0;
                } else {
                    // integer images of the vector offsets (broadcast LDS.128, 16 columns at a time to keep the
                    // register count down), folded into 4 independent VIADDMNMX chains
                    const int *vq = vq_s + vb * kBatchN + col0;
                    constexpr int kInit = MODE == kFiltIntMax ? INT_MIN : INT_MAX;
                    int a[4] = {kInit, kInit, kInit, kInit};
#pragma unroll
                    for (int hb = 0; hb < 2; hb++) {
                        int v[16];
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const int4 t4 = reinterpret_cast<const int4 *>(vq)[4 * hb + i];
                            v[4 * i] = t4.x;
                            v[4 * i + 1] = t4.y;
                            v[4 * i + 2] = t4.z;
                            v[4 * i + 3] = t4.w;
                        }
#pragma unroll
                        for (int i = 0; i < 4; i++) {
#pragma unroll
                            for (int s = 0; s < 4; s++) a[s] = viaddmm<MODE>((int)rc[16 * hb + 4 * i + s], v[4 * i + s], a[s]);
                        }
                    }
                    bool h[4];
#pragma unroll
                    for (int s = 0; s < 4; s++) h[s] = MODE == kFiltIntMax ? a[s] >= A : a[s] <= A;
                    if (__any_sync(0xffffffffu, h[0] | h[1] | h[2] | h[3])) {
                        // some lane has a column at or beyond its bound in some chain: per-column tests of the
                        // flagged chains only (uniform branches)
#pragma unroll
                        for (int s = 0; s < 4; s++) {
                            if (!__any_sync(0xffffffffu, h[s])) continue;
#pragma unroll
                            for (int i = 0; i < 8; i++) {
                                const int j = 4 * i + s;
                                const int sv = (int)rc[j] + vq[j];
                                const bool ok = MODE == kFiltIntMax ? sv >= A : sv <= A;
                                mask |= ok ? (1u << j) : 0u;
                            }
                        }
                    }
                }
                if (left < 32) mask &= (1u << left) - 1;
                }
#ifdef QZ_TUNING
                if (p.dbg & 2) mask = 0;
#endif
                // Flagged accumulators are re-scored exactly, ONE PER LANE per round: every lane picks its own
                // lowest flagged column, fetches that accumulator from its registers through a 5-level select tree
                // (31 SEL; registers cannot be indexed dynamically), forms the reference score and key and, if the
                // key beats the lane's threshold, appends it to the warp's shared-memory queue (one ballot per
                // round).  Rounds = the largest number of flagged columns of any lane (1 in the steady state).
                while (__any_sync(0xffffffffu, mask != 0)) {  // rolled on purpose (instruction cache)
                    const bool has = mask != 0;
                    const uint32_t j = has ? __ffs(mask) - 1 : 0u;
                    mask &= mask - 1;
                    uint32_t s16[16], s8[8], s4[4], s2[2];
#pragma unroll
                    for (int i = 0; i < 16; i++) s16[i] = (j & 1) ? rc[2 * i + 1] : rc[2 * i];
#pragma unroll
                    for (int i = 0; i < 8; i++) s8[i] = (j & 2) ? s16[2 * i + 1] : s16[2 * i];
#pragma unroll
                    for (int i = 0; i < 4; i++) s4[i] = (j & 4) ? s8[2 * i + 1] : s8[2 * i];
#pragma unroll
                    for (int i = 0; i < 2; i++) s2[i] = (j & 8) ? s4[2 * i + 1] : s4[2 * i];
                    const int S = (int)((j & 16) ? s2[1] : s2[0]);
                    const float vj = vo[col0 + j];
                    // multiplier * S + query.offset + vector_offset, left to right, no FMA (u8.rs:347)
                    const float sx = __fadd_rn(__fadd_rn(__fmul_rn(p.multiplier, __int2float_rn(S)), qo), vj);
                    const uint64_t key = make_key(sx, p.id_base + r0 + col0 + j, p.flip);
                    const bool ok = has && key > thr_key;
                    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
                    if (bal) {
                        if (qcount + 32 > (uint32_t)kQueueCap) {
                            queue_flush(p.cnt, p.cand, p.cap, p.overflow, qkeys, qq, qcount, lane);
                            qcount = 0;
                        }
                        if (ok) {
                            const uint32_t w = qcount + __popc(bal & ((1u << lane) - 1));
                            qkeys[w] = key;
                            qq[w] = q;
                        }
                        qcount += __popc(bal);
                    }
                }
                if (hh == 0) {
                    // this register set is free: fetch its next group; after the second fetch every accumulator of
                    // the tile is in registers and the MMA warp may overwrite the buffer
                    tmem_ld32_nowait(taddr + (2 + ss) * 32, rs[ss]);
                    if (ss == 1) {
                        tmem_ld_wait(rs[0]);
                        tmem_ld_wait(rs[1]);
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&tempty[buf]);
                    }
                }
              }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&vempty[vb]);
        }
        queue_flush(p.cnt, p.cand, p.cap, p.overflow, qkeys, qq, qcount, lane);
    }

    tc_fence_before();
    __syncthreads();
    if (CL > 1) cluster_sync_all();  // no CTA leaves while a peer may still write its shared memory or barriers
    if (warp == 1) {
        __syncwarp();
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

// bitonic sort of P (power of two) keys in shared memory, descending, by the whole CTA
__device__ __forceinline__ void block_sort_desc(uint64_t *ks, uint32_t P, uint32_t tid, uint32_t T) {
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
}

// Cross-shard merge for query batches: per query, G lists of k (score, id) results (e.g. the all-gathered
// outputs of qz_u8_search_batch_dev on G ranks) -> the k best overall, same ordering contract.
struct MergeBatchParams {
    const float *scores_in;   // [G][nq][k]
    const uint32_t *ids_in;   // [G][nq][k]
    uint32_t G, nq, k, flip;
    float *scores;            // [nq][k]
    uint32_t *ids;            // [nq][k]
};

__global__ void __launch_bounds__(512) topk_merge_batch_kernel(MergeBatchParams p) {
    extern __shared__ __align__(16) uint64_t ks[];
    const uint32_t q = blockIdx.x, tid = threadIdx.x, T = blockDim.x;
    const uint32_t n = p.G * p.k;
    uint32_t P = 32;
    while (P < n) P <<= 1;
    for (uint32_t i = tid; i < P; i += T) {
        uint64_t key = 0;
        if (i < n) {
            const uint32_t g = i / p.k, j = i - g * p.k;
            const size_t at = ((size_t)g * p.nq + q) * p.k + j;
            const uint32_t id = p.ids_in[at];
            if (id != 0xFFFFFFFFu) key = make_key(p.scores_in[at], id, p.flip);  // 0xFFFFFFFF marks an empty slot
        }
        ks[i] = key;
    }
    __syncthreads();
    block_sort_desc(ks, P, tid, T);
    for (uint32_t i = tid; i < p.k; i += T) {
        float sc;
        uint32_t id;
        split_key(ks[i], p.flip, sc, id);
        p.scores[(size_t)q * p.k + i] = sc;
        p.ids[(size_t)q * p.k + i] = id;
    }
}

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
    block_sort_desc(ks, P, tid, T);
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

// k <= 128: the same compaction with the register sorting networks of topk.cuh.  4 warps per query: windows of
// 512 candidates are reduced to their 128 best (each warp sorts 128 keys in registers, two max-merge levels), the
// window heads are packed and reduced again until one window is left.  Shared memory traffic is one load and one
// store per key per level instead of one per compare-exchange stage (the generic kernel above spent 87 us per
// call on 4096 x ~350 candidates; this one ~10 us).
constexpr int kCompactWarps = 4;
constexpr uint32_t kCompactWindow = kCompactWarps * 128;
__global__ void __launch_bounds__(kCompactWarps * 32) u8_batch_compact128_kernel(CompactParams p) {
    extern __shared__ __align__(16) uint64_t ks[];  // [round_up(n, 512)] <= cap
    const uint32_t q = blockIdx.x, tid = threadIdx.x, T = kCompactWarps * 32, warp = tid >> 5, lane = tid & 31;
    uint64_t *cq = p.cand + (size_t)q * p.cap;
    uint32_t n = p.cnt[q];
    if (n > p.cap) n = p.cap;
    if (!p.final_out && n <= p.k) return;  // nothing to drop yet; threshold stays
    for (uint32_t i = tid; i < n; i += T) ks[i] = cq[i];
    __syncthreads();
    uint32_t nc = n;
    for (;;) {
        const uint32_t nw = (nc + kCompactWindow - 1) / kCompactWindow;
        for (uint32_t w = 0; w < nw; w++) {
            const uint32_t nk = nc - w * kCompactWindow < kCompactWindow ? nc - w * kCompactWindow : kCompactWindow;
            coop_top128_t<kCompactWarps>(ks + (size_t)w * kCompactWindow, nk, (nk + 127) / 128, warp, lane);
        }
        if (nw <= 1) break;  // (nw == 0: a query without a single candidate, e.g. an empty index)
        // head of window w -> ks[128w .. 128w+128): below every window not yet read (512w' >= 128w + 128 for w' >= w >= 1)
        for (uint32_t i = 128 + tid; i < nw * 128; i += T) ks[i] = ks[(size_t)(i >> 7) * kCompactWindow + (i & 127)];
        __syncthreads();
        nc = nw * 128;
    }
    // ks[0..128): the best keys, descending (zeros past the candidates)
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

// 2-D u8 tensor [rows][K] (K contiguous), box = 128 bytes x box_rows, 128-byte swizzle, OOB reads give 0
static bool make_map(CUtensorMap *map, const void *base, uint64_t rows, uint32_t K, uint32_t box_rows) {
    return make_tensor_map_u8_2d(map, base, rows, K, kBatchKC, box_rows, CU_TENSOR_MAP_SWIZZLE_128B);
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static qz_status batch_search(qz_index *ix, const float *d_queries, uint32_t nq, uint32_t k, int largest, float *d_scores,
                              uint32_t *d_ids) {
    qz_ctx *ctx = ix->ctx;
    const uint32_t K = ix->actual_dim;
    const uint64_t N = ix->count;
    // ---- SoA mirror of the corpus (once; rebuilt after qz_u8_index_set_params) ----
    if (!ix->soa_valid && N) {
        // (+2 KB: the kernel always stages whole 256-row blocks of offsets; the last 4 bytes hold max |offset|)
        const size_t vb = align_up((size_t)N * 4, 256) + 2048;
        if (!ix->d_codes_soa) {
            QZ_CUDA(ctx, cudaMalloc((void **)&ix->d_codes_soa, align_up((size_t)N * K, 256)));
            QZ_CUDA(ctx, cudaMalloc((void **)&ix->d_voff_soa, vb));
            QZ_CUDA(ctx, cudaMalloc((void **)&ix->d_vq_soa, vb));
        }
        QZ_CUDA(ctx, cudaMemsetAsync(ix->d_voff_soa, 0, vb, ctx->stream));
        QZ_CUDA(ctx, cudaMemsetAsync(ix->d_vq_soa, 0, vb, ctx->stream));
        u8_split_rows_kernel<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(ix->d_storage, N, K, ix->d_codes_soa, ix->d_voff_soa);
        QZ_CUDA(ctx, cudaGetLastError());
        uint32_t *d_amax = reinterpret_cast<uint32_t *>(reinterpret_cast<uint8_t *>(ix->d_voff_soa) + vb - 4);
        absmax_kernel<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(ix->d_voff_soa, N, d_amax);
        QZ_CUDA(ctx, cudaGetLastError());
        QZ_CUDA(ctx, cudaMemcpyAsync(&ix->voff_absmax, d_amax, 4, cudaMemcpyDeviceToHost, ctx->stream));
        QZ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        ctx->launches += 2;
        // Integer image of the vector offsets for the integer filter (see kFiltIntMax): needs a finite non-zero
        // multiplier and |offset / multiplier| < 2^30; otherwise the float filter is used for this index.
        const double M = (double)ix->multiplier;
        ix->vq_ok = std::isfinite(M) && M != 0.0 && std::isfinite((double)ix->voff_absmax) &&
                    (double)ix->voff_absmax / fabs(M) < 1073741824.0;
        if (ix->vq_ok) {
            u8_vq_kernel<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(ix->d_voff_soa, N, M, ix->d_vq_soa);
            QZ_CUDA(ctx, cudaGetLastError());
            ctx->launches++;
        }
        ix->soa_valid = true;
    }
    // ---- scratch arena ----
    const uint32_t cap = tune_u32("QZ_BATCH_CAP", k <= 256 ? 4096u : 8192u);
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

    // ---- filter mode and cluster size (2 or 1 CTAs share each corpus tile), chosen once per context = device ----
    int mode = kFiltFloat;
    if (ix->vq_ok && !tune_flag("QZ_BATCH_FLOAT_FILTER"))
        mode = ((largest != 0) == (ix->multiplier > 0.0f)) ? kFiltIntMax : kFiltIntMin;
    typedef void (*batch_fn)(const CUtensorMap, const CUtensorMap, const U8BatchParams);
    static const batch_fn fns[3][3] = {{nullptr, u8_batch_kernel<kFiltFloat, 1>, u8_batch_kernel<kFiltFloat, 2>},
                                       {nullptr, u8_batch_kernel<kFiltIntMax, 1>, u8_batch_kernel<kFiltIntMax, 2>},
                                       {nullptr, u8_batch_kernel<kFiltIntMin, 1>, u8_batch_kernel<kFiltIntMin, 2>}};
    for (int md = 0; md < 3; md++)
        for (int cl = 1; cl <= 2; cl++) QZ_TRY(ensure_smem_attr(ctx, (const void *)fns[md][cl], (int)kBatchSmem));
    QZ_TRY(ensure_smem_attr(ctx, (const void *)u8_batch_compact_kernel, 8192 * 8));
    if (cap * 8 > 48 * 1024) QZ_TRY(ensure_smem_attr(ctx, (const void *)u8_batch_compact128_kernel, (int)cap * 8));
    {
        std::lock_guard<std::mutex> ga(ctx->attr_mu);
        if (!ctx->batch_cl_pick) {
            ctx->batch_cl_grid[1] = ctx->sm_count;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(ctx->sm_count / 2 * 2);
            cfg.blockDim = dim3(kBatchThreads);
            cfg.dynamicSmemBytes = kBatchSmem;
            cudaLaunchAttribute at;
            at.id = cudaLaunchAttributeClusterDimension;
            at.val.clusterDim.x = 2;
            at.val.clusterDim.y = at.val.clusterDim.z = 1;
            cfg.attrs = &at;
            cfg.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, (const void *)fns[1][2], &cfg) != cudaSuccess) {
                cudaGetLastError();
                n = 0;
            }
            if (n > ctx->sm_count / 2) n = ctx->sm_count / 2;
            ctx->batch_cl_grid[2] = n * 2;
            // CTA pairs when they keep at least ~95% of the SMs busy
            ctx->batch_cl_pick = ctx->batch_cl_grid[2] * 100 >= ctx->sm_count * 95 ? 2 : 1;
            const uint32_t v = tune_u32("QZ_BATCH_CLUSTER", 0);
            if ((v == 1 || v == 2) && ctx->batch_cl_grid[v] > 0) ctx->batch_cl_pick = (int)v;
        }
    }
    const int CL = ctx->batch_cl_pick;
    const int *cl_grid = ctx->batch_cl_grid;
    CUtensorMap map_q, map_v;
    if (!make_map(&map_q, d_qcodes, nq, K, kBatchM) || (N && !make_map(&map_v, ix->d_codes_soa, N, K, kBatchN / CL)))
        return fail(ctx, QZ_ENCODING, "cuTensorMapEncodeTiled failed");
    if (!N) memset(&map_v, 0, sizeof map_v);

    U8BatchParams p{};
    p.nq = nq;
    p.k_chunks = (K + kBatchKC - 1) / kBatchKC;
    p.n_qb = ((nq + kBatchM - 1) / kBatchM + CL - 1) / CL * CL;  // padded: whole groups of CL query blocks
    p.qoff = d_qoff;
    p.voff = ix->d_voff_soa;
    p.vq = ix->d_vq_soa;
    p.voff_absmax = ix->voff_absmax;
    p.multiplier = ix->multiplier;
    p.rmult = mode == kFiltFloat ? 0.0f : 1.0f / ix->multiplier;
#ifdef QZ_TUNING
    p.dbg = tune_u32("QZ_BATCH_DBG", 0);
#endif
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
        // (fast schedule: the first chunk is only cap/4 rows: all of them become candidates, and the first
        // compaction sorts that many keys per query)
        const uint64_t first = safe ? ((cap - k) / kBatchN * kBatchN ? (cap - k) / kBatchN * kBatchN : kBatchN)
                                    : cap / tune_u32("QZ_BATCH_FIRST_DIV", 4);
        do {
            uint64_t R = first;
            if (!safe && done) {
                const uint64_t margin = tune_u32("QZ_BATCH_MARGIN", 16);
                const uint64_t grow = done * cap / (margin * k);
                R = grow > first ? grow / kBatchN * kBatchN : first;
            }
            if (safe && done) R = (cap - k) / kBatchN * kBatchN ? (cap - k) / kBatchN * kBatchN : kBatchN;
            if (done + R > N) R = N - done;
            const bool last = done + R >= N;
            if (R) {
                p.row_begin = (uint32_t)done;
                p.row_end = (uint32_t)(done + R);
                p.n_ct = (uint32_t)((R + kBatchN - 1) / kBatchN);
                const uint32_t items = p.n_qb / CL * p.n_ct;
                const uint32_t max_clusters = (uint32_t)cl_grid[CL] / CL;
                const uint32_t grid = (items < max_clusters ? items : max_clusters) * CL;
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(grid);
                cfg.blockDim = dim3(kBatchThreads);
                cfg.dynamicSmemBytes = kBatchSmem;
                cfg.stream = ctx->stream;
                cudaLaunchAttribute at;
                at.id = cudaLaunchAttributeClusterDimension;
                at.val.clusterDim.x = CL;
                at.val.clusterDim.y = at.val.clusterDim.z = 1;
                cfg.attrs = &at;
                cfg.numAttrs = 1;
                timing_begin(ctx);
                const cudaError_t le = cudaLaunchKernelEx(&cfg, fns[mode][CL], map_q, map_v, p);
                timing_end(ctx);
                QZ_CUDA(ctx, le);
                ctx->launches++;
            }
            cp.final_out = last;
            if (k <= 128 && !tune_flag("QZ_BATCH_COMPACT_SORT"))
                u8_batch_compact128_kernel<<<nq, kCompactWarps * 32, (size_t)cap * 8, ctx->stream>>>(cp);
            else
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

qz_status qz_merge_topk_batch_dev(qz_ctx *ctx, const float *d_scores_in, const uint32_t *d_ids_in, uint32_t n_lists,
                                  uint32_t nq, uint32_t k, int32_t largest, float *d_scores, uint32_t *d_ids) {
    if (!ctx) return QZ_ARGUMENTS;
    if (k == 0 || k > QZ_MAX_K) return fail(ctx, QZ_ARGUMENTS, "merge_topk_batch: k must be in [1, QZ_MAX_K]");
    if (nq == 0 || n_lists == 0) return QZ_OK;
    if (!d_scores_in || !d_ids_in || !d_scores || !d_ids) return fail(ctx, QZ_ARGUMENTS, "merge_topk_batch: NULL argument");
    if ((uint64_t)n_lists * k > 16384) return fail(ctx, QZ_ARGUMENTS, "merge_topk_batch: n_lists * k must be <= 16384");
    QZ_CUDA(ctx, cudaSetDevice(ctx->device));
    QZ_TRY(ensure_smem_attr(ctx, (const void *)topk_merge_batch_kernel, 16384 * 8));
    uint32_t P = 32;
    while (P < n_lists * k) P <<= 1;
    MergeBatchParams mp{d_scores_in, d_ids_in, n_lists, nq, k, largest ? 0u : 0xFFFFFFFFu, d_scores, d_ids};
    topk_merge_batch_kernel<<<nq, 512, (size_t)P * 8, ctx->stream>>>(mp);
    QZ_CUDA(ctx, cudaGetLastError());
    ctx->launches++;
    return QZ_OK;
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
