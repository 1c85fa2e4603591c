// topk.cuh — exact, deterministic top-k over u64 keys (see common.cuh make_key).
//
//  * Inside a scan CTA: every consumer warp filters scores against a CTA-wide threshold and
//    stages survivors in a warp-private region; a full region is merged (under a CTA lock)
//    into the CTA's sorted top-Kp list with bitonic networks, which raises the threshold.
//    Any interleaving gives the same final list because keys are unique and the filter is
//    conservative (only keys <= the current k-th best are ever dropped).
//  * Across CTAs / shards: merge_lists_kernel reduces sorted lists pairwise
//    (elementwise max of A[i], B[Kp-1-i] = bitonic sequence holding the top Kp, then a
//    bitonic merge), log2(#lists) levels, in shared memory.
//
// Replaces the caller-side BinaryHeap of demos/src/ann_benchmark_data.rs:150-167.
#pragma once
#include "common.cuh"

namespace qz {

constexpr int kCandPerWarp = 128;  // warp-private staging region (u64 keys)

__device__ __forceinline__ void ce_desc(uint64_t *a, int lo, int hi) {
    uint64_t x = a[lo], y = a[hi];
    if (x < y) {
        a[lo] = y;
        a[hi] = x;
    }
}
__device__ __forceinline__ void ce_asc(uint64_t *a, int lo, int hi) {
    uint64_t x = a[lo], y = a[hi];
    if (x > y) {
        a[lo] = y;
        a[hi] = x;
    }
}

// one warp sorts n (power of two) keys in shared memory, descending
__device__ __forceinline__ void warp_sort_desc(uint64_t *a, int n, int lane) {
    for (int size = 2; size <= n; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = lane; t < (n >> 1); t += 32) {
                int lo = 2 * t - (t & (stride - 1));
                int hi = lo + stride;
                if ((lo & size) == 0 || size == n) ce_desc(a, lo, hi);
                else ce_asc(a, lo, hi);
            }
            __syncwarp();
        }
    }
}
// one warp turns a bitonic sequence of n keys into a sorted one
template <bool DESC>
__device__ __forceinline__ void warp_bitonic_merge(uint64_t *a, int n, int lane) {
    for (int stride = n >> 1; stride > 0; stride >>= 1) {
        for (int t = lane; t < (n >> 1); t += 32) {
            int lo = 2 * t - (t & (stride - 1));
            int hi = lo + stride;
            if (DESC) ce_desc(a, lo, hi);
            else ce_asc(a, lo, hi);
        }
        __syncwarp();
    }
}

// CTA-wide top-k state living in shared memory
struct CtaTopK {
    uint64_t *top;            // [Kp] sorted descending, 0 = empty
    volatile uint64_t *thr;   // current k-th best key of this CTA (monotone non-decreasing)
    int *lock;
    uint32_t k, Kp;
};

// Merge one warp's staged candidates cand[0..cnt) into the CTA list.  Called by a full warp.
__device__ __forceinline__ void cta_topk_flush(const CtaTopK &st, uint64_t *cand, int cnt, int lane) {
    for (int i = cnt + lane; i < kCandPerWarp; i += 32) cand[i] = 0;
    __syncwarp();
    warp_sort_desc(cand, kCandPerWarp, lane);
    if (lane == 0) {
        while (atomicCAS(st.lock, 0, 1) != 0) {
        }
    }
    __syncwarp();
    __threadfence_block();
    uint64_t *tail = st.top + (st.Kp - kCandPerWarp);
    for (int i = lane; i < kCandPerWarp; i += 32) {
        uint64_t a = tail[i], b = cand[kCandPerWarp - 1 - i];
        tail[i] = a > b ? a : b;
    }
    __syncwarp();
    if (st.Kp == (uint32_t)kCandPerWarp) {
        warp_bitonic_merge<true>(st.top, kCandPerWarp, lane);
    } else {
        warp_bitonic_merge<false>(tail, kCandPerWarp, lane);  // tail ascending => [head desc | tail asc] is bitonic
        warp_bitonic_merge<true>(st.top, (int)st.Kp, lane);
    }
    if (lane == 0) *st.thr = st.top[st.k - 1];
    __syncwarp();
    __threadfence_block();
    if (lane == 0) atomicExch(st.lock, 0);
}

// ---------------------------------------------------------------------------------------------
// Merge of sorted lists.  Each CTA reduces `lists_per_cta` (power of two) consecutive input
// lists to one list of Kp keys.  Input list l occupies in[l*in_stride .. +in_len), best first,
// padded here with 0.  Shared memory: (lists_per_cta/2) * Kp keys.
// If `final_out` the single result is also decoded into scores / ids (first k entries).
// ---------------------------------------------------------------------------------------------
struct MergeParams {
    const uint64_t *in;
    uint32_t n_lists, in_len, in_stride;
    uint32_t Kp, lists_per_cta;
    uint64_t *out;  // [grid][out_stride] ; may be NULL when final
    uint32_t out_stride, out_len;
    int final_out;
    uint32_t k, flip;
    uint64_t *keys;  // final outputs (each may be NULL)
    float *scores;
    uint32_t *ids;
};

template <int kUnused = 0>
__global__ void __launch_bounds__(1024, 1) merge_lists_kernel(MergeParams p) {
    extern __shared__ __align__(16) uint64_t s[];
    const uint32_t tid = threadIdx.x, T = blockDim.x;
    const uint32_t Kp = p.Kp, half = Kp >> 1;
    const uint32_t first = blockIdx.x * p.lists_per_cta;
    uint32_t nl = p.lists_per_cta >> 1;  // lists resident in smem after the load level
    // level 0: fused load + pairwise max
    for (uint32_t idx = tid; idx < nl * Kp; idx += T) {
        uint32_t pr = idx / Kp, i = idx - pr * Kp;
        uint32_t la = first + 2 * pr, lb = la + 1, j = Kp - 1 - i;
        uint64_t a = (la < p.n_lists && i < p.in_len) ? p.in[(size_t)la * p.in_stride + i] : 0ull;
        uint64_t b = (lb < p.n_lists && j < p.in_len) ? p.in[(size_t)lb * p.in_stride + j] : 0ull;
        s[idx] = a > b ? a : b;
    }
    __syncthreads();
    uint32_t step = 1;  // distance (in lists) between live lists
    for (;;) {
        // bitonic merge of every live list (descending)
        for (uint32_t stride = half; stride > 0; stride >>= 1) {
            for (uint32_t t = tid; t < nl * half; t += T) {
                uint32_t sub = t / half, tt = t - sub * half;
                uint32_t lo = sub * step * Kp + 2 * tt - (tt & (stride - 1));
                ce_desc(s, (int)lo, (int)(lo + stride));
            }
            __syncthreads();
        }
        if (nl == 1) break;
        // pairwise max of live lists (in place: result lands in the first list of each pair)
        nl >>= 1;
        for (uint32_t idx = tid; idx < nl * Kp; idx += T) {
            uint32_t pr = idx / Kp, i = idx - pr * Kp;
            uint64_t *A = s + (size_t)(2 * pr) * step * Kp, *B = A + (size_t)step * Kp;
            uint64_t a = A[i], b = B[Kp - 1 - i];
            if (b > a) A[i] = b;
        }
        step <<= 1;
        __syncthreads();
    }
    if (p.out) {
        for (uint32_t i = tid; i < p.out_len; i += T) p.out[(size_t)blockIdx.x * p.out_stride + i] = i < Kp ? s[i] : 0ull;
    }
    if (p.final_out) {
        for (uint32_t i = tid; i < p.k; i += T) {
            uint64_t key = i < Kp ? s[i] : 0ull;
            if (p.keys) p.keys[i] = key;
            float sc;
            uint32_t id;
            split_key(key, p.flip, sc, id);
            if (p.scores) p.scores[i] = sc;
            if (p.ids) p.ids[i] = id;
        }
    }
}

}  // namespace qz
