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

// ---------------------------------------------------------------------------------------------
// Register-resident bitonic networks over 128 keys held 4 per lane ("blocked": element e lives in
// lane e / 4, register e % 4).  Strides >= 4 exchange whole registers between lanes with
// shuffles, strides 1 and 2 stay inside the lane.  No shared memory, no __syncwarp.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t shfl_xor_u64(uint64_t v, int mask) {
    const uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)v, mask);
    const uint32_t hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(v >> 32), mask);
    return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ uint64_t shfl_idx_u64(uint64_t v, int src) {
    const uint32_t lo = __shfl_sync(0xffffffffu, (uint32_t)v, src);
    const uint32_t hi = __shfl_sync(0xffffffffu, (uint32_t)(v >> 32), src);
    return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ void ce_reg(uint64_t &a, uint64_t &b, bool desc) {  // a is the lower index
    const bool swap = desc ? (a < b) : (a > b);
    const uint64_t x = swap ? b : a, y = swap ? a : b;
    a = x;
    b = y;
}
// one (size, stride) step of the network; `desc_all` forces one direction (a merge step)
template <bool MERGE_DESC>
__device__ __forceinline__ void reg128_step(uint64_t (&v)[4], int size, int stride, int lane) {
    if (stride >= 4) {
        const int lm = stride >> 2;
        const bool lower = (lane & lm) == 0;
        const bool desc = MERGE_DESC || (((lane << 2) & size) == 0);
        const bool keep_max = lower == desc;
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint64_t o = shfl_xor_u64(v[r], lm);
            v[r] = keep_max ? (v[r] > o ? v[r] : o) : (v[r] < o ? v[r] : o);
        }
    } else if (stride == 2) {
        const bool desc = MERGE_DESC || (((lane << 2) & size) == 0);
        ce_reg(v[0], v[2], desc);
        ce_reg(v[1], v[3], desc);
    } else {
        // stride 1: elements (0,1) and (2,3); for size 2 the two pairs sort in opposite directions
        const bool d0 = MERGE_DESC || ((((lane << 2) | 0) & size) == 0);
        const bool d1 = MERGE_DESC || ((((lane << 2) | 2) & size) == 0);
        ce_reg(v[0], v[1], d0);
        ce_reg(v[2], v[3], d1);
    }
}
// full sort, descending
__device__ __forceinline__ void reg128_sort_desc(uint64_t (&v)[4], int lane) {
#pragma unroll
    for (int size = 2; size <= 128; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            if (size == 128) reg128_step<true>(v, size, stride, lane);
            else reg128_step<false>(v, size, stride, lane);
        }
    }
}
// bitonic sequence -> sorted descending
__device__ __forceinline__ void reg128_merge_desc(uint64_t (&v)[4], int lane) {
#pragma unroll
    for (int stride = 64; stride > 0; stride >>= 1) reg128_step<true>(v, 128, stride, lane);
}

// Cooperative reduction of n_chunks (<= 16) chunks of 128 keys in shared memory to the 128 largest,
// sorted descending, in chunk 0.  Called by all kCoopWarps consumer warps (named barrier 1);
// keys[n_keys .. n_chunks*128) is zero-filled here.  Chunk w is sorted by warp w in registers, then
// a pairwise max-merge tree (log2(n_chunks) levels) runs with one warp per surviving pair.
constexpr int kCoopWarps = 16;
// NW = warps taking part (all of them call; they must be warps 0..NW-1 of the CTA and the only users of named
// barrier 1); n_chunks <= NW.
template <int NW>
__device__ __forceinline__ void coop_top128_t(uint64_t *keys, uint32_t n_keys, uint32_t n_chunks, uint32_t warp,
                                              uint32_t lane) {
    for (uint32_t i = n_keys + warp * 32 + lane; i < n_chunks * 128; i += NW * 32) keys[i] = 0ull;
    asm volatile("bar.sync 1, %0;" ::"r"(NW * 32) : "memory");
    if (warp < n_chunks) {
        uint64_t c[4];
        uint64_t *ch = keys + warp * 128;
#pragma unroll
        for (int r = 0; r < 4; r++) c[r] = ch[lane * 4 + r];
        reg128_sort_desc(c, (int)lane);
#pragma unroll
        for (int r = 0; r < 4; r++) ch[lane * 4 + r] = c[r];
    }
    for (uint32_t stride = 1; stride < n_chunks; stride <<= 1) {
        asm volatile("bar.sync 1, %0;" ::"r"(NW * 32) : "memory");
        if ((warp & (2 * stride - 1)) == 0 && warp + stride < n_chunks) {
            uint64_t *A = keys + warp * 128, *B = keys + (warp + stride) * 128;
            uint64_t t[4];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const uint64_t a = A[lane * 4 + r], b = B[127 - (lane * 4 + r)];
                t[r] = a > b ? a : b;
            }
            reg128_merge_desc(t, (int)lane);
#pragma unroll
            for (int r = 0; r < 4; r++) A[lane * 4 + r] = t[r];
        }
    }
    asm volatile("bar.sync 1, %0;" ::"r"(NW * 32) : "memory");
}
__device__ __forceinline__ void coop_top128(uint64_t *keys, uint32_t n_keys, uint32_t n_chunks, uint32_t warp,
                                            uint32_t lane) {
    coop_top128_t<kCoopWarps>(keys, n_keys, n_chunks, warp, lane);
}

// Merge one warp's staged candidates cand[0..cnt) into the CTA list.  Called by a full warp.
__device__ __forceinline__ void cta_topk_flush(const CtaTopK &st, uint64_t *cand, int cnt, int lane) {
    if (st.Kp == (uint32_t)kCandPerWarp) {
        // fast path (k <= 128): everything in registers; the lock covers ~12 network steps
        uint64_t c[4];
#pragma unroll
        for (int r = 0; r < 4; r++) c[r] = (lane * 4 + r) < cnt ? cand[lane * 4 + r] : 0ull;
        reg128_sort_desc(c, lane);
        // reversed copy: element e <- element 127 - e  (lane 31 - lane, register 3 - r)
        uint64_t rv[4];
#pragma unroll
        for (int r = 0; r < 4; r++) rv[r] = shfl_idx_u64(c[3 - r], 31 - lane);
        if (lane == 0) {
            while (atomicCAS(st.lock, 0, 1) != 0) {
            }
        }
        __syncwarp();
        __threadfence_block();
        uint64_t t[4];
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint64_t a = st.top[lane * 4 + r];
            t[r] = a > rv[r] ? a : rv[r];  // top-128 of the union, as a bitonic sequence
        }
        reg128_merge_desc(t, lane);
#pragma unroll
        for (int r = 0; r < 4; r++) st.top[lane * 4 + r] = t[r];
        const uint32_t ke = st.k - 1;
        const uint32_t kr = ke & 3;
        const uint64_t kth = kr == 0 ? t[0] : (kr == 1 ? t[1] : (kr == 2 ? t[2] : t[3]));
        if ((uint32_t)lane == (ke >> 2)) *st.thr = kth;
        __syncwarp();
        __threadfence_block();
        if (lane == 0) atomicExch(st.lock, 0);
        __syncwarp();
        return;
    }
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
    warp_bitonic_merge<false>(tail, kCandPerWarp, lane);  // tail ascending => [head desc | tail asc] is bitonic
    warp_bitonic_merge<true>(st.top, (int)st.Kp, lane);
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

// The merge bodies below run either as a whole CTA (BAR == 0: __syncthreads) or as a thread group of
// T threads with ranks tid = 0..T-1 inside a bigger CTA (BAR > 0: named barrier BAR), which is how the
// single-query scan runs its final merge in its last CTA (row_scan.cuh).  CG: read the input lists with
// ld.global.cg (they were written by other CTAs of the same launch).
template <int BAR>
__device__ __forceinline__ void grp_sync(uint32_t T) {
    if (BAR == 0) __syncthreads();
    else named_bar_sync(BAR, T);
}
template <bool CG>
__device__ __forceinline__ uint64_t ld_list(const uint64_t *ptr) {
    return CG ? (uint64_t)__ldcg(reinterpret_cast<const unsigned long long *>(ptr)) : *ptr;
}

// first k keys of s[0..have) -> keys / scores / ids outputs (slots past `have` are empty)
__device__ __forceinline__ void write_final(const MergeParams &p, const uint64_t *s, uint32_t have, uint32_t tid,
                                            uint32_t T) {
    for (uint32_t i = tid; i < p.k; i += T) {
        const uint64_t key = i < have ? s[i] : 0ull;
        if (p.keys) p.keys[i] = key;
        float sc;
        uint32_t id;
        split_key(key, p.flip, sc, id);
        if (p.scores) p.scores[i] = sc;
        if (p.ids) p.ids[i] = id;
    }
}

template <int BAR, bool CG>
__device__ __forceinline__ void tree_merge(const MergeParams &p, uint64_t *s, uint32_t tid, uint32_t T, uint32_t blk) {
    const uint32_t Kp = p.Kp, half = Kp >> 1;
    const uint32_t first = blk * p.lists_per_cta;
    uint32_t nl = p.lists_per_cta >> 1;  // lists resident in smem after the load level
    // level 0: fused load + pairwise max
    for (uint32_t idx = tid; idx < nl * Kp; idx += T) {
        uint32_t pr = idx / Kp, i = idx - pr * Kp;
        uint32_t la = first + 2 * pr, lb = la + 1, j = Kp - 1 - i;
        uint64_t a = (la < p.n_lists && i < p.in_len) ? ld_list<CG>(p.in + (size_t)la * p.in_stride + i) : 0ull;
        uint64_t b = (lb < p.n_lists && j < p.in_len) ? ld_list<CG>(p.in + (size_t)lb * p.in_stride + j) : 0ull;
        s[idx] = a > b ? a : b;
    }
    grp_sync<BAR>(T);
    uint32_t step = 1;  // distance (in lists) between live lists
    for (;;) {
        // bitonic merge of every live list (descending)
        for (uint32_t stride = half; stride > 0; stride >>= 1) {
            for (uint32_t t = tid; t < nl * half; t += T) {
                uint32_t sub = t / half, tt = t - sub * half;
                uint32_t lo = sub * step * Kp + 2 * tt - (tt & (stride - 1));
                ce_desc(s, (int)lo, (int)(lo + stride));
            }
            grp_sync<BAR>(T);
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
        grp_sync<BAR>(T);
    }
    if (p.out) {
        for (uint32_t i = tid; i < p.out_len; i += T) p.out[(size_t)blk * p.out_stride + i] = i < Kp ? s[i] : 0ull;
    }
    if (p.final_out) write_final(p, s, Kp, tid, T);
}

template <int kUnused = 0>
__global__ void __launch_bounds__(1024, 1) merge_lists_kernel(MergeParams p) {
    extern __shared__ __align__(16) uint64_t s[];
    tree_merge<0, false>(p, s, threadIdx.x, blockDim.x, blockIdx.x);
}

// ---------------------------------------------------------------------------------------------
// Final merge with pruning (one thread group).  The lists are sorted, so for any position j the m-th
// largest of {list_g[j]} has at least m*(j+1) elements at or above it: with m*(j+1) >= k that is
// a lower bound T on the k-th best key overall.  Only keys >= T (typically a few hundred out of
// n_lists*k) are gathered and sorted.  If more than kPruneCap survive (adversarial inputs) the exact
// tree merge above runs instead, in the same launch.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kPruneCap = 2048;
constexpr uint32_t kPruneMaxLists = 256;

struct PruneShared {
    uint64_t vals[kPruneMaxLists];
    uint64_t wmax[kPruneMaxLists / 32];
    uint64_t thrA;
    uint32_t ranks[kPruneMaxLists];
    uint32_t cnt, pad;
};
// shared memory the pruned path needs: sort area + staged survivors + their ranks + the block above
constexpr size_t kPruneSmem = (size_t)kPruneCap * (8 + 8 + 4) + sizeof(PruneShared);

// s: max(kPruneSmem, tree-merge need) bytes, 16-byte aligned.  Needs 256 <= T, T a multiple of 256... (T = 512, 1024)
template <int BAR, bool CG>
__device__ __forceinline__ void prune_merge(const MergeParams &p, uint64_t *s, uint32_t tid, uint32_t T) {
    uint64_t *cand = s + kPruneCap;
    uint32_t *rk = reinterpret_cast<uint32_t *>(cand + kPruneCap);
    PruneShared &sh = *reinterpret_cast<PruneShared *>(rk + kPruneCap);
    const uint32_t lane = tid & 31, warp = tid >> 5;
    const uint32_t G = p.n_lists;
    // probe A: position jp1-1 and the m-th largest value there, m*jp1 >= k, m <= G
    // probe B: position k-1, m = 1 (the best list's k-th key)
    const uint32_t m0 = G < p.k ? G : p.k;
    const uint32_t jp1 = m0 ? (p.k + m0 - 1) / m0 : 1;
    const uint32_t m = (p.k + jp1 - 1) / jp1;
    const bool probeA = G > 0 && jp1 <= p.in_len && m <= G;
    const bool probeB = G > 0 && p.k <= p.in_len;
    if (tid == 0) {
        sh.thrA = 0ull;
        sh.cnt = 0;
    }
    if (tid < kPruneMaxLists) {
        uint64_t a = 0, b = 0;
        if (tid < G) {
            if (probeA) a = ld_list<CG>(p.in + (size_t)tid * p.in_stride + (jp1 - 1));
            if (probeB) b = ld_list<CG>(p.in + (size_t)tid * p.in_stride + (p.k - 1));
        }
        sh.vals[tid] = a;
        sh.ranks[tid] = 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const uint64_t b2 = __shfl_xor_sync(0xffffffffu, b, o);
            b = b2 > b ? b2 : b;
        }
        if (lane == 0) sh.wmax[warp] = b;
    }
    // first 32 positions of this warp's first 8 lists: in flight while the threshold is computed
    const uint32_t nw = T >> 5;
    uint64_t kk[8];
#pragma unroll
    for (int u = 0; u < 8; u++) {
        const uint32_t l = warp + u * nw;
        kk[u] = (l < G && lane < p.in_len) ? ld_list<CG>(p.in + (size_t)l * p.in_stride + lane) : 0ull;
    }
    grp_sync<BAR>(T);
    uint64_t thrB = 0;
#pragma unroll
    for (uint32_t w = 0; w < kPruneMaxLists / 32; w++) thrB = sh.wmax[w] > thrB ? sh.wmax[w] : thrB;
    if (probeA) {  // rank every probe value; the comparison loop is split over T / Gp threads per value
        uint32_t Gp = 32;
        while (Gp < G) Gp <<= 1;
        const uint32_t parts = T / Gp, g = tid % Gp, part = tid / Gp;
        if (g < G) {
            const uint64_t mine = sh.vals[g];
            const uint32_t per = (G + parts - 1) / parts;
            const uint32_t h0 = part * per, h1 = min(h0 + per, G);
            uint32_t r = 0;
            for (uint32_t h = h0; h < h1; h++) {
                const uint64_t v = sh.vals[h];
                r += (v > mine) || (v == mine && h < g);
            }
            if (r) atomicAdd(&sh.ranks[g], r);
        }
        grp_sync<BAR>(T);
        if (tid < G && sh.ranks[tid] == m - 1) sh.thrA = sh.vals[tid];
    }
    grp_sync<BAR>(T);
    const uint64_t Tkey = sh.thrA > thrB ? sh.thrA : thrB;
    // the survivors of a sorted list are a prefix: one warp per list, 32 positions per step
    for (uint32_t l0 = warp; l0 < G; l0 += nw * 8) {
        if (l0 != warp) {
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const uint32_t l = l0 + u * nw;
                kk[u] = (l < G && lane < p.in_len) ? ld_list<CG>(p.in + (size_t)l * p.in_stride + lane) : 0ull;
            }
        }
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const uint32_t l = l0 + u * nw;
            if (l >= G) break;
            const uint64_t *lp = p.in + (size_t)l * p.in_stride;
            uint64_t key = kk[u];
            for (uint32_t i0 = 0;;) {
                const bool pass = key != 0 && key >= Tkey;
                const uint32_t mk = __ballot_sync(0xffffffffu, pass);
                if (mk) {
                    uint32_t pos0 = 0;
                    if (lane == 0) pos0 = atomicAdd(&sh.cnt, (uint32_t)__popc(mk));
                    pos0 = __shfl_sync(0xffffffffu, pos0, 0);
                    const uint32_t pos = pos0 + __popc(mk & ((1u << lane) - 1));
                    if (pass && pos < kPruneCap) cand[pos] = key;
                }
                i0 += 32;
                if (mk != 0xffffffffu || i0 >= p.in_len) break;
                key = i0 + lane < p.in_len ? ld_list<CG>(lp + i0 + lane) : 0ull;
            }
        }
    }
    grp_sync<BAR>(T);
    const uint32_t n = sh.cnt;
    if (n > kPruneCap) {  // not prunable: the exact tree merge, in the same launch (uniform branch)
        grp_sync<BAR>(T);  // everyone has read n before the tree merge overwrites the area
        tree_merge<BAR, CG>(p, s, tid, T, 0);
        return;
    }
    if (n <= 512) {
        // rank sort: survivor e goes to slot #{survivors greater than it} (keys are unique)
        for (uint32_t i = tid; i < kPruneCap; i += T) {
            if (i < p.k) s[i] = 0ull;
            if (i < 512) rk[i] = 0;
        }
        grp_sync<BAR>(T);
        uint32_t np = 32;
        while (np < n) np <<= 1;
        const uint32_t parts = T / np, e = tid % np, part = tid / np;
        if (e < n) {
            const uint64_t mine = cand[e];
            const uint32_t per = (n + parts - 1) / parts;
            const uint32_t h0 = part * per, h1 = min(h0 + per, n);
            uint32_t r = 0;
            for (uint32_t h = h0; h < h1; h++) r += cand[h] > mine;
            if (r) atomicAdd(&rk[e], r);
        }
        grp_sync<BAR>(T);
        if (tid < n && rk[tid] < p.k) s[rk[tid]] = cand[tid];
        grp_sync<BAR>(T);
        write_final(p, s, p.k, tid, T);
        return;
    }
    uint32_t P = 1024;
    while (P < n) P <<= 1;
    for (uint32_t i = tid; i < P; i += T) s[i] = i < n ? cand[i] : 0ull;
    grp_sync<BAR>(T);
    for (uint32_t size = 2; size <= P; size <<= 1) {
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            for (uint32_t t = tid; t < (P >> 1); t += T) {
                const uint32_t lo = 2 * t - (t & (stride - 1));
                if ((lo & size) == 0 || size == P) ce_desc(s, (int)lo, (int)(lo + stride));
                else ce_asc(s, (int)lo, (int)(lo + stride));
            }
            grp_sync<BAR>(T);
        }
    }
    write_final(p, s, P, tid, T);
}

// Final merge of a FEW sorted lists (the world's lists after the cross-GPU exchange: G <= 16, at most 128 keys each):
// list g becomes chunk g of the work area, then a pairwise max-merge tree in registers, one warp per pair, log2(G)
// levels.  About a third of prune_merge's latency at G = 8 (no probes, no ranking, no sort).  T = 512 threads.
template <int BAR, bool CG>
__device__ __forceinline__ void few_lists_merge(const MergeParams &p, uint64_t *s, uint32_t tid, uint32_t T) {
    const uint32_t G = p.n_lists, lane = tid & 31, warp = tid >> 5;
    uint32_t Gp = 1;
    while (Gp < G) Gp <<= 1;
    for (uint32_t i = tid; i < Gp * 128; i += T) {
        const uint32_t g = i >> 7, j = i & 127;
        s[i] = (g < G && j < p.in_len) ? ld_list<CG>(p.in + (size_t)g * p.in_stride + j) : 0ull;
    }
    grp_sync<BAR>(T);
    for (uint32_t stride = 1; stride < G; stride <<= 1) {
        if ((warp & (2 * stride - 1)) == 0 && warp + stride < G) {
            uint64_t *A = s + warp * 128, *B = s + (warp + stride) * 128;
            uint64_t t[4];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const uint64_t a = A[lane * 4 + r], b = B[127 - (lane * 4 + r)];
                t[r] = a > b ? a : b;
            }
            reg128_merge_desc(t, (int)lane);
#pragma unroll
            for (int r = 0; r < 4; r++) A[lane * 4 + r] = t[r];
        }
        grp_sync<BAR>(T);
    }
    write_final(p, s, 128, tid, T);
}

template <int kUnused = 0>
__global__ void __launch_bounds__(1024, 1) prune_merge_kernel(MergeParams p) {
    extern __shared__ __align__(16) uint64_t s[];  // max(kPruneSmem, tree-merge need)
    prune_merge<0, false>(p, s, threadIdx.x, blockDim.x);
}

}  // namespace qz
