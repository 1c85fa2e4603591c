// peer_exchange.cuh — the cross-GPU top-k exchange executed by the last thread block of a scan launch
// (row_scan.cuh, pq_scan2.cu); host side in peer.cu.
#pragma once
#include "common.cuh"
#include "topk.cuh"

namespace qz {

// Cross-GPU exchange fused into the scan launch (one process per GPU, shards = point-id ranges).  Every GPU owns a
// symmetric buffer mapped into its peers (CUDA IPC over NVLink): gather[2][world][128] u64 keys and flags[2][world]
// u32.  After the in-launch merge the last CTA stores its k best keys into slot [buf][rank] of EVERY GPU's buffer,
// fences system-wide, raises flag [buf][rank] = seq on every GPU, waits until its own flags [buf][*] all read seq, and
// merges the world's lists: all ranks end the launch with the same global top-k, with no host-side collective.
// buf = seq & 1: a rank can be at most one step ahead of a peer (it needs the peer's flag of the step), so two
// buffers are enough.
constexpr int kMaxPeers = 8;
constexpr uint32_t kPeerStride = 128;  // keys per gather slot
struct PeerExchange {
    uint32_t world, rank, seq, buf;
    int32_t root;        // < 0: all-to-all (every rank ends with the global top-k); else the gathering rank
    // Flow control.  A rank that READS its buffer in a step (every rank of an all-to-all step, the root of a rooted
    // one) acknowledges the step to every peer once its merge is done: ack[src] in the peer's memory.  Before a rank
    // writes into peer g's buffer half `buf` it waits until g has acknowledged the last step in which g read that
    // half (need_ack[g], kept by the host: every rank makes the same sequence of calls).  In a pure all-to-all
    // sequence the wait is always already satisfied (nobody can be two steps ahead); it matters when rooted steps,
    // in which the non-root ranks wait for nobody, are in flight.
    uint32_t need_ack[kMaxPeers];
    uint64_t *gather[kMaxPeers];  // gather[g] = base of GPU g's buffer (own = local pointer)
    uint32_t *flags[kMaxPeers];   // flags[2][world] u32, then ack[world] u32 from index kPeerAckIndex
    MergeParams merge;  // world lists of k keys from the local buffer -> final outputs
};
constexpr uint32_t kPeerAckIndex = 32;

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// wait until peer `g` has acknowledged step `need` (0 = nothing to wait for); called by one thread per target
__device__ __forceinline__ void peer_wait_ack(const PeerExchange &x, uint32_t g, uint32_t need) {
    if (!need) return;
    const uint32_t *a = x.flags[x.rank] + kPeerAckIndex + g;
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(a) - need) < 0) {
        if (++spins > (1u << 26)) __trap();
    }
}
// tell every peer that this rank has finished reading its buffer for step x.seq
__device__ __forceinline__ void peer_send_acks(const PeerExchange &x) {
    if (threadIdx.x < x.world && threadIdx.x != x.rank) st_release_sys(x.flags[threadIdx.x] + kPeerAckIndex + x.rank, x.seq);
}


// Called by the 512 consumer threads (ranks threadIdx.x = 0..511, named barrier 1) of the launch's last CTA after
// the local final merge has written this GPU's k best keys to its own gather slot.  `work` = shared memory for the
// merge (>= kPruneSmem bytes).
__device__ __forceinline__ void peer_exchange_merge(const PeerExchange &x, uint32_t k, uint64_t *work) {
    constexpr uint32_t T = 512;
    const size_t slot = ((size_t)x.buf * x.world + x.rank) * kPeerStride;
    __threadfence();
    named_bar_sync(1, T);
    if (x.root >= 0) {
        // ---- rooted gather (north_star: "merged with a gather"): only the root waits and merges ----
        if (x.rank != (uint32_t)x.root) {
            // A non-root rank never waits for the slowest rank (which the all-to-all form pays every step); it may
            // run ahead of the root by as many steps as the root's acknowledgements allow (two buffer halves).
            if (threadIdx.x == 0) peer_wait_ack(x, (uint32_t)x.root, x.need_ack[x.root]);
            named_bar_sync(1, T);
            for (uint32_t j = threadIdx.x; j < k; j += T)
                x.gather[x.root][slot + j] = __ldcg(reinterpret_cast<const unsigned long long *>(x.gather[x.rank] + slot + j));
            __threadfence_system();
            named_bar_sync(1, T);
            if (threadIdx.x == 0) st_release_sys(x.flags[x.root] + x.buf * x.world + x.rank, x.seq);
            return;
        }
        if (threadIdx.x < x.world && threadIdx.x != x.rank) {
            const uint32_t *f = x.flags[x.rank] + x.buf * x.world + threadIdx.x;
            uint32_t spins = 0;
            while (ld_acquire_sys(f) != x.seq) {
                if (++spins > (1u << 26)) __trap();
            }
        }
        named_bar_sync(1, T);
        few_lists_merge<1, true>(x.merge, work, threadIdx.x, T);  // world <= kMaxPeers lists of k <= 128 keys
        named_bar_sync(1, T);  // every list has been read: the peers may overwrite this buffer half
        peer_send_acks(x);
        return;
    }
    // publish: this GPU's k keys (just written to its own slot) to the same slot on every peer
    if (threadIdx.x < x.world && threadIdx.x != x.rank) peer_wait_ack(x, threadIdx.x, x.need_ack[threadIdx.x]);
    named_bar_sync(1, T);
    for (uint32_t i = threadIdx.x; i < x.world * k; i += T) {
        const uint32_t g = i / k, j = i - g * k;
        if (g != x.rank)
            x.gather[g][slot + j] = __ldcg(reinterpret_cast<const unsigned long long *>(x.gather[x.rank] + slot + j));
    }
    __threadfence_system();
    named_bar_sync(1, T);
    if (threadIdx.x < x.world) {
        st_release_sys(x.flags[threadIdx.x] + x.buf * x.world + x.rank, x.seq);
        // wait for peer `threadIdx.x`'s keys to have landed here (bounded: a missing peer traps)
        const uint32_t *f = x.flags[x.rank] + x.buf * x.world + threadIdx.x;
        uint32_t spins = 0;
        while (ld_acquire_sys(f) != x.seq) {
            if (++spins > (1u << 26)) __trap();
        }
    }
    named_bar_sync(1, T);
    few_lists_merge<1, true>(x.merge, work, threadIdx.x, T);  // world <= kMaxPeers lists of k <= 128 keys
    named_bar_sync(1, T);
    peer_send_acks(x);
}

}  // namespace qz
