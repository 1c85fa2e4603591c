"""Point-id sharding of an encoded corpus over the GPUs of one box.

One process per GPU.  Rank g holds rows [floor(g*N/G), floor((g+1)*N/G)) of the SAME host storage
(a byte-range slice, no re-encoding) with id_base = its first row, so point ids stay global.  The
only exchange per query is the all-gather of each rank's k best u64 keys (8*k bytes per rank) over
`torch.distributed` (NCCL on GPUs), followed by the G*k -> k merge on every rank.  Keys order
(score, then lower id) identically everywhere, so the result does not depend on G.

The scan and the merge are libqz_cuda calls (qz_scan_topk_dev / qz_merge_topk_dev); this module is
the host plumbing around them: shard arithmetic, buffers, the collective.  The two device steps are
injectable (`local_topk`, `merge`) so that the plumbing can be exercised with world_size > 1 on a
CPU box over gloo (tests/test_sharded_gloo.py), where no CUDA device exists.
"""
import ctypes as C

import numpy as np


def shard_range(count, rank, world):
    """rows [lo, hi) of rank `rank`: contiguous, sizes differ by at most 1, union = [0, count)"""
    return (rank * count) // world, ((rank + 1) * count) // world


def shard_storage(storage, row_bytes, count, rank, world):
    """(byte slice of a flat EncodedStorage for this rank, rows in it, id_base)"""
    lo, hi = shard_range(count, rank, world)
    return storage[lo * row_bytes: hi * row_bytes], hi - lo, lo


def f32_order_bits(u):
    u = np.asarray(u, np.uint32)
    return np.where(u & np.uint32(0x80000000), ~u, u | np.uint32(0x80000000)).astype(np.uint32)


def keys_to_results(keys, largest):
    """decode u64 keys (csrc/common.cuh make_key) into (scores f32, ids u32); empty slots (key 0) are dropped"""
    keys = np.asarray(keys, np.uint64)
    keys = keys[keys != 0]
    o = (keys >> np.uint64(32)).astype(np.uint32)
    if not largest:
        o = ~o
    u = np.where(o & np.uint32(0x80000000), o & np.uint32(0x7FFFFFFF), ~o).astype(np.uint32)
    ids = (np.uint32(0xFFFFFFFF) - (keys & np.uint64(0xFFFFFFFF)).astype(np.uint32)).astype(np.uint32)
    return u.view(np.float32), ids


def results_to_keys(scores, ids, largest):
    """inverse of keys_to_results (host-side packing, used by the tests)"""
    o = f32_order_bits(np.ascontiguousarray(scores, np.float32).view(np.uint32))
    if not largest:
        o = ~o
    return (o.astype(np.uint64) << np.uint64(32)) | (np.uint32(0xFFFFFFFF) - np.asarray(ids, np.uint32)).astype(np.uint64)


class ShardedSearch:
    """search_top_k over all ranks' shards; every rank returns the same (scores, ids)."""

    def __init__(self, index=None, group=None, device=None, *, local_topk=None, merge=None, peer=False, root=None):
        """peer=True: exchange through peer memory inside the scan launch (qz_scan_topk_peer_dev: one kernel per
        query per GPU, no collective call) instead of all-gather + merge.  k <= 128.
        root=r (with peer=True): the rooted gather (qz_scan_topk_gather_dev) — only rank r ends with the global top-k,
        the other ranks publish their keys and never wait."""
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.index = index
        self.device = device if device is not None else torch.device("cpu")
        self._local_topk = local_topk if local_topk is not None else self._gpu_local_topk
        self._merge = merge if merge is not None else self._gpu_merge
        self._bufs = {}
        self._peer = None
        self.root = root
        if peer and self.world > 1:
            self._peer_setup()

    # ---- device steps (the product path) -----------------------------------------------------------
    def _buffers(self, k):
        if k not in self._bufs:
            t = self.torch
            self._bufs[k] = (t.zeros(k, dtype=t.int64, device=self.device),
                             t.zeros(self.world * k, dtype=t.int64, device=self.device),
                             t.zeros(k, dtype=t.float32, device=self.device),
                             t.zeros(k, dtype=t.int32, device=self.device))
        return self._bufs[k]

    def _gpu_local_topk(self, d_query, k, largest, d_keys):
        self.index.search_top_k_dev(d_query, k, largest, d_keys.data_ptr(), 0, 0)

    def _gpu_merge(self, d_all, n_lists, k, largest, d_scores, d_ids):
        ctx = self.index.ctx
        ctx.check(ctx._L.qz_merge_topk_dev(ctx.handle, C.c_void_p(d_all.data_ptr()), n_lists, k, int(bool(largest)),
                                           None, C.c_void_p(d_scores.data_ptr()), C.c_void_p(d_ids.data_ptr())))

    # ---- peer-memory exchange (csrc/peer.cu) ----------------------------------------------------------
    PEER_HANDLE_BYTES = 64

    def _peer_setup(self):
        """create this rank's symmetric buffer, all-gather the CUDA IPC handles, map the peers"""
        t, ctx = self.torch, self.index.ctx
        mine = np.zeros(self.PEER_HANDLE_BYTES, np.uint8)
        h = C.c_void_p()
        ctx.check(ctx._L.qz_peer_create(ctx.handle, self.world, self.rank, C.c_void_p(mine.ctypes.data), C.byref(h)))
        d_mine = t.from_numpy(mine).to(self.device)
        d_all = t.zeros(self.world * self.PEER_HANDLE_BYTES, dtype=t.uint8, device=self.device)
        self.dist.all_gather_into_tensor(d_all, d_mine, group=self.group)
        handles = np.ascontiguousarray(d_all.cpu().numpy())
        ctx.check(ctx._L.qz_peer_connect(h, C.c_void_p(handles.ctypes.data)))
        self.dist.barrier(group=self.group)  # every rank has mapped every buffer before the first scan
        self._peer = h

    def close(self):
        if self._peer is not None:
            self.dist.barrier(group=self.group)  # no peer may still be writing into this rank's buffer
            self.index.ctx._L.qz_peer_destroy(self._peer)
            self._peer = None

    def _order(self, ctx_first):
        """Make work on one of (the qz Context's stream, torch's current stream) wait for what has been enqueued on
        the other.  A no-op when they are the same stream (bench.py creates the Context on torch's stream) and on
        the CPU test path; with the default Context (its own non-blocking stream) the all-gather could otherwise
        read the keys before the scan has written them."""
        if self.index is None or self.device.type != "cuda":
            return
        t = self.torch
        cur = t.cuda.current_stream(self.device)
        cs = self.index.ctx.stream_handle
        if cs == cur.cuda_stream:
            return
        ext = t.cuda.ExternalStream(cs, device=self.device)
        ev = t.cuda.Event()
        if ctx_first:
            ev.record(ext)
            cur.wait_event(ev)
        else:
            ev.record(cur)
            ext.wait_event(ev)

    # ---- the step ------------------------------------------------------------------------------------
    def search_top_k_dev(self, d_query, k, largest=True):
        """asynchronous on the current stream; returns device tensors (scores[k], ids[k]) valid on every rank"""
        d_keys, d_all, d_scores, d_ids = self._buffers(k)
        if self._peer is not None:
            ctx = self.index.ctx
            if self.root is not None:
                ctx.check(ctx._L.qz_scan_topk_gather_dev(self.index.handle, self._peer, self.root, C.c_void_p(d_query), k,
                                                         int(bool(largest)), None, C.c_void_p(d_scores.data_ptr()),
                                                         C.c_void_p(d_ids.data_ptr())))
                return d_scores, d_ids  # (the global result on rank `root` only)
            ctx.check(ctx._L.qz_scan_topk_peer_dev(self.index.handle, self._peer, C.c_void_p(d_query), k,
                                                   int(bool(largest)), None, C.c_void_p(d_scores.data_ptr()),
                                                   C.c_void_p(d_ids.data_ptr())))
            return d_scores, d_ids
        self._local_topk(d_query, k, largest, d_keys)
        self._order(ctx_first=True)   # the collective (torch's stream) reads what the scan (context stream) wrote
        if self.world > 1:
            self.dist.all_gather_into_tensor(d_all, d_keys, group=self.group)
        else:
            d_all.copy_(d_keys)
        self._order(ctx_first=False)  # the merge (context stream) reads what the collective wrote
        self._merge(d_all, self.world, k, largest, d_scores, d_ids)
        self._order(ctx_first=True)   # the caller reads the results on torch's stream
        return d_scores, d_ids

    def search(self, query_f32, k, largest=True):
        """host f32 query -> the global (scores, ids) on every rank: encode_query, scan, exchange and the copies
        in one library call (peer mode only)"""
        if self._peer is None:
            raise RuntimeError("ShardedSearch.search needs peer=True and world > 1")
        from .encoded_vectors import ptr
        ctx = self.index.ctx
        q = np.ascontiguousarray(query_f32, np.float32).reshape(-1)
        scores, ids, n, p_scores, p_ids, p_n = self.index._topk_out(k)
        if self.root is not None:
            ctx.check(ctx._L.qz_search_topk_gather(self.index.handle, self._peer, self.root, ptr(q), k,
                                                   int(bool(largest)), p_scores, p_ids, p_n))
            return scores[: n.value].copy(), ids[: n.value].copy()  # empty on the ranks that are not the root
        ctx.check(ctx._L.qz_search_topk_peer(self.index.handle, self._peer, ptr(q), k, int(bool(largest)), p_scores,
                                             p_ids, p_n))
        return scores[: n.value].copy(), ids[: n.value].copy()

    def search_top_k(self, d_query, k, largest=True, total_count=None):
        s, i = self.search_top_k_dev(d_query, k, largest)
        s, i = s.cpu().numpy(), i.cpu().numpy().view(np.uint32)
        n = k if total_count is None else min(k, total_count)
        return s[:n].copy(), i[:n].copy()


# =====================================================================================================
# sharded encode ("Encode: shard rows the same way; min/max needs one all-reduce of 2 floats")
# =====================================================================================================
QUANTILE_SAMPLE = 100_000  # quantile.rs:30


def _splitmix64(z):
    z = z + np.uint64(0x9E3779B97F4A7C15)
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def sample_ids(count, n, seed=0):
    """The library's default sample (csrc/encode_api.cu pick_sample_ids, same arithmetic): n distinct ascending ids —
    the first n outputs of a 4-round Feistel permutation keyed by `seed`, cycle-walked into [0, count).  The
    reference draws them from a randomly keyed permutation (quantile.rs:32-34)."""
    if n >= count:
        return np.arange(count, dtype=np.uint64)
    bits = 2
    while bits < 64 and (1 << bits) < count:
        bits += 2
    half, mask = np.uint64(bits // 2), np.uint64((1 << (bits // 2)) - 1)
    with np.errstate(over="ignore"):
        keys = [_splitmix64(np.uint64((seed * 4 + r) & 0xFFFFFFFFFFFFFFFF)) for r in range(4)]

        def rounds(x):
            l, r = x >> half, x & mask
            for k in keys:
                l, r = r, l ^ (_splitmix64(r ^ k) & mask)
            return (l << half) | r

        x = rounds(np.arange(n, dtype=np.uint64))
        while True:
            out = x >= np.uint64(count)
            if not out.any():
                break
            x[out] = rounds(x[out])
    return np.sort(x)


def all_gather_sample(self, d_rows, ids, count, dim):
    """rows `ids` (ascending, global) of a corpus whose rows are spread over the ranks -> the same [len(ids), dim]
    matrix on every rank.  Equal-size all-gather: every rank's share is padded to the largest share and the padding
    dropped afterwards; ranks own ascending id ranges, so concatenation in rank order keeps the sample in id order."""
    t, dist = self.torch, self.dist
    lo, hi = shard_range(count, self.rank, self.world)
    mine = ids[(ids >= lo) & (ids < hi)] - np.uint64(lo)
    d_mine = self._gather_rows(d_rows, mine)
    if self.world == 1:
        return d_mine.contiguous()
    counts = [int(((ids >= a) & (ids < b)).sum()) for a, b in (shard_range(count, r, self.world) for r in range(self.world))]
    cap = max(max(counts), 1)
    d_pad = t.zeros((cap, dim), dtype=t.float32, device=self.device)
    d_pad[:counts[self.rank]] = d_mine
    d_all = t.zeros((self.world * cap, dim), dtype=t.float32, device=self.device)
    dist.all_gather_into_tensor(d_all, d_pad, group=self.group)
    return t.cat([d_all[r * cap: r * cap + counts[r]] for r in range(self.world)]).contiguous()


class ShardedEncodeU8:
    """EncodedVectorsU8::encode (u8.rs:34-140) for a corpus whose rows are spread over the ranks (rank g holds the
    f32 rows shard_range(count, g, G)).  Global quantities are combined across ranks, then every rank encodes its own
    rows; the shards are byte-identical to slices of a single-GPU encode and (alpha, offset, multiplier) are the same
    on every rank.  Exchanges: one all-reduce(MIN) + all-reduce(MAX) of a float each; with a quantile, one
    all-gather of the sampled rows (<= 100 000 x dim f32 in total) so that every rank selects the same two order
    statistics.  The device steps are injectable like ShardedSearch's, for the gloo tests."""

    def __init__(self, ctx=None, group=None, device=None, *, local_minmax=None, local_quantile=None, local_encode=None,
                 gather_rows=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.ctx, self.group = torch, dist, ctx, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = device if device is not None else torch.device("cpu")
        self._minmax = local_minmax or self._gpu_minmax
        self._quantile = local_quantile or self._gpu_quantile
        self._encode = local_encode or self._gpu_encode
        self._gather_rows = gather_rows or (lambda d_rows, ids: d_rows[self.torch.as_tensor(ids.astype(np.int64), device=d_rows.device)])

    # ---- device steps (libqz_cuda) -------------------------------------------------------------------
    def _gpu_minmax(self, d_rows):
        mn, mx = C.c_float(), C.c_float()
        n, dim = d_rows.shape
        self.ctx.check(self.ctx._L.qz_u8_minmax_dev(self.ctx.handle, C.c_void_p(d_rows.data_ptr()), n, dim,
                                                    C.byref(mn), C.byref(mx)))
        return mn.value, mx.value

    def _gpu_quantile(self, d_sample, slice_size, quantile):
        found, lo, hi = C.c_int32(), C.c_float(), C.c_float()
        n, dim = d_sample.shape
        self.ctx.check(self.ctx._L.qz_u8_quantile_dev(self.ctx.handle, C.c_void_p(d_sample.data_ptr()), n, dim, slice_size,
                                                      quantile, C.byref(found), C.byref(lo), C.byref(hi)))
        return (lo.value, hi.value) if found.value else None

    def _gpu_encode(self, d_rows, vp_local, mn, mx, id_base):
        from .encoded_vectors_u8 import EncodedVectorsU8
        ix = EncodedVectorsU8(self.ctx, None, vp_local, 0.0, 0.0, 0.0, id_base)
        a, o, m = C.c_float(), C.c_float(), C.c_float()
        n, dim = d_rows.shape
        self.ctx.check(self.ctx._L.qz_u8_encode_range_dev(self.ctx.handle, C.c_void_p(d_rows.data_ptr()), n, dim,
                                                          int(vp_local.distance_type), int(vp_local.invert), mn, mx, None,
                                                          C.c_void_p(ix.storage_dev), C.byref(a), C.byref(o), C.byref(m)))
        self.ctx.check(self.ctx._L.qz_u8_index_set_params(ix.handle, a.value, o.value, m.value))
        ix.alpha, ix.offset, ix.multiplier = np.float32(a.value), np.float32(o.value), np.float32(m.value)
        return ix

    # ---- the encode ------------------------------------------------------------------------------------
    def encode(self, d_rows, vector_parameters, quantile=None, ids=None, seed=0):
        """d_rows: this rank's [hi - lo, dim] f32 rows (a torch tensor on self.device; the caller has synchronised
        its producer); vector_parameters describes the WHOLE corpus.  Returns the rank's shard (id_base = lo)."""
        from .encoded_vectors import VectorParameters
        t, dist = self.torch, self.dist
        vp = vector_parameters
        lo, hi = shard_range(vp.count, self.rank, self.world)
        assert tuple(d_rows.shape) == (hi - lo, vp.dim), "rank holds rows shard_range(count, rank, world)"
        mn, mx = self._minmax(d_rows)
        if self.world > 1:
            # The reference folds with strict < and > in id order (quantile.rs:5-19): of equal extremes — i.e. zeros
            # of different sign — it keeps the first.  Ranks hold ascending id ranges, so the per-rank (min, max)
            # pairs are gathered and folded in rank order with the same strict comparisons (a MIN / MAX all-reduce
            # would leave the sign of a zero to the collective's implementation).
            d_mine = t.tensor([mn, mx], dtype=t.float32, device=self.device)
            d_all = t.zeros(2 * self.world, dtype=t.float32, device=self.device)
            dist.all_gather_into_tensor(d_all, d_mine, group=self.group)
            pairs = d_all.cpu().numpy().reshape(self.world, 2)
            mn, mx = np.float32(np.finfo(np.float32).max), np.float32(np.finfo(np.float32).min)
            for r in range(self.world):
                if shard_range(vp.count, r, self.world)[0] == shard_range(vp.count, r, self.world)[1]:
                    continue  # an empty shard contributes the fold's identity
                if pairs[r, 0] < mn:
                    mn = pairs[r, 0]
                if pairs[r, 1] > mx:
                    mx = pairs[r, 1]
            mn, mx = float(mn), float(mx)
        if quantile is not None and vp.count >= 127 and quantile < 1.0:  # quantile.rs:27-29
            slice_size = min(vp.count, QUANTILE_SAMPLE)
            ids = sample_ids(vp.count, slice_size, seed) if ids is None else np.sort(np.asarray(ids, np.uint64))
            d_sample = all_gather_sample(self, d_rows, ids, vp.count, vp.dim)
            if self.device.type == "cuda":
                t.cuda.synchronize(self.device)  # the context runs on its own stream
            iv = self._quantile(d_sample, slice_size, float(quantile))
            if iv is not None:
                mn, mx = iv
        vp_local = VectorParameters(vp.dim, hi - lo, vp.distance_type, vp.invert)
        return self._encode(d_rows, vp_local, mn, mx, lo)


KMEANS_SAMPLE = 10_000  # pq.rs:286


class ShardedEncodePQ:
    """EncodedVectorsPQ::encode (pq.rs:56-107) for rows spread over the ranks: the <= 10 000 sampled rows of
    find_centroids (pq.rs:278-342) are all-gathered, every rank runs the same deterministic k-means on them (768
    chunk problems side by side, ~0.1 s — cheaper than broadcasting and waiting for one rank), then assigns its own
    rows.  Centroids are bit-identical on every rank and to the single-GPU encode with the same sample ids, and the
    code shards are byte slices of the single-GPU codes."""

    def __init__(self, ctx=None, group=None, device=None, *, local_train=None, local_assign=None, gather_rows=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.ctx, self.group = torch, dist, ctx, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = device if device is not None else torch.device("cpu")
        self._train = local_train or self._gpu_train
        self._assign = local_assign or self._gpu_assign
        self._gather_rows = gather_rows or (lambda d_rows, ids: d_rows[self.torch.as_tensor(ids.astype(np.int64), device=d_rows.device)])

    def _gpu_train(self, d_sample, chunk_size, seed):
        n, dim = d_sample.shape
        cen = np.zeros((256, dim), np.float32)
        self.ctx.check(self.ctx._L.qz_pq_train_dev(self.ctx.handle, C.c_void_p(d_sample.data_ptr()), n, dim, chunk_size, None, 0,
                                                   seed, 100, 1e-5, None, C.c_void_p(cen.ctypes.data), None))
        return cen

    def _gpu_assign(self, d_rows, vp_local, chunk_size, centroids, id_base):
        from .encoded_vectors_pq import EncodedVectorsPQ
        ix = EncodedVectorsPQ(self.ctx, None, vp_local, chunk_size, centroids, id_base)
        n, dim = d_rows.shape
        self.ctx.check(self.ctx._L.qz_pq_assign_dev(self.ctx.handle, C.c_void_p(d_rows.data_ptr()), n, dim, chunk_size,
                                                    C.c_void_p(ix.centroids.ctypes.data), None, C.c_void_p(ix.storage_dev)))
        return ix

    def encode(self, d_rows, vector_parameters, chunk_size, ids=None, seed=0):
        from .encoded_vectors import VectorParameters
        vp = vector_parameters
        lo, hi = shard_range(vp.count, self.rank, self.world)
        assert tuple(d_rows.shape) == (hi - lo, vp.dim), "rank holds rows shard_range(count, rank, world)"
        ids = sample_ids(vp.count, min(vp.count, KMEANS_SAMPLE), seed) if ids is None else np.sort(np.asarray(ids, np.uint64))
        d_sample = all_gather_sample(self, d_rows, ids, vp.count, vp.dim)
        if self.device.type == "cuda":
            self.torch.cuda.synchronize(self.device)  # the context runs on its own stream
        centroids = self._train(d_sample, chunk_size, seed)
        return self._assign(d_rows, VectorParameters(vp.dim, hi - lo, vp.distance_type, vp.invert), chunk_size, centroids, lo)
