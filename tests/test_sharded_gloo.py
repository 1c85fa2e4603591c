"""world_size-2 (and 3) runs of the sharded search plumbing over gloo on the CPU box: shard arithmetic,
id_base, key packing, the all-gather and the merge order.  The two device steps of ShardedSearch are
replaced by oracle-driven stand-ins (no CUDA device here); on the GPU box the same class runs with the
libqz_cuda calls (tests/test_gpu_sharded.py, bench.py --gpus N)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle as O
from quantization_b200 import sharded as S


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, codec, count, dim, k, largest, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        if codec == "u8":
            data = O.uniform(42, count * dim).reshape(count, dim)
            full = O.OracleU8.encode(data, O.DOT, False)
            rb = full.stride
            sl, n, base = S.shard_storage(full.storage, rb, count, rank, world)
            shard = O.OracleU8(sl, n, dim, O.DOT, False, full.alpha, full.offset, full.multiplier)
            q = full.encode_query(O.uniform(43, dim))
        else:
            data = O.signs(42, count * dim).reshape(count, dim)
            full = O.OracleBin.encode(data, O.DOT, False, 16)
            rb = full.row_bytes
            sl, n, base = S.shard_storage(full.storage, rb, count, rank, world)
            shard = O.OracleBin(sl, n, dim, 16, O.DOT, False)
            q = full.encode_query(O.signs(43, dim))

        def local_topk(query, kk, lg, d_keys):  # stand-in for qz_scan_topk_dev
            s, i = shard.search(query, kk, lg, id_base=base)
            keys = np.zeros(kk, np.uint64)
            keys[: s.size] = S.results_to_keys(s, i, lg)
            d_keys.copy_(torch.from_numpy(keys.view(np.int64)))

        def merge(d_all, n_lists, kk, lg, d_scores, d_ids):  # stand-in for qz_merge_topk_dev
            keys = np.sort(d_all.numpy().view(np.uint64))[::-1][:kk]
            s, i = S.keys_to_results(keys, lg)
            d_scores.zero_(); d_ids.zero_()
            d_scores[: s.size] = torch.from_numpy(s.copy())
            d_ids[: i.size] = torch.from_numpy(i.view(np.int32).copy())

        ss = S.ShardedSearch(None, local_topk=local_topk, merge=merge)
        assert ss.world == world and ss.rank == rank
        s, i = ss.search_top_k(q, k, largest, total_count=count)
        ws, wi = full.search(q, k, largest)
        ok = np.array_equal(i, wi) and np.array_equal(s.view(np.uint32), ws.view(np.uint32))
        out[rank] = int(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("codec,count,dim,k,largest", [("u8", 1000, 65, 100, True), ("u8", 257, 32, 30, False),
                                                       ("bin", 2000, 128, 100, True), ("u8", 50, 16, 100, True)])
def test_sharded_search_matches_single_shard(world, codec, count, dim, k, largest):
    out = mp.get_context("spawn").Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), codec, count, dim, k, largest, out), nprocs=world, join=True)
    assert [out.get(r) for r in range(world)] == [1] * world


def test_shard_ranges_cover():
    for count in (0, 1, 7, 1000, 10_000_001):
        for world in (1, 2, 3, 4, 8):
            r = [S.shard_range(count, g, world) for g in range(world)]
            assert r[0][0] == 0 and r[-1][1] == count
            assert all(r[g][1] == r[g + 1][0] for g in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1


def test_key_roundtrip_and_order():
    s = np.array([3.5, -0.0, 0.0, -2.0, np.inf, -np.inf, 3.5], np.float32)
    i = np.array([7, 1, 2, 3, 4, 5, 6], np.uint32)
    for lg in (True, False):
        k = S.results_to_keys(s, i, lg)
        s2, i2 = S.keys_to_results(k, lg)
        assert np.array_equal(s2.view(np.uint32), s.view(np.uint32)) and np.array_equal(i2, i)
        assert all(int(k[j]) == O.make_key(float(s[j]), int(i[j]), lg) for j in range(len(s)))
    k = S.results_to_keys(s, i, True)
    assert k[6] > k[0]  # equal scores: the lower id has the larger key


# ---- sharded encode: min/max all-reduce, sample all-gather, byte-identical shards ------------------
def _encode_worker(rank, world, port, count, dim, quantile, out):
    import ctypes as C
    from oracle import oracle as OM
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        data = O.uniform(42, count * dim, -1.0, 3.0).reshape(count, dim)
        ids = np.sort(np.random.default_rng(5).permutation(count)[: min(count, 300)]).astype(np.uint64) if quantile else None
        full = O.OracleU8.encode(data, O.L2, True, quantile, ids)
        lo, hi = S.shard_range(count, rank, world)
        got = {}

        def local_encode(rows, vp_local, mn, mx, id_base):  # stand-in for qz_u8_encode_range_dev
            a, o = C.c_float(), C.c_float()
            OM._lib.qzo_u8_alpha_offset(C.c_float(mn), C.c_float(mx), C.byref(a), C.byref(o))
            r = np.ascontiguousarray(rows.numpy())
            st = np.empty(r.shape[0] * (O.u8_actual_dim(dim) + 4), np.uint8)
            OM._lib.qzo_u8_encode(OM._p(r, OM._f32p), C.c_uint64(r.shape[0]), C.c_uint32(dim), C.c_int(O.L2), C.c_int(1), a, o,
                                  OM._p(st, OM._u8p))
            got.update(alpha=a.value, offset=o.value, id_base=id_base)
            return st

        enc = S.ShardedEncodeU8(
            None, local_minmax=lambda rows: O.minmax(rows.numpy()),
            local_quantile=lambda sample, slice_size, q: O.quantile_interval(sample.numpy(), count, slice_size, q),
            local_encode=local_encode)
        from quantization_b200 import VectorParameters, DistanceType
        st = enc.encode(torch.from_numpy(data[lo:hi].copy()), VectorParameters(dim, count, DistanceType.L2, True), quantile, ids)
        rb = full.stride
        ok = (np.array_equal(st, full.storage[lo * rb: hi * rb]) and np.float32(got["alpha"]) == full.alpha
              and np.float32(got["offset"]) == full.offset and got["id_base"] == lo)
        out[rank] = int(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("count,dim,quantile", [(1000, 33, None), (1000, 33, 0.9), (131, 16, 0.99), (5, 8, None)])
def test_sharded_encode_is_a_slice_of_the_global_encode(world, count, dim, quantile):
    out = mp.Manager().list([0] * world)
    mp.spawn(_encode_worker, args=(world, _free_port(), count, dim, quantile, out), nprocs=world, join=True)
    assert list(out) == [1] * world


def _pq_encode_worker(rank, world, port, count, dim, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        data = O.uniform(42, count * dim).reshape(count, dim)
        ids = np.sort(np.random.default_rng(5).permutation(count)[:300]).astype(np.uint64)
        bounds = O.pq_division(dim, 2)
        full = O.OraclePQ.encode(data, O.DOT, False, 2, sample_ids=ids)
        lo, hi = S.shard_range(count, rank, world)
        seen = {}

        def train(sample, chunk_size, seed):  # stand-in for qz_pq_train_dev on the gathered sample
            seen["sample_ok"] = np.array_equal(sample.numpy(), data[ids.astype(np.int64)])
            return O.OraclePQ.find_centroids(sample.numpy(), bounds)

        def assign(rows, vp_local, chunk_size, cen, id_base):  # stand-in for qz_pq_assign_dev
            seen["id_base"] = id_base
            return cen, O.OraclePQ.assign(rows.numpy(), bounds, cen)

        from quantization_b200 import VectorParameters, DistanceType
        enc = S.ShardedEncodePQ(None, local_train=train, local_assign=assign)
        cen, codes = enc.encode(torch.from_numpy(data[lo:hi].copy()), VectorParameters(dim, count, DistanceType.Dot, False), 2, ids)
        nch = dim // 2
        ok = (seen["sample_ok"] and seen["id_base"] == lo and np.array_equal(cen, full.centroids)
              and np.array_equal(codes, full.codes[lo * nch: hi * nch]))
        out[rank] = int(ok)
    finally:
        dist.destroy_process_group()


def test_sharded_pq_encode_gathers_the_sample_and_assigns_locally():
    world = 2
    out = mp.Manager().list([0] * world)
    mp.spawn(_pq_encode_worker, args=(world, _free_port(), 700, 8, out), nprocs=world, join=True)
    assert list(out) == [1] * world
