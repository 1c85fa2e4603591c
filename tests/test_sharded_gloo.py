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
