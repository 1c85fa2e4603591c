"""Sharded search on the GPU: G shards of one corpus resident on one device (a 1-GPU box can still run the
whole path: per-shard qz_scan_topk_dev, concatenated keys = what the all-gather delivers, qz_merge_topk_dev),
compared with the single-shard oracle.  Multi-process NCCL runs are bench.py --gpus N."""
import ctypes as C

import numpy as np
import pytest
import torch

import oracle as O
import quantization_b200 as qz
from quantization_b200 import sharded as S

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("G", [1, 2, 4, 8])
@pytest.mark.parametrize("codec", ["u8", "bin"])
def test_shards_on_one_device(ctx, G, codec):
    count, dim, k = 20011, 256, 100
    dev = torch.device("cuda", 0)
    if codec == "u8":
        data = O.uniform(42, count * dim).reshape(count, dim)
        full = O.OracleU8.encode(data, O.DOT, False)
        rb, q = full.stride, full.encode_query(O.uniform(43, dim))
    else:
        data = O.signs(42, count * dim).reshape(count, dim)
        full = O.OracleBin.encode(data, O.DOT, False, 16)
        rb, q = full.row_bytes, full.encode_query(O.signs(43, dim))
    d_q = torch.from_numpy(q.copy()).to(dev)
    d_all = torch.zeros(G * k, dtype=torch.int64, device=dev)
    shards = []
    for g in range(G):
        sl, n, base = S.shard_storage(full.storage, rb, count, g, G)
        vp = qz.VectorParameters(dim, n, qz.DistanceType.Dot, False)
        ix = (qz.EncodedVectorsU8(ctx, sl.copy(), vp, full.alpha, full.offset, full.multiplier, id_base=base)
              if codec == "u8" else qz.EncodedVectorsBin(ctx, sl.copy(), vp, id_base=base))
        ix.search_top_k_dev(d_q.data_ptr(), k, True, d_all[g * k:].data_ptr(), 0, 0)
        shards.append(ix)
    d_s = torch.zeros(k, dtype=torch.float32, device=dev)
    d_i = torch.zeros(k, dtype=torch.int32, device=dev)
    ctx.check(ctx._L.qz_merge_topk_dev(ctx.handle, C.c_void_p(d_all.data_ptr()), G, k, 1, None,
                                       C.c_void_p(d_s.data_ptr()), C.c_void_p(d_i.data_ptr())))
    ctx.synchronize()
    ws, wi = full.search(q, k, True)
    assert np.array_equal(d_i.cpu().numpy().view(np.uint32), wi)
    assert np.array_equal(d_s.cpu().numpy().view(np.uint32), ws.view(np.uint32))
    # host decode of the gathered keys agrees with the device merge
    keys = np.sort(d_all.cpu().numpy().view(np.uint64))[::-1][:k]
    hs, hi = S.keys_to_results(keys, True)
    assert np.array_equal(hi, wi)
    for ix in shards:
        ix.close()


def test_sharded_search_class_world1():
    """ShardedSearch mixes torch ops (copy / all-gather) with library launches, so the library context
    must run on torch's current stream (as bench.py does)."""
    count, dim, k = 5000, 128, 50
    stream = torch.cuda.Stream(device=0)
    torch.cuda.set_stream(stream)
    ctx = qz.Context(0, stream=stream.cuda_stream)
    data = O.uniform(1, count * dim).reshape(count, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.L2, True)
    ix = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, lambda: False, ctx=ctx)
    o = O.OracleU8.encode(data, O.L2, True)
    q = ix.encode_query(O.uniform(2, dim))
    dev = torch.device("cuda", 0)
    ss = S.ShardedSearch(ix, device=dev)
    d_q = torch.from_numpy(q.copy()).to(dev)  # (kept alive: the call takes a raw device pointer)
    s, i = ss.search_top_k(d_q.data_ptr(), k, True, total_count=count)
    ws, wi = o.search(q, k, True)
    assert np.array_equal(i, wi) and np.array_equal(s.view(np.uint32), ws.view(np.uint32))
    torch.cuda.set_stream(torch.cuda.default_stream(0))
    ix.close()


@pytest.mark.parametrize("G", [2, 3])
@pytest.mark.parametrize("count,dim,quantile,dist", [(20_000, 1536, None, qz.DistanceType.Dot), (150_000, 64, 0.99, qz.DistanceType.L2),
                                                     (3000, 33, 0.9, qz.DistanceType.L1)])
def test_sharded_encode_steps_on_one_device(ctx, G, count, dim, quantile, dist):
    """the three C-ABI steps of a sharded encode (qz_u8_minmax_dev / qz_u8_quantile_dev / qz_u8_encode_range_dev), run
    for G row shards of one device-resident matrix with the reductions done on the host: every shard is a byte
    slice of the single-GPU qz_u8_encode_dev result and the Metadata scalars agree"""
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(3)
    data = torch.rand((count, dim), generator=g, device=dev) * 4 - 1
    torch.cuda.synchronize()
    vp = qz.VectorParameters(dim, count, dist, True)
    whole = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, quantile, ctx=ctx, seed=1)
    ref = whole.download()
    enc = S.ShardedEncodeU8(ctx, device=dev)
    parts = [data[lo:hi] for lo, hi in (S.shard_range(count, r, G) for r in range(G))]
    mm = [enc._gpu_minmax(p) for p in parts]
    mn, mx = min(m[0] for m in mm), max(m[1] for m in mm)
    if quantile is not None:
        slice_size = min(count, S.QUANTILE_SAMPLE)
        ids = S.sample_ids(count, slice_size, 1)
        sample = data[torch.as_tensor(ids.astype(np.int64), device=dev)].contiguous()
        torch.cuda.synchronize()
        iv = enc._gpu_quantile(sample, slice_size, quantile)
        assert iv is not None
        mn, mx = iv
    rb = qz.EncodedVectorsU8.get_quantized_vector_size(vp)
    for r, p in enumerate(parts):
        lo, hi = S.shard_range(count, r, G)
        ix = enc._gpu_encode(p, qz.VectorParameters(dim, hi - lo, dist, True), mn, mx, lo)
        assert (ix.alpha, ix.offset, ix.multiplier) == (whole.alpha, whole.offset, whole.multiplier)
        assert np.array_equal(ix.download(), ref[lo * rb: hi * rb]), f"shard {r} of {G} is not a slice of the global encode"
        # and it scans with global ids
        q = whole.encode_query(O.uniform(9, dim))
        s, i = ix.search_top_k(q, 10, True)
        assert i.min() >= lo and i.max() < hi


@pytest.mark.parametrize("G", [2, 3])
def test_sharded_pq_encode_steps_on_one_device(ctx, G):
    """ShardedEncodePQ's device steps for G row shards of one matrix: k-means on the gathered sample gives the same
    centroids as the single-GPU encode with the same sample ids, and the code shards are byte slices of its codes"""
    count, dim, chunk = 30_000, 64, 2
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(5)
    data = torch.rand((count, dim), generator=g, device=dev)
    torch.cuda.synchronize()
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    whole = qz.EncodedVectorsPQ.encode_dev(data.data_ptr(), vp, chunk, ctx=ctx, seed=2)
    ref = whole.download()
    enc = S.ShardedEncodePQ(ctx, device=dev)
    ids = S.sample_ids(count, S.KMEANS_SAMPLE, 2)
    sample = data[torch.as_tensor(ids.astype(np.int64), device=dev)].contiguous()
    torch.cuda.synchronize()
    cen = enc._gpu_train(sample, chunk, 2)
    assert np.array_equal(cen.view(np.uint32), whole.centroids.view(np.uint32)), "k-means on the gathered sample differs"
    nch = dim // chunk
    for r in range(G):
        lo, hi = S.shard_range(count, r, G)
        part = data[lo:hi].contiguous(); torch.cuda.synchronize()
        ix = enc._gpu_assign(part, qz.VectorParameters(dim, hi - lo, qz.DistanceType.Dot, False), chunk, cen, lo)
        assert np.array_equal(ix.download(), ref[lo * nch: hi * nch])
