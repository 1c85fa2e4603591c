"""Batched multi-query u8 search (tcgen05 kind::i8 path, BASELINE config 5) vs the CPU oracle: for every
query the (scores, ids) must equal search_top_k(encode_query(q)) of the oracle bit for bit — the integer
contraction is exact and the score map runs in the reference's order (u8.rs:347)."""
import numpy as np
import pytest

import oracle as O
import quantization_b200 as qz

pytestmark = pytest.mark.gpu


def check(ctx, count, dim, nq, k, dist, invert, largest, lo=0.0):
    data = O.uniform(42, count * dim, lo, 1.0).reshape(count, dim)
    queries = O.uniform(43, nq * dim, lo, 1.0).reshape(nq, dim)
    vp = qz.VectorParameters(dim, count, dist, invert)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, lambda: False, ctx=ctx)
    o = O.OracleU8.encode(data, int(dist), invert)
    s, ids = g.search_batch(queries, k, largest)
    assert s.shape == (nq, k) and ids.shape == (nq, k)
    n = min(k, count)
    for qi in range(nq):
        ws, wi = o.search(o.encode_query(queries[qi]), k, largest, threads=4)
        assert np.array_equal(ids[qi, :n], wi), f"query {qi}: ids differ"
        assert np.array_equal(s[qi, :n].view(np.uint32), ws.view(np.uint32)), f"query {qi}: scores differ"
        if n < k:
            assert np.isnan(s[qi, n:]).all() and (ids[qi, n:] == 0xFFFFFFFF).all()
    g.close()


@pytest.mark.parametrize("count,dim,nq,k", [(3000, 128, 5, 10), (3000, 65, 130, 100), (5000, 768, 64, 100),
                                            (20000, 1536, 33, 100), (300, 16, 3, 100), (9000, 200, 257, 30)])
def test_batch_dot(ctx, count, dim, nq, k):
    check(ctx, count, dim, nq, k, qz.DistanceType.Dot, False, True)


@pytest.mark.parametrize("dist,invert,largest", [(qz.DistanceType.L2, False, False), (qz.DistanceType.L2, True, True),
                                                 (qz.DistanceType.Dot, True, False)])
def test_batch_metrics(ctx, dist, invert, largest):
    check(ctx, 6000, 100, 40, 50, dist, invert, largest, lo=-1.0)


def test_batch_matches_single_query_scan_large(ctx):
    """mid-size: several chunks of the growth schedule; compare with the GPU single-query scan"""
    count, dim, nq, k = 200_000, 256, 96, 100
    data = O.uniform(7, count * dim).reshape(count, dim)
    queries = O.uniform(8, nq * dim).reshape(nq, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, lambda: False, ctx=ctx)
    s, ids = g.search_batch(queries, k, True)
    for qi in range(0, nq, 7):
        ws, wi = g.search_top_k(g.encode_query(queries[qi]), k, True)
        assert np.array_equal(ids[qi], wi) and np.array_equal(s[qi].view(np.uint32), ws.view(np.uint32))
    g.close()


def test_batch_sorted_order_takes_safe_path(ctx):
    """rows sorted by increasing score: every row beats the threshold, the fast chunk schedule overflows
    the candidate buffers and the exact safe schedule must take over"""
    count, dim, nq, k = 60_000, 32, 4, 10
    base = np.linspace(0.0, 1.0, count, dtype=np.float32)[:, None]
    data = np.repeat(base, dim, axis=1).astype(np.float32)
    queries = np.ones((nq, dim), np.float32)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, lambda: False, ctx=ctx)
    o = O.OracleU8.encode(data, 0, False)
    s, ids = g.search_batch(queries, k, True)
    ws, wi = o.search(o.encode_query(queries[0]), k, True)
    assert np.array_equal(ids[0], wi) and np.array_equal(s[0].view(np.uint32), ws.view(np.uint32))
    g.close()


def test_batch_argument_errors(ctx):
    data = O.uniform(1, 100 * 16).reshape(100, 16)
    vp = qz.VectorParameters(16, 100, qz.DistanceType.L1, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, lambda: False, ctx=ctx)
    with pytest.raises(qz.EncodingError):
        g.search_batch(data[:4], 10, True)  # L1 is not a contraction
    g.close()
