"""Out-of-core scan (qz_scan_topk_chunked): the storage streamed from host memory through two device buffers must
give exactly what the reference's caller loop gives over the same bytes (the oracle: score_point for every row + the
caller-side heap, demos/src/ann_benchmark_data.rs:150-167) — ragged last chunk, one chunk, many chunks, all three
codecs; and, as a second check, what one resident index gives."""
import ctypes as C

import numpy as np
import pytest

import oracle as O
import quantization_b200 as qz

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("chunk_rows", [7_000, 50_021, 80_000, 512])
def test_chunked_u8(ctx, chunk_rows):
    count, dim, k = 50_021, 128, 100
    data = O.uniform(21, count * dim).reshape(count, dim)
    query = O.uniform(22, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, ctx=ctx)
    vpc = qz.VectorParameters(dim, chunk_rows, qz.DistanceType.Dot, False)
    a = qz.EncodedVectorsU8(ctx, None, vpc, g.alpha, g.offset, g.multiplier)
    b = qz.EncodedVectorsU8(ctx, None, vpc, g.alpha, g.offset, g.multiplier)
    q = g.encode_query(query)
    host = np.asarray(g.encoded_vectors)
    o = O.OracleU8(host, count, dim, 0, False, g.alpha, g.offset, g.multiplier)
    for largest in (True, False):
        s, i = a.search_top_k_chunked(b, host, count, q, k, largest)
        os_, oi = o.search(o.encode_query(query), k, largest, threads=4)
        assert np.array_equal(i, oi) and np.array_equal(bits(s), bits(os_)), "chunked scan differs from the oracle"
        rs, ri = g.search_top_k(q, k, largest)
        assert np.array_equal(i, ri) and np.array_equal(bits(s), bits(rs))
    # pinned storage: same answer
    L = ctx._L
    ctx.check(L.qz_host_pin(ctx.handle, C.c_void_p(host.ctypes.data), host.nbytes))
    try:
        s, i = a.search_top_k_chunked(b, host, count, q, 17, True)
        rs, ri = g.search_top_k(q, 17, True)
        assert np.array_equal(i, ri) and np.array_equal(bits(s), bits(rs))
    finally:
        ctx.check(L.qz_host_unpin(ctx.handle, C.c_void_p(host.ctypes.data)))
    # fewer rows than k, and argument errors
    s, i = a.search_top_k_chunked(b, host, 40, q, k, True)
    assert s.size == 40
    with pytest.raises(qz.EncodingError):
        a.search_top_k_chunked(a, host, count, q, k, True)


def test_chunked_binary_and_pq(ctx):
    count, dim, k = 30_007, 512, 64
    data = O.uniform(23, count * dim, -1.0, 1.0).reshape(count, dim)
    query = O.uniform(24, dim, -1.0, 1.0)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsBin.encode(data, qz.VecStorageBuilder(), vp, ctx=ctx)
    vpc = qz.VectorParameters(dim, 4096, qz.DistanceType.Dot, False)
    a, b = qz.EncodedVectorsBin(ctx, None, vpc), qz.EncodedVectorsBin(ctx, None, vpc)
    q = g.encode_query(query)
    s, i = a.search_top_k_chunked(b, np.asarray(g.encoded_vectors), count, q, k, True)
    ob = O.OracleBin(np.asarray(g.encoded_vectors), count, dim, 16, 0, False)
    os_, oi = ob.search(ob.encode_query(query), k, True, threads=4)
    assert np.array_equal(i, oi) and np.array_equal(bits(s), bits(os_)), "chunked binary scan differs from the oracle"
    rs, ri = g.search_top_k(q, k, True)
    assert np.array_equal(i, ri) and np.array_equal(bits(s), bits(rs))

    n_chunks = dim // 2
    codes = np.random.default_rng(25).integers(0, 256, size=(count, n_chunks), dtype=np.uint8)
    cen = O.uniform(26, 256 * dim).reshape(256, dim)
    p = qz.EncodedVectorsPQ(ctx, codes.reshape(-1), vp, 2, cen)
    pa, pb = qz.EncodedVectorsPQ(ctx, None, vpc, 2, cen), qz.EncodedVectorsPQ(ctx, None, vpc, 2, cen)
    lut = p.encode_query(query)
    s, i = pa.search_top_k_chunked(pb, codes.reshape(-1), count, lut, k, True)
    rs, ri = p.search_top_k(lut, k, True)
    assert np.array_equal(i, ri) and np.array_equal(bits(s), bits(rs))
    # the oracle sums in the x86 order (pq.rs:405-440): scores within the 1e-5 relative contract, same rows up to
    # ties within that tolerance
    op = O.OraclePQ(codes.reshape(-1), count, dim, O.pq_division(dim, 2), cen, 0, False)
    os_, oi = op.search(lut.view(np.float32), k, True, threads=4)
    assert np.allclose(s, os_, rtol=1e-5, atol=0), "chunked PQ scan beyond 1e-5 relative of the oracle"
    assert len(set(i.tolist()) & set(oi.tolist())) >= k - 2
    # a k for which the per-chunk lists would not fit the merge scratch is refused, not overrun
    tiny = qz.VectorParameters(dim, 4, qz.DistanceType.Dot, False)
    ta, tb = qz.EncodedVectorsPQ(ctx, None, tiny, 2, cen), qz.EncodedVectorsPQ(ctx, None, tiny, 2, cen)
    with pytest.raises(qz.EncodingError):
        ta.search_top_k_chunked(tb, codes.reshape(-1), count, lut, 1000, True)  # 7502 chunks of 4 rows, Kp = 1024: at most 4096
