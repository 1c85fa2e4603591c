"""Out-of-core scan (qz_scan_topk_chunked): the storage streamed from host memory through two device buffers must
give exactly the results of one resident index — ragged last chunk, one chunk, many chunks, all three codecs."""
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
    for largest in (True, False):
        s, i = a.search_top_k_chunked(b, host, count, q, k, largest)
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
