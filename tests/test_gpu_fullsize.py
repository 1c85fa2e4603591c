"""Parity at BASELINE.json's FULL sizes (configs 2, 3, 4): the encoded storage the GPU produced is downloaded once
and the CPU oracle (the reference's own AVX2 / SSE scorers from oracle/_ref where they exist) scans ALL of it —
every score and the top-100 must match bit for bit (u8, binary) or within the 1e-5 relative contract (PQ).  The
encode side is pinned on row slices (the f32 source lives on the device; three slices are brought back and
re-encoded by the oracle with the GPU's alpha/offset, which are themselves checked against the data's min/max).
Size-independent properties on top: top-k == top-k of the GPU's own full score vector, two half shards merge to
the global answer."""
import os

import ctypes as C
import numpy as np
import pytest

import oracle as O
from oracle import oracle as OM

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import quantization_b200 as qz  # noqa: E402

K = 100
THREADS = max(1, min(32, os.cpu_count() or 1))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _slices(count, n=2048):
    return [(0, n), (count // 2 - 5, n), (count - n, n)]


@pytest.mark.parametrize("dist", [qz.DistanceType.Dot, qz.DistanceType.L2])
def test_c2_u8_1m_x_1536(ctx, dist):
    count, dim = 1_000_000, 1536
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(42)
    data = torch.rand((count, dim), generator=g, device=dev)
    torch.cuda.synchronize()  # the context has its own stream
    vp = qz.VectorParameters(dim, count, dist, False)
    ix = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, None, ctx=ctx)
    # alpha / offset from the global min / max, u8.rs:228-232
    mn, mx = np.float32(data.min().item()), np.float32(data.max().item())
    assert ix.offset == mn and ix.alpha == np.float32(np.float32(mx - mn) / np.float32(127.0))
    storage = ix.download()
    stride = dim + 4
    # encode parity on row slices: oracle re-encodes the same f32 rows with the same alpha / offset
    for first, n in _slices(count):
        rows = np.ascontiguousarray(data[first:first + n].cpu().numpy())
        out = np.empty(n * stride, np.uint8)
        OM._lib.qzo_u8_encode(OM._p(rows, OM._f32p), C.c_uint64(n), C.c_uint32(dim), C.c_int(int(dist)), C.c_int(0),
                              C.c_float(ix.alpha), C.c_float(ix.offset), OM._p(out, OM._u8p))
        assert np.array_equal(out, storage[first * stride:(first + n) * stride]), f"u8 rows {first}.. differ"
    del data
    o = O.OracleU8(storage, count, dim, int(dist), False, ix.alpha, ix.offset, ix.multiplier)
    for seed in (43, 44):
        q = O.uniform(seed, dim)
        enc = ix.encode_query(q)
        assert np.array_equal(np.frombuffer(bytes(enc), np.uint8), o.encode_query(q))
        largest = dist == qz.DistanceType.Dot
        gs, gi = ix.search_top_k(enc, K, largest)
        os_, oi = o.search(o.encode_query(q), K, largest, threads=THREADS)
        assert np.array_equal(gi, oi), "top-100 ids differ from the reference scorers at 1M x 1536"
        assert np.array_equal(bits(gs), bits(os_))
        full = ix.score_all(enc)
        if seed == 43:  # every one of the 1M scores, against the reference's AVX2 kernel
            assert np.array_equal(bits(full), bits(o.scan(o.encode_query(q))))
        es, ei = O.topk(full, K, largest)
        assert np.array_equal(gi, ei) and np.array_equal(bits(gs), bits(es))
    # two shards of the same storage merge to the global answer
    half = count // 2
    a = qz.EncodedVectorsU8(ctx, storage[:half * stride], qz.VectorParameters(dim, half, dist, False), ix.alpha, ix.offset,
                            ix.multiplier, id_base=0)
    b = qz.EncodedVectorsU8(ctx, storage[half * stride:], qz.VectorParameters(dim, count - half, dist, False), ix.alpha,
                            ix.offset, ix.multiplier, id_base=half)
    sa, ia = a.search_top_k(enc, K, largest)
    sb, ib = b.search_top_k(enc, K, largest)
    ms, mi = O.merge_topk(np.concatenate([sa, sb]), np.concatenate([ia, ib]), K, largest)
    assert np.array_equal(mi, gi) and np.array_equal(bits(ms), bits(gs))


def test_c3_binary_10m_x_1536(ctx):
    count, dim, chunk = 10_000_000, 1536, 1_000_000
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(7)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    ix = qz.EncodedVectorsBin.empty(vp, ctx=ctx)
    kept = {}
    for r in range(0, count, chunk):
        part = torch.rand((chunk, dim), generator=g, device=dev) - 0.5
        torch.cuda.synchronize()
        ix.encode_rows_dev(part.data_ptr(), r, chunk)
        ctx.synchronize()
        kept[r] = np.ascontiguousarray(part[:1024].cpu().numpy())
        del part
    storage = ix.download()
    rb = 192
    assert ix.row_bytes == rb
    for r, rows in kept.items():  # encode parity on 1024 rows of every chunk
        ob = O.OracleBin.encode(rows, 0, False, 16)
        assert np.array_equal(ob.storage, storage[r * rb:(r + 1024) * rb]), f"binary rows {r}.. differ"
    o = O.OracleBin(storage, count, dim, 16, 0, False)
    q = O.signs(8, dim)
    enc = ix.encode_query(q)
    assert np.array_equal(np.frombuffer(bytes(enc), np.uint8), o.encode_query(q))
    gs, gi = ix.search_top_k(enc, K, True)
    os_, oi = o.search(o.encode_query(q), K, True, threads=THREADS)
    assert np.array_equal(gi, oi) and np.array_equal(bits(gs), bits(os_)), "top-100 differs at 10M x 1536 bit"
    full = ix.score_all(enc)
    assert np.array_equal(bits(full), bits(o.scan(o.encode_query(q)))), "xor-popcount scores differ from sse.c"
    es, ei = O.topk(full, K, True)
    assert np.array_equal(gi, ei)


def test_c4_pq_1m_x_768_codes(ctx):
    count, dim, chunk_size = 1_000_000, 1536, 2
    n_chunks = dim // chunk_size
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(9)
    data = torch.rand((count, dim), generator=g, device=dev)
    cen = O.uniform(77, 256 * dim).reshape(256, dim)
    torch.cuda.synchronize()
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    ix = qz.EncodedVectorsPQ.encode_dev(data.data_ptr(), vp, chunk_size, ctx=ctx, centroids=cen)
    codes = ix.download()
    bounds = O.pq_division(dim, chunk_size)
    for first, n in _slices(count, 1024):  # assignment is bit-exact given the centroids (pq.rs:237-265)
        rows = np.ascontiguousarray(data[first:first + n].cpu().numpy())
        assert np.array_equal(O.OraclePQ.assign(rows, bounds, cen), codes[first * n_chunks:(first + n) * n_chunks])
    del data
    o = O.OraclePQ(codes, count, dim, bounds, cen, 0, False)
    q = O.uniform(78, dim)
    glut = ix.encode_query(q)
    lut = o.encode_query(q)
    assert np.allclose(np.frombuffer(bytes(glut), np.float32), lut, rtol=1e-5, atol=0)
    full = ix.score_all(glut)
    ref = o.scan(lut)  # x86 summation order of score_point_sse, all 1M rows
    assert np.allclose(full, ref, rtol=1e-5, atol=0), "PQ scores beyond 1e-5 relative of the reference order"
    gs, gi = ix.search_top_k(glut, K, True)
    es, ei = O.topk(full, K, True)  # exact top-k of the GPU's own scores
    assert np.array_equal(gi, ei) and np.array_equal(bits(gs), bits(es))
    # against the oracle's top-k: the same rows except where scores tie within the tolerance
    os_, oi = o.search(lut, K, True, threads=THREADS)
    assert np.allclose(np.sort(gs), np.sort(os_), rtol=1e-5, atol=0)
    assert len(set(gi.tolist()) & set(oi.tolist())) >= K - 2


@pytest.mark.parametrize("dim", [128, 1024])
def test_c1_u8_100k_host_path(ctx, dim):
    """BASELINE config 1 (the reference's own bench shape, demos/benches/encode.rs:15-16, 42-49, 91-102): host f32
    data -> encode through the host C-ABI call -> linear scan, random-access score_point and score_internal.  Inputs are
    SURVEY 8d's (SplitMix64 seeds 42 / 43); everything is compared with the oracle on ALL 100k rows."""
    count = 100_000
    data = O.uniform(42, count * dim).reshape(count, dim)
    query = O.uniform(43, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, lambda: False, ctx=ctx)
    o = O.OracleU8.encode(data, 0, False)
    assert np.array_equal(np.asarray(g.encoded_vectors), o.storage)
    assert (g.alpha, g.offset, g.multiplier) == (o.alpha, o.offset, o.multiplier)
    enc = g.encode_query(query)
    assert np.array_equal(np.frombuffer(bytes(enc), np.uint8), o.encode_query(query))
    assert np.array_equal(bits(g.score_all(enc)), bits(o.scan(o.encode_query(query))))
    gs, gi = g.search_top_k(enc, K, True)
    os_, oi = o.search(o.encode_query(query), K, True, threads=THREADS)
    assert np.array_equal(gi, oi) and np.array_equal(bits(gs), bits(os_))
    rng = np.random.default_rng(1)
    ids = rng.integers(0, count, 5000).astype(np.uint32)  # the bench's random access pattern
    assert np.array_equal(bits(g.score_points(enc, ids)), bits(o.scan(o.encode_query(query))[ids]))
    ia, ib = rng.integers(0, count, 2000).astype(np.uint32), rng.integers(0, count, 2000).astype(np.uint32)
    assert np.array_equal(bits(g.score_pairs(ia, ib)), bits(o.score_pairs(ia, ib)))
    assert bits(np.float32(g.score_point(enc, 123)))[()] == bits(np.float32(o.score_point(o.encode_query(query), 123)))[()]


def test_c5_batch_1p25m_x_768(ctx):
    """BASELINE config 5 at one GPU's share of its configured size: 4096 queries x 1.25M rows x 768-d through
    qz_u8_search_batch_dev (tcgen05 int8 contraction).  This is the size at which the batched path walks its whole
    chunk schedule (bootstrap chunks, growing chunks, threshold refresh between them).  Every query's top-100 must be
    bit-identical to the single-query scan (row_scan_kernel, itself pinned against the reference scorers in
    test_c2_*), and 32 of them are compared with the oracle directly — encode_query (u8.rs:290-329), the reference's
    AVX2 dot kernel (cpp/avx2.c:31-69), the score map of u8.rs:347 and the caller-side heap — on all 1.25M rows."""
    count, dim, nq = 1_250_000, 768, 4096
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(42)
    data = torch.rand((count, dim), generator=g, device=dev)
    queries = torch.rand((nq, dim), generator=g, device=dev)
    torch.cuda.synchronize()  # the context has its own stream
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    ix = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, None, ctx=ctx)
    del data
    d_s = torch.zeros((nq, K), dtype=torch.float32, device=dev)
    d_i = torch.zeros((nq, K), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    for rep in range(2):  # the second call reuses the device-side mirror and scratch of the first
        ix.search_batch_dev(queries.data_ptr(), nq, K, True, d_s.data_ptr(), d_i.data_ptr())
    ctx.synchronize()
    bs, bi = d_s.cpu().numpy(), d_i.cpu().numpy().view(np.uint32)
    # every query against the single-query scan on the device
    d_enc = torch.zeros((nq, dim + 4), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    ix.encode_query_dev(queries.data_ptr(), nq, d_enc.data_ptr())
    d_ss = torch.zeros((nq, K), dtype=torch.float32, device=dev)
    d_si = torch.zeros((nq, K), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()  # (torch's zero fills run on its own stream)
    for qi in range(nq):
        ix.search_top_k_dev(d_enc[qi].data_ptr(), K, True, 0, d_ss[qi].data_ptr(), d_si[qi].data_ptr())
    ctx.synchronize()
    ss, si = d_ss.cpu().numpy(), d_si.cpu().numpy().view(np.uint32)
    bad = np.nonzero((bi != si).any(axis=1) | (bits(bs) != bits(ss)).any(axis=1))[0]
    assert bad.size == 0, f"batched top-100 differs from the single-query scan for queries {bad[:10]}"
    # 32 queries against the oracle (the reference's own AVX2 scorer over all rows)
    storage = ix.download()
    o = O.OracleU8(storage, count, dim, 0, False, ix.alpha, ix.offset, ix.multiplier)
    qh = queries.cpu().numpy()
    for qi in range(0, nq, nq // 32):
        enc = o.encode_query(qh[qi])
        os_, oi = o.search(enc, K, True, threads=THREADS)
        assert np.array_equal(bi[qi], oi), f"query {qi}: batched ids differ from the reference scorers"
        assert np.array_equal(bits(bs[qi]), bits(os_)), f"query {qi}: batched scores differ"
    # smallest-first over the same corpus (the other filter direction), 128 queries
    ix.search_batch_dev(queries.data_ptr(), 128, K, False, d_s.data_ptr(), d_i.data_ptr())
    ctx.synchronize()
    bs, bi = d_s[:128].cpu().numpy(), d_i[:128].cpu().numpy().view(np.uint32)
    for qi in (0, 77, 127):
        os_, oi = o.search(o.encode_query(qh[qi]), K, False, threads=THREADS)
        assert np.array_equal(bi[qi], oi) and np.array_equal(bits(bs[qi]), bits(os_))
    ix.close()
