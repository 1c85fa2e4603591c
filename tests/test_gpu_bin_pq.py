"""GPU parity for the binary and product quantizers vs the CPU oracle.
Binary: bit-exact bytes, popcounts, scores, top-k (shapes of quantization/tests/test_binary.rs:
dims {0,1,8,33,65,387} for u8 words, {1,387} for u128 words, 128 vectors of +-1).
PQ: bit-exact codes (given centroids), bit-exact LUT, bit-exact score_point in the x86 order,
scan sums within 1e-5 relative (different but fixed summation order), k-means within the
reference's own tolerance (quantization/tests/test_pq.rs: 513 x 65, |score - exact| < 0.05*dim)."""
import numpy as np
import pytest

import oracle as O
import quantization_b200 as qz

pytestmark = pytest.mark.gpu
DISTS = [qz.DistanceType.Dot, qz.DistanceType.L1, qz.DistanceType.L2]


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


# ------------------------------------------------------------------------------------------ binary
@pytest.mark.parametrize("store,dim", [(1, 0), (1, 1), (1, 8), (1, 33), (1, 65), (1, 387), (16, 1), (16, 387),
                                       (16, 1536), (1, 1536), (16, 128), (1, 100)])
@pytest.mark.parametrize("dist", DISTS)
@pytest.mark.parametrize("invert", [False, True])
def test_binary_reference_shapes(ctx, store, dim, dist, invert):
    count = 128
    data = O.signs(42, count * dim).reshape(count, dim)
    query = O.signs(43, dim)
    vp = qz.VectorParameters(dim, count, dist, invert)
    g = qz.EncodedVectorsBin.encode(data, qz.VecStorageBuilder(), vp, bits_store_type=store, ctx=ctx)
    o = O.OracleBin.encode(data, int(dist), invert, store)
    assert g.row_bytes == o.row_bytes
    assert np.array_equal(np.asarray(g.encoded_vectors), o.storage)
    gq, oq = g.encode_query(query), o.encode_query(query)
    assert np.array_equal(gq, oq)
    gs = g.score_all(gq)
    assert np.array_equal(bits(gs), bits(o.scan(oq)))
    assert np.array_equal(bits(gs), bits(o.scan(oq, use_ref=False)))
    if dist == qz.DistanceType.Dot and dim:
        exact = data @ query * (-1 if invert else 1)
        assert np.abs(gs - exact).max() <= 0.01 * dim  # test_binary.rs:41,69
    ia = np.arange(count, dtype=np.uint32)
    ib = (ia * 5 + 1) % count
    assert np.array_equal(bits(g.score_pairs(ia, ib)), bits(o.score_pairs(ia, ib)))
    assert np.array_equal(bits(g.score_points(gq, ib)), bits(o.scan(oq))[ib])
    largest = (dist == qz.DistanceType.Dot) != invert
    gs_k, gi = g.search_top_k(gq, 20, largest)
    os_k, oi = o.search(oq, 20, largest)
    assert np.array_equal(gi, oi) and np.array_equal(bits(gs_k), bits(os_k))


@pytest.mark.parametrize("dim", [256, 384, 768, 1024, 1536, 2048, 2560, 3072, 4096, 4224])
def test_binary_vector_mappings(ctx, dim):
    """u128 rows of 2..33 16-byte vectors: every lanes-per-row / vectors-per-lane variant of the scan
    (full and masked), several tiles per CTA with a ragged last tile, scores and top-k bit-exact"""
    count = 40_037
    data = O.uniform(11 + dim, count * dim, -1.0, 1.0).reshape(count, dim)
    query = O.uniform(12 + dim, dim, -1.0, 1.0)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsBin.encode(data, qz.VecStorageBuilder(), vp, ctx=ctx)
    o = O.OracleBin.encode(data, 0, False, 16)
    assert np.array_equal(np.asarray(g.encoded_vectors), o.storage)
    gq, oq = g.encode_query(query), o.encode_query(query)
    ref = o.scan(oq)
    assert np.array_equal(bits(g.score_all(gq)), bits(ref))
    for largest in (True, False):
        for k in (1, 100, 128, 300):
            gs_k, gi = g.search_top_k(gq, k, largest)
            os_k, oi = o.search(oq, k, largest)
            assert np.array_equal(gi, oi) and np.array_equal(bits(gs_k), bits(os_k))


def test_binary_massive_ties_large(ctx):
    """<= dim+1 distinct scores over 200k rows: the id tie-break is load bearing"""
    count, dim = 200_000, 256
    data = O.uniform(7, count * dim, -1.0, 1.0).reshape(count, dim)
    query = O.uniform(8, dim, -1.0, 1.0)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsBin.encode(data, qz.VecStorageBuilder(), vp, ctx=ctx)
    o = O.OracleBin.encode(data, 0, False, 16)
    assert np.array_equal(np.asarray(g.encoded_vectors), o.storage)
    gq = g.encode_query(query)
    for k, largest in ((100, True), (1000, True), (100, False)):
        gs, gi = g.search_top_k(gq, k, largest)
        os_, oi = o.search(gq, k, largest, threads=4)
        assert np.array_equal(gi, oi) and np.array_equal(bits(gs), bits(os_))


# ---------------------------------------------------------------------------------------------- pq
def pq_pair(ctx, data, dist, invert, chunk, centroids=None):
    count, dim = data.shape
    o = O.OraclePQ.encode(data, int(dist), invert, chunk, centroids=centroids)
    vp = qz.VectorParameters(dim, count, dist, invert)
    g = qz.EncodedVectorsPQ.encode(data, qz.VecStorageBuilder(), vp, chunk, 1, ctx=ctx, centroids=o.centroids)
    return g, o


@pytest.mark.parametrize("dist", DISTS)
@pytest.mark.parametrize("invert", [False, True])
def test_pq_given_centroids_513x65(ctx, dist, invert):
    """assignment, LUT and exact-order scoring are bit-exact once the centroids are fixed"""
    data = O.uniform(42, 513 * 65).reshape(513, 65)
    query = O.uniform(43, 65)
    g, o = pq_pair(ctx, data, dist, invert, 1)
    assert np.array_equal(np.asarray(g.encoded_vectors), o.codes)
    glut = g.encode_query(query).view(np.float32)
    olut = o.encode_query(query)
    assert np.array_equal(bits(glut), bits(olut))
    rows = np.arange(513, dtype=np.uint32)
    ref = o.scan(olut)  # x86 order, pq.rs:405-440
    assert np.array_equal(bits(g.score_points(glut.view(np.uint8), rows)), bits(ref))
    gs = g.score_all(glut.view(np.uint8))
    scale = np.abs(olut).reshape(-1, 256).max(1).sum()
    assert np.abs(gs - ref).max() <= 1e-5 * scale
    ia, ib = rows, (rows * 11 + 2) % 513
    assert np.array_equal(bits(g.score_pairs(ia, ib)), bits(o.score_pairs(ia, ib)))
    exact = {0: data @ query, 1: np.abs(data - query).sum(1), 2: ((data - query) ** 2).sum(1)}[int(dist)]
    assert np.abs(gs - (-exact if invert else exact)).max() < 0.05 * 65  # test_pq.rs:16-18,48


@pytest.mark.parametrize("dim,chunk", [(64, 2), (96, 4), (130, 3), (512, 2), (1536, 2), (100, 8), (48, 16)])
def test_pq_shapes_topk(ctx, dim, chunk):
    count = 3000
    data = O.uniform(dim, count * dim).reshape(count, dim)
    query = O.uniform(dim + 1, dim)
    bounds = O.pq_division(dim, chunk)
    cen = O.uniform(99, 256 * dim).reshape(256, dim)  # arbitrary fixed centroids: parity does not need training
    g, o = pq_pair(ctx, data, qz.DistanceType.Dot, False, chunk, centroids=cen)
    assert np.array_equal(np.asarray(g.encoded_vectors), o.codes)
    glut = g.encode_query(query)
    olut = o.encode_query(query)
    assert np.array_equal(bits(glut.view(np.float32)), bits(olut))
    gs = g.score_all(glut)
    ref = o.scan(olut)
    assert np.abs(gs - ref).max() <= 1e-5 * np.abs(ref).max()
    # top-k must agree with a top-k over the GPU's own scores (ordering contract), ids exact
    k_s, k_i = g.search_top_k(glut, 50, True)
    e_s, e_i = O.topk(gs, 50, True)
    assert np.array_equal(k_i, e_i) and np.array_equal(bits(k_s), bits(e_s))
    assert bounds.size - 1 == g.row_bytes


@pytest.mark.parametrize("dim,chunk,count", [(32, 2, 1_300_003), (2048, 8, 60_001), (320, 1, 9_000)])
def test_pq_streaming_scan_blocks(ctx, dim, chunk, count):
    """the streaming scan kernel at sizes where a CTA owns several row blocks (count > 148 x block), with 1, 2
    and 3 LUT slices and a ragged last tile: scores within 1e-5 relative of the oracle's x86-order sums, and the
    in-kernel top-k (k <= 128) / generic top-k (k > 128) exactly the top-k of the GPU's own scores"""
    n_chunks = (dim + chunk - 1) // chunk
    rng = np.random.default_rng(dim + count)
    codes = rng.integers(0, 256, size=(count, n_chunks), dtype=np.uint8)
    cen = O.uniform(77, 256 * dim).reshape(256, dim)
    query = O.uniform(78, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsPQ(ctx, codes.reshape(-1), vp, chunk, cen)
    glut = g.encode_query(query)
    lut = glut.view(np.float32).reshape(n_chunks, 256)
    gs = g.score_all(glut)
    ref = lut[np.arange(n_chunks)[None, :], codes].astype(np.float64).sum(1)
    assert np.abs(gs - ref).max() <= 1e-5 * np.abs(lut).max(1).sum()
    for largest in (True, False):
        for k in (1, 100, 128, 200):
            k_s, k_i = g.search_top_k(glut, k, largest)
            e_s, e_i = O.topk(gs, k, largest)
            assert np.array_equal(k_i, e_i) and np.array_equal(bits(k_s), bits(e_s))


def test_pq_scan_score_does_not_depend_on_row_position(ctx):
    """a shard of a corpus and the whole corpus must agree bit for bit (tools/peer_check.py): the scan
    sums a row's lookups in an order that is the same wherever the row sits (any offset mod 16 / 64, any CTA)"""
    dim, chunk, count = 512, 2, 40_000
    n_chunks = dim // chunk
    rng = np.random.default_rng(11)
    codes = rng.integers(0, 256, size=(count, n_chunks), dtype=np.uint8)
    cen = O.uniform(7, 256 * dim).reshape(256, dim)
    query = O.uniform(8, dim)
    whole = qz.EncodedVectorsPQ(ctx, codes.reshape(-1), qz.VectorParameters(dim, count, qz.DistanceType.Dot, False), chunk, cen)
    lut = whole.encode_query(query)
    ref = whole.score_all(lut)
    for lo, hi in ((1, count), (13, 30_001), (20_003, count), (64, 128)):
        part = qz.EncodedVectorsPQ(ctx, codes[lo:hi].reshape(-1), qz.VectorParameters(dim, hi - lo, qz.DistanceType.Dot, False),
                                   chunk, cen, id_base=lo)
        assert np.array_equal(bits(part.score_all(lut)), bits(ref[lo:hi])), (lo, hi)
        s, i = part.search_top_k(lut, 50, True)
        es, ei = O.topk(ref[lo:hi], 50, True, id_base=lo)
        assert np.array_equal(i, ei) and np.array_equal(bits(s), bits(es))
        part.close()
    whole.close()


def test_pq_kmeans_training(ctx):
    """real k-means on the device (count > 256): reference tolerance + quantization error no worse than
    the CPU oracle's k-means by more than 2%"""
    data = O.uniform(42, 2000 * 32).reshape(2000, 32)
    query = O.uniform(43, 32)
    vp = qz.VectorParameters(32, 2000, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsPQ.encode(data, qz.VecStorageBuilder(), vp, 2, 1, ctx=ctx)
    o = O.OraclePQ.encode(data, 0, False, 2)
    gs = g.score_all(g.encode_query(query))
    assert np.abs(gs - data @ query).max() < 0.05 * 32

    def qerr(cen, codes):
        rec = np.zeros_like(data)
        c = np.asarray(codes).reshape(2000, 16)
        for ch in range(16):
            rec[:, 2 * ch:2 * ch + 2] = cen[c[:, ch], 2 * ch:2 * ch + 2]
        return float(((rec - data) ** 2).sum())
    ge, oe = qerr(g.centroids, g.encoded_vectors), qerr(o.centroids, o.codes)
    assert ge <= 1.02 * oe
    # no empty-cluster reseeds on this data => the device k-means is the oracle's, bit for bit
    assert np.array_equal(bits(g.centroids), bits(o.centroids))


def test_pq_few_points_and_empty(ctx):
    """count <= 256: centroids are the points (pq.rs:290-297); count = 0 (empty_storage.rs)"""
    data = O.uniform(1, 100 * 8).reshape(100, 8)
    vp = qz.VectorParameters(8, 100, qz.DistanceType.L2, False)
    g = qz.EncodedVectorsPQ.encode(data, qz.VecStorageBuilder(), vp, 2, 1, ctx=ctx)
    assert np.array_equal(g.centroids[:100], data) and not g.centroids[100:].any()
    s = g.score_all(g.encode_query(data[3]))
    assert int(np.argmin(s)) == 3
    vp0 = qz.VectorParameters(8, 0, qz.DistanceType.Dot, False)
    g0 = qz.EncodedVectorsPQ.encode(np.zeros((0, 8), np.float32), qz.VecStorageBuilder(), vp0, 2, 1, ctx=ctx)
    assert g0.count == 0 and g0.search_top_k(g0.encode_query(np.zeros(8, np.float32)), 3, True)[1].size == 0
