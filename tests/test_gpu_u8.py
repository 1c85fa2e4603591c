"""GPU parity, u8 scalar quantizer: the CUDA path behind the C-ABI vs the CPU oracle (which runs the
reference's own AVX2 integer kernels where oracle/_ref exists).  Bit-exact everywhere: storage
bytes, Metadata scalars, encoded queries, scores, top-k ids.  Shapes follow the reference's tests
(129 x 65 -> actual_dim 80, quantization/tests/test_simple.rs, test_avx2.rs) plus larger ones."""
import numpy as np
import pytest

import oracle as O
import quantization_b200 as qz

pytestmark = pytest.mark.gpu

DISTS = [qz.DistanceType.Dot, qz.DistanceType.L1, qz.DistanceType.L2]


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def make(count, dim, lo=0.0, hi=1.0, seed=42):
    data = O.uniform(seed, count * dim, lo, hi).reshape(count, dim)
    query = O.uniform(seed + 1, dim, lo, hi)
    return data, query


def encode_both(ctx, data, dist, invert, quantile=None):
    count, dim = data.shape
    vp = qz.VectorParameters(dim, count, dist, invert)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, quantile, lambda: False, ctx=ctx)
    o = O.OracleU8.encode(data, int(dist), invert, quantile)
    return g, o


@pytest.mark.parametrize("dist", DISTS)
@pytest.mark.parametrize("invert", [False, True])
def test_reference_shape_129x65(ctx, dist, invert):
    """test_simple.rs / test_avx2.rs shapes; reference tolerance |score - exact| < 0.1*dim plus bit parity"""
    data, query = make(129, 65)
    g, o = encode_both(ctx, data, dist, invert)
    assert g.actual_dim == 80
    assert np.array_equal(np.asarray(g.encoded_vectors), o.storage)
    assert bits([g.alpha, g.offset, g.multiplier]).tolist() == bits([o.alpha, o.offset, o.multiplier]).tolist()
    gq, oq = g.encode_query(query), o.encode_query(query)
    assert np.array_equal(gq, oq)
    gs = g.score_all(gq)
    assert np.array_equal(bits(gs), bits(o.scan(oq)))             # vs reference AVX2 kernels
    assert np.array_equal(bits(gs), bits(o.scan(oq, kernel="scalar")))  # vs scalar restatement
    exact = {0: data @ query, 1: np.abs(data - query).sum(1), 2: ((data - query) ** 2).sum(1)}[int(dist)]
    if invert:
        exact = -exact
    assert np.abs(gs - exact).max() < 0.1 * 65
    # score_point one id at a time, as the trait is called
    for i in (0, 1, 64, 128):
        assert np.float32(g.score_point(gq, i)).view(np.uint32) == np.float32(o.score_point(oq, i)).view(np.uint32)


@pytest.mark.parametrize("dist", DISTS)
@pytest.mark.parametrize("invert", [False, True])
def test_score_internal(ctx, dist, invert):
    data, _ = make(129, 65, -1.0, 1.0)
    g, o = encode_both(ctx, data, dist, invert)
    ia = np.arange(129, dtype=np.uint32)
    ib = (ia * 7 + 3) % 129
    assert np.array_equal(bits(g.score_pairs(ia, ib)), bits(o.score_pairs(ia, ib)))
    assert np.float32(g.score_internal(5, 9)).view(np.uint32) == np.float32(o.score_internal(5, 9)).view(np.uint32)


@pytest.mark.parametrize("dim", [1, 15, 16, 17, 64, 100, 128, 384, 768, 1000, 1536, 2048, 4096])
def test_dims(ctx, dim):
    count = 777
    data, query = make(count, dim, -1.0, 1.0, seed=dim)
    for dist in DISTS:
        g, o = encode_both(ctx, data, dist, False)
        assert np.array_equal(np.asarray(g.encoded_vectors), o.storage)
        gq = g.encode_query(query)
        assert np.array_equal(gq, o.encode_query(query))
        assert np.array_equal(bits(g.score_all(gq)), bits(o.scan(gq, kernel="scalar")))
        largest = dist == qz.DistanceType.Dot
        gs, gi = g.search_top_k(gq, 10, largest)
        os_, oi = o.search(gq, 10, largest)
        assert np.array_equal(gi, oi) and np.array_equal(bits(gs), bits(os_))


@pytest.mark.parametrize("k", [1, 2, 30, 100, 128, 129, 500, 1024])
@pytest.mark.parametrize("largest", [True, False])
def test_topk_sizes(ctx, k, largest):
    data, query = make(20000, 128)
    g, o = encode_both(ctx, data, qz.DistanceType.Dot, False)
    gq = g.encode_query(query)
    gs, gi = g.search_top_k(gq, k, largest)
    os_, oi = o.search(gq, k, largest, threads=4)
    assert np.array_equal(gi, oi)
    assert np.array_equal(bits(gs), bits(os_))


def test_topk_more_than_count_and_tiny(ctx):
    data, query = make(7, 33)
    g, o = encode_both(ctx, data, qz.DistanceType.L2, False)
    gq = g.encode_query(query)
    gs, gi = g.search_top_k(gq, 100, False)
    os_, oi = o.search(gq, 100, False)
    assert gi.size == 7 and np.array_equal(gi, oi) and np.array_equal(bits(gs), bits(os_))


def test_topk_ties_prefer_lower_id(ctx):
    """identical rows => identical scores; the contract resolves ties by lower id"""
    row = O.uniform(5, 96)
    data = np.tile(row, (5000, 1))
    data[1234] += 0.25  # one distinct row so that min != max
    g, o = encode_both(ctx, data, qz.DistanceType.Dot, False)
    gq = g.encode_query(row)
    gs, gi = g.search_top_k(gq, 50, True)
    os_, oi = o.search(gq, 50, True)
    assert np.array_equal(gi, oi)
    assert sorted(gi[1:].tolist()) == gi[1:].tolist()


def test_adversarial_order_sorted_scores(ctx):
    """scores increasing with id defeats any sampled threshold: every row is a new best"""
    count, dim = 30000, 32
    base = np.linspace(0.0, 1.0, count, dtype=np.float32)[:, None]
    data = np.repeat(base, dim, axis=1)
    g, o = encode_both(ctx, data, qz.DistanceType.Dot, False)
    gq = g.encode_query(np.ones(dim, np.float32))
    for largest in (True, False):
        gs, gi = g.search_top_k(gq, 100, largest)
        os_, oi = o.search(gq, 100, largest)
        assert np.array_equal(gi, oi) and np.array_equal(bits(gs), bits(os_))


def test_quantile(ctx):
    """quantile.rs:21-71 incl. the vector-count quirk; count <= 100000 => sample = all => deterministic"""
    data, query = make(1000, 65, -1.0, 1.0)
    for q in (0.99, 0.9, 0.5):
        g, o = encode_both(ctx, data, qz.DistanceType.Dot, False, quantile=q)
        assert bits([g.alpha, g.offset]).tolist() == bits([o.alpha, o.offset]).tolist()
        assert np.array_equal(np.asarray(g.encoded_vectors), o.storage)
    # quantile >= 1.0 and count < 127 fall back to min/max (quantile.rs:27-29)
    g, o = encode_both(ctx, data, qz.DistanceType.Dot, False, quantile=1.0)
    g2, _ = encode_both(ctx, data, qz.DistanceType.Dot, False)
    assert g.alpha == g2.alpha and g.offset == g2.offset == np.float32(o.offset)
    g, o = encode_both(ctx, data[:100], qz.DistanceType.Dot, False, quantile=0.9)
    assert bits([g.alpha, g.offset]).tolist() == bits([o.alpha, o.offset]).tolist()


def test_quantile_explicit_sample(ctx):
    data, _ = make(3000, 40, -2.0, 2.0)
    ids = np.sort(np.random.default_rng(0).choice(3000, 500, replace=False)).astype(np.uint64)
    vp = qz.VectorParameters(40, 3000, qz.DistanceType.L2, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, 0.95, ctx=ctx, sample_ids=ids)
    # oracle: slice_size is min(count, 100000) in the reference; an explicit sample replaces the permutation
    iv = O.quantile_interval(data[ids], 3000, ids.size, 0.95)
    a, off = (iv[1] - iv[0]), iv[0]
    assert np.float32(g.offset) == np.float32(off)
    assert np.float32(g.alpha) == np.float32(np.float32(np.float32(iv[1]) - np.float32(iv[0])) / np.float32(127.0))


def test_empty_and_degenerate(ctx):
    """empty_storage.rs (count = 0) and stop_condition.rs's all-zero data (alpha = 0 -> NaN -> code 0)"""
    vp = qz.VectorParameters(256, 0, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsU8.encode(np.zeros((0, 256), np.float32), qz.VecStorageBuilder(), vp, None, ctx=ctx)
    assert (g.alpha, g.offset, g.multiplier) == (0.0, 0.0, 0.0) and g.count == 0
    assert g.search_top_k(np.zeros(260, np.uint8), 5, True)[1].size == 0
    data = np.zeros((300, 8), np.float32)
    g, o = encode_both(ctx, data, qz.DistanceType.Dot, False)
    assert np.array_equal(np.asarray(g.encoded_vectors), o.storage) and g.multiplier == 0.0
    gq = g.encode_query(np.zeros(8, np.float32))
    assert np.array_equal(bits(g.score_all(gq)), bits(o.scan(gq)))


def test_stop_condition(ctx):
    """stop_condition.rs: a stop request surfaces as EncodingError::Stopped"""
    data, _ = make(1000, 8)
    vp = qz.VectorParameters(8, 1000, qz.DistanceType.Dot, False)
    with pytest.raises(qz.EncodingError) as e:
        qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, lambda: True, ctx=ctx)
    assert e.value == qz.EncodingError.Stopped


def test_argument_errors(ctx):
    data, _ = make(10, 8)
    with pytest.raises(qz.EncodingError) as e:
        qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), qz.VectorParameters(8, 11, qz.DistanceType.Dot, False), ctx=ctx)
    assert e.value.kind == "ArgumentsError"
    g, _ = encode_both(ctx, data, qz.DistanceType.Dot, False)
    with pytest.raises(qz.EncodingError):
        g.search_top_k(g.encode_query(data[0]), 0, True)
    with pytest.raises(qz.EncodingError):
        g.score_point(g.encode_query(data[0]), 10)


def test_save_load_roundtrip(ctx, tmp_path):
    data, query = make(200, 24)
    g, _ = encode_both(ctx, data, qz.DistanceType.L2, True)
    g.save(str(tmp_path / "d" / "data.bin"), str(tmp_path / "m" / "meta.json"))
    vp = qz.VectorParameters(24, 200, qz.DistanceType.L2, True)
    h = qz.EncodedVectorsU8.load(str(tmp_path / "d" / "data.bin"), str(tmp_path / "m" / "meta.json"), vp, ctx=ctx)
    q = g.encode_query(query)
    assert np.array_equal(q, h.encode_query(query))
    assert np.array_equal(bits(g.score_all(q)), bits(h.score_all(q)))
    import json
    meta = json.load(open(tmp_path / "m" / "meta.json"))
    assert set(meta) == {"actual_dim", "alpha", "offset", "multiplier", "vector_parameters"}  # u8.rs:24-31
    assert meta["vector_parameters"] == {"dim": 24, "count": 200, "distance_type": "L2", "invert": True}


def test_sharded_ids(ctx):
    """two shards with id_base reproduce the single-index top-k after a merge"""
    data, query = make(9000, 64)
    g, o = encode_both(ctx, data, qz.DistanceType.Dot, False)
    row = g.get_quantized_vector_size(g.vector_parameters)
    st = np.asarray(g.encoded_vectors)
    parts = []
    for a, b in ((0, 4000), (4000, 9000)):
        vp = qz.VectorParameters(64, b - a, qz.DistanceType.Dot, False)
        parts.append(qz.EncodedVectorsU8(ctx, st[a * row:b * row], vp, g.alpha, g.offset, g.multiplier, id_base=a))
    gq = g.encode_query(query)
    res = [p.search_top_k(gq, 64, True) for p in parts]
    ms, mi = O.merge_topk(np.concatenate([r[0] for r in res]), np.concatenate([r[1] for r in res]), 64, True)
    os_, oi = o.search(gq, 64, True)
    assert np.array_equal(mi, oi) and np.array_equal(bits(ms), bits(os_))


def test_search_fused_call(ctx):
    """qz_search_topk == encode_query + scan_topk"""
    data, query = make(3000, 200)
    for dist in DISTS:
        g, o = encode_both(ctx, data, dist, False)
        s, i = g.search(query, 50, dist == qz.DistanceType.Dot)
        ws, wi = o.search(o.encode_query(query), 50, dist == qz.DistanceType.Dot)
        assert np.array_equal(i, wi) and np.array_equal(bits(s), bits(ws))
