"""The reference's own integration tests (quantization/tests/*.rs) restated against the CPU oracle.

The Rust-only arithmetic (encode, headers, LUT, k-means) has no golden vectors in the reference; what
pins it there are these tolerance / ordering properties, so the oracle must satisfy them with the
reference's shapes and tolerances.  Data comes from the SplitMix64 stream (StdRng is ChaCha12 and not
reproducible outside Rust).  CPU only."""
import numpy as np
import pytest

import oracle as O

DISTS = [O.DOT, O.L1, O.L2]


def exact_metric(data, q, dist, invert):
    """DistanceType::distance, encoded_vectors.rs:37-45"""
    e = {O.DOT: data @ q, O.L1: np.abs(data - q).sum(1), O.L2: ((data - q) ** 2).sum(1)}[dist]
    return -e if invert else e


# ---- u8: tests/test_simple.rs:26-351, test_avx2.rs:15-123, test_sse.rs ------------------------------
@pytest.mark.parametrize("dist", DISTS)
@pytest.mark.parametrize("invert", [False, True])
@pytest.mark.parametrize("kernel", ["scalar", "avx", "sse"])
def test_u8_score_point_tolerance(dist, invert, kernel):
    count, dim = 129, 65
    # data ranges of the reference tests: gen() in [0,1) for Dot and L2, gen_range(-1..=1) for L1
    # (test_simple.rs:24,98).  L2 over [-1,1] would expose the uncancelled actual_dim*offset^2 shift of
    # u8.rs:101 (quirk A.6.3, replicated and pinned in test_u8_l2_constant_shift below).
    lo = -1.0 if dist == O.L1 else 0.0
    data = O.uniform(42, count * dim, lo, 1.0).reshape(count, dim)
    query = O.uniform(43, dim, lo, 1.0)
    o = O.OracleU8.encode(data, dist, invert)
    assert o.actual_dim == 80 and o.stride == 84
    s = o.scan(o.encode_query(query), kernel=kernel)
    assert np.abs(s - exact_metric(data, query, dist, invert)).max() < 0.1 * dim  # test_simple.rs:29,58


@pytest.mark.parametrize("invert", [False, True])
def test_u8_score_internal_tolerance(invert):
    """test_simple.rs: score_internal Dot x invert"""
    count, dim = 129, 65
    data = O.uniform(42, count * dim).reshape(count, dim)
    o = O.OracleU8.encode(data, O.DOT, invert)
    ia = np.arange(1, count, dtype=np.uint32)
    s = o.score_pairs(np.zeros_like(ia), ia)
    ex = data[1:] @ data[0]
    assert np.abs(s - (-ex if invert else ex)).max() < 0.1 * dim


def test_u8_l2_constant_shift():
    """quirk A.6.3: L2 scores carry + actual_dim*offset^2 (u8.rs:101); ordering is unaffected"""
    count, dim = 129, 65
    data = O.uniform(42, count * dim, -1.0, 1.0).reshape(count, dim)
    query = O.uniform(43, dim, -1.0, 1.0)
    o = O.OracleU8.encode(data, O.L2, False)
    s = o.scan(o.encode_query(query))
    shift = 80.0 * float(o.offset) ** 2
    assert np.abs(s - shift - exact_metric(data, query, O.L2, False)).max() < 0.1 * dim


def test_u8_quantile_tolerance():
    """test_simple.rs:318-351: quantile just below 1.0 behaves like min/max; and the documented cut rule (A.2)"""
    count, dim = 129, 65
    data = O.uniform(42, count * dim).reshape(count, dim)
    query = O.uniform(43, dim)
    o = O.OracleU8.encode(data, O.DOT, False, quantile=1.0 - 1e-7)
    s = o.scan(o.encode_query(query))
    assert np.abs(s - data @ query).max() < 0.1 * dim
    # count < 127 or quantile >= 1.0 -> None (quantile.rs:27-29)
    assert O.quantile_interval(data[:100], 100, 100, 0.99) is None
    assert O.quantile_interval(data, count, count, 1.0) is None
    # cut uses the VECTOR count (quirk A.6.5): cut = max(1, trunc(129 * 0.01 / 2)) = 1
    lo, hi = O.quantile_interval(data, count, count, 0.99)
    srt = np.sort(data.reshape(-1))
    assert lo == srt[2] and hi == srt[-2]


def test_u8_layout_and_header():
    """A.1/A.2: row = [f32 vector_offset][codes][pad]; pad = code(0.0) for Dot, 0 for L1/L2; header formula"""
    data = O.uniform(7, 40 * 20, -1.0, 1.0).reshape(40, 20)
    for dist in DISTS:
        for inv in (False, True):
            o = O.OracleU8.encode(data, dist, inv)
            rows = o.storage.reshape(40, 36)
            codes = rows[:, 4:]
            assert codes.max() <= 127
            a, off = np.float32(o.alpha), np.float32(o.offset)
            want = np.clip(np.trunc(((data - off) / a).astype(np.float32)), 0, 127).astype(np.uint8)
            assert np.array_equal(codes[:, :20], want)
            pad = np.uint8(np.clip(np.trunc((np.float32(0.0) - off) / a), 0, 127)) if dist == O.DOT else 0
            assert (codes[:, 20:] == pad).all()
            hdr = rows[:, :4].copy().view(np.float32).reshape(-1)
            s1 = codes.astype(np.float32).sum(1, dtype=np.float32)
            s2 = (codes.astype(np.float32) ** 2).sum(1, dtype=np.float32)
            base = np.float32(np.float32(32.0) * off) * off
            want_h = {O.DOT: base + np.float32(s1 * a) * off, O.L1: np.zeros(40, np.float32),
                      O.L2: base + np.float32(s2 * a) * a}[dist].astype(np.float32)
            assert np.array_equal(hdr, -want_h if inv else want_h)


# ---- binary: tests/test_binary.rs:28-679 ---------------------------------------------------------------
@pytest.mark.parametrize("store,dim", [(1, 0), (1, 1), (1, 8), (1, 33), (1, 65), (1, 387), (16, 1), (16, 387)])
@pytest.mark.parametrize("invert", [False, True])
def test_binary_dot_exact(store, dim, invert):
    count = 128
    data = O.signs(42, count * dim).reshape(count, dim)
    query = O.signs(43, dim)
    o = O.OracleBin.encode(data, O.DOT, invert, store)
    s = o.scan(o.encode_query(query))
    assert np.abs(s - exact_metric(data, query, O.DOT, invert)).max() <= 0.01 * dim  # test_binary.rs:41
    si = o.score_pairs(np.zeros(count - 1, np.uint32), np.arange(1, count, dtype=np.uint32))
    assert np.abs(si - exact_metric(data[1:], data[0], O.DOT, invert)).max() <= 0.01 * dim


@pytest.mark.parametrize("store,dim", [(1, 1), (1, 8), (1, 33), (1, 65), (1, 387), (16, 1), (16, 387)])
@pytest.mark.parametrize("dist", [O.L1, O.L2])
@pytest.mark.parametrize("invert", [False, True])
def test_binary_l1_l2_ordering(store, dim, dist, invert):
    """test_binary.rs:243-263: argsort by binary score == argsort by the f32 metric (stable)"""
    count = 128
    data = O.signs(42, count * dim).reshape(count, dim)
    query = O.signs(43, dim)
    o = O.OracleBin.encode(data, dist, invert, store)
    s = o.scan(o.encode_query(query))
    ex = exact_metric(data, query, dist, invert)
    assert np.array_equal(np.argsort(s, kind="stable"), np.argsort(ex, kind="stable"))


def test_binary_layout():
    """A.7: bit i = (v[i] > 0.0) at byte i/8, bit i%8; u8 and u128 stores agree byte for byte where sizes agree"""
    dim = 387
    data = O.uniform(3, 10 * dim, -1.0, 1.0).reshape(10, dim)
    data[0, :5] = [0.0, -0.0, np.nan, 1e-30, -1e-30]
    a = O.OracleBin.encode(data, O.DOT, False, 16)
    b = O.OracleBin.encode(data, O.DOT, False, 1)
    assert a.row_bytes == 64 and b.row_bytes == 64 and np.array_equal(a.storage, b.storage)
    bits = np.unpackbits(a.storage.reshape(10, 64), axis=1, bitorder="little")[:, :dim]
    assert np.array_equal(bits.astype(bool), data > 0.0)
    assert [O.bin_row_bytes(d, 1) for d in (0, 1, 8, 32, 33, 64, 65, 128, 129, 387)] == [0, 1, 1, 4, 8, 8, 16, 16, 32, 64]


# ---- PQ: tests/test_pq.rs:21-270 -------------------------------------------------------------------------
@pytest.mark.parametrize("dist", DISTS)
@pytest.mark.parametrize("invert", [False, True])
def test_pq_score_point_tolerance(dist, invert):
    count, dim = 513, 65  # > 256 points so k-means really runs; chunk_size 1, 1 thread
    data = O.uniform(42, count * dim).reshape(count, dim)
    query = O.uniform(43, dim)
    o = O.OraclePQ.encode(data, dist, invert, 1)
    lut = o.encode_query(query)
    for order in ("sse", "simple"):
        s = o.scan(lut, order=order)
        assert np.abs(s - exact_metric(data, query, dist, invert)).max() < 0.05 * dim  # test_pq.rs:16-18,48
    assert np.allclose(o.scan(lut, order="sse"), o.scan(lut, order="simple"), rtol=1e-5, atol=0)


@pytest.mark.parametrize("invert", [False, True])
def test_pq_score_internal_tolerance(invert):
    count, dim = 513, 65
    data = O.uniform(42, count * dim).reshape(count, dim)
    o = O.OraclePQ.encode(data, O.DOT, invert, 1)
    s = o.score_pairs(np.zeros(count - 1, np.uint32), np.arange(1, count, dtype=np.uint32))
    assert np.abs(s - exact_metric(data[1:], data[0], O.DOT, invert)).max() < 0.05 * dim


def test_pq_assignment_rule_and_small_counts():
    """pq.rs:237-265 (strict <, lowest index wins) and pq.rs:290-297 (count <= 256: the points themselves)"""
    data = O.uniform(9, 100 * 8).reshape(100, 8)
    b = O.pq_division(8, 3)
    assert b.tolist() == [0, 3, 6, 8]  # last chunk short, pq.rs:116-121
    cen = O.OraclePQ.find_centroids(data, b)
    assert np.array_equal(cen[:100], data) and not cen[100:].any()
    cen2 = cen.copy()
    cen2[50] = cen2[10]  # duplicate centroid: the lower index must win
    codes = O.OraclePQ.assign(data, b, cen2).reshape(100, 3)
    assert (codes[10] == 10).all()  # centroid 50 is an exact duplicate of 10: strict < keeps the lower index
    d = ((data[:, None, 0:3] - cen2[None, :, 0:3]) ** 2).sum(-1)
    assert np.array_equal(codes[:, 0], d.argmin(1))


def test_kmeans_properties():
    """kmeans.rs:7-136: init = first 256 points; Lloyd lowers the quantisation error; <= max_iter iterations"""
    data = O.uniform(21, 4000 * 2).reshape(4000, 2)
    c0, it0 = O.kmeans(data, 256, 0)
    assert np.array_equal(c0, data[:256]) and it0 == 0
    c, it = O.kmeans(data, 256, 100)
    assert 1 <= it <= 100

    def err(cen):
        return ((data[:, None, :] - cen[None]) ** 2).sum(-1).min(1).sum()

    assert err(c) < 0.7 * err(c0)
    c_mt, _ = O.kmeans(data, 256, 100, max_threads=4)  # f64 merge order differs, result statistically the same
    assert abs(err(c_mt) - err(c)) < 0.05 * err(c)


# ---- empty storage / degenerate: tests/empty_storage.rs, tests/stop_condition.rs:36 ------------------------
def test_empty_and_constant_data():
    o = O.OracleU8.encode(np.zeros((0, 256), np.float32), O.DOT, False)
    assert o.count == 0 and o.storage.size == 0 and (o.alpha, o.offset, o.multiplier) == (0.0, 0.0, 0.0)
    p = O.OraclePQ.encode(np.zeros((0, 8), np.float32), O.DOT, False, 2)
    assert p.count == 0 and p.codes.size == 0
    z = O.OracleU8.encode(np.zeros((10, 8), np.float32), O.DOT, False)  # alpha = 0 -> NaN -> code 0 (A.6.8)
    assert z.alpha == 0.0 and z.multiplier == 0.0 and not z.storage.reshape(10, 20)[:, 4:].any()


# ---- top-k contract ---------------------------------------------------------------------------
def test_topk_order_and_merge():
    s = np.array([1.0, 3.0, 3.0, -0.0, 0.0, 2.0, 3.0], np.float32)
    sc, ids = O.topk(s, 4, True)
    assert ids.tolist() == [1, 2, 6, 5]                    # ties -> lower id first
    sc, ids = O.topk(s, 3, False)
    assert ids.tolist() == [3, 4, 0]                       # -0.0 < +0.0 by bit pattern, then 1.0
    sc, ids = O.topk(s, 100, True, id_base=10)
    assert ids.size == 7 and ids[0] == 11
    big = O.uniform(5, 10000)
    a = O.topk(big[:5000], 50, True, 0)
    b = O.topk(big[5000:], 50, True, 5000)
    m = O.merge_topk(np.concatenate([a[0], b[0]]), np.concatenate([a[1], b[1]]), 50, True)
    w = O.topk(big, 50, True)
    assert np.array_equal(m[1], w[1]) and np.array_equal(m[0], w[0])
