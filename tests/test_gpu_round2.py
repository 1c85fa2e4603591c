"""Boundary and edge behaviour the reference's own tests pin and round 1 left uncovered:

  * stop flag flipped INTO a running encode, for u8, binary, PQ k-means and PQ assignment
    (quantization/tests/stop_condition.rs:18-86: a timer thread sets the AtomicBool 100 / 300 ms into a 1M x 8 encode
    and the call must come back with EncodingError::Stopped)
  * save -> load -> score round trips for PQ and binary with the serde JSON shape of
    encoded_vectors_pq.rs:39-44, 498-523 and encoded_vectors_binary.rs:260-286
  * id-list scoring on device pointers (qz_score_points_dev / qz_score_pairs_dev) against the oracle
  * one context used from two threads; two contexts on two devices in one process
  * min / max of data whose extreme is a zero of mixed sign (quantile.rs:5-19 keeps the first one it sees)
  * the batched search on an empty index and after qz_u8_index_set_params
"""
import ctypes as C
import json
import threading
import time

import numpy as np
import pytest

import oracle as O
import quantization_b200 as qz

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


# ------------------------------------------------------------------------------------------ stop flag
def _timer_stop(delay_s):
    """the reference test's timer thread: false until `delay_s` after the first poll, then true"""
    state = {"t0": None, "fired_at": None}

    def cond():
        now = time.perf_counter()
        if state["t0"] is None:
            state["t0"] = now
        if now - state["t0"] >= delay_s:
            if state["fired_at"] is None:
                state["fired_at"] = now
            return True
        return False

    return cond, state


def _expect_stopped(fn):
    with pytest.raises(qz.EncodingError) as e:
        fn()
    assert e.value == qz.EncodingError.Stopped


def test_stop_flag_mid_encode_u8_and_binary(ctx):
    """2M x 128 f32 (1 GB) from pageable host memory: the encode streams it through the device in 256 MB pieces and
    polls the flag between pieces (u8.rs:74-76, binary.rs:178-180); the flag flips 10 ms in"""
    count, dim = 2_000_000, 128
    data = np.random.default_rng(1).random((count, dim), dtype=np.float32)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    for encode in (lambda c: qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, c, ctx=ctx),
                   lambda c: qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, 0.99, c, ctx=ctx),
                   lambda c: qz.EncodedVectorsBin.encode(data, qz.VecStorageBuilder(), vp, c, ctx=ctx)):
        cond, state = _timer_stop(0.010)
        t0 = time.perf_counter()
        _expect_stopped(lambda: encode(cond))
        assert state["fired_at"] is not None and time.perf_counter() - t0 < 5.0
    # and the same call runs to completion when nobody stops it (the flag is really what ended the others)
    g = qz.EncodedVectorsBin.encode(data[:50_000], qz.VecStorageBuilder(), qz.VectorParameters(dim, 50_000, qz.DistanceType.Dot, False),
                                    lambda: False, ctx=ctx)
    assert g.count == 50_000
    g.close()


def test_stop_flag_mid_pq_train_and_assign(ctx):
    """k-means polls between Lloyd iterations (kmeans.rs:29-31), the assignment between row pieces (pq.rs:198-200)"""
    L = ctx._L
    count, dim, chunk = 20_000, 512, 2
    data = np.random.default_rng(2).random((count, dim), dtype=np.float32)
    cen = np.zeros((256, dim), np.float32)
    # accuracy 0 never converges early: 100 iterations of 256 chunk problems, the flag flips 5 ms in
    flag = C.c_int32(0)
    timer = threading.Timer(0.005, lambda: setattr(flag, "value", 1))
    timer.start()
    st = L.qz_pq_train(ctx.handle, C.c_void_p(data.ctypes.data), count, dim, chunk, None, 0, 7, 100, C.c_float(0.0),
                       C.c_void_p(C.addressof(flag)), C.c_void_p(cen.ctypes.data), None)
    timer.join()
    assert st == 3, f"qz_pq_train returned {st}, expected QZ_STOPPED"  # EncodingError::Stopped
    # assignment: 1.5M x 128 (768 MB pageable) against 256 centroids
    count, dim = 1_500_000, 128
    data = np.random.default_rng(3).random((count, dim), dtype=np.float32)
    cen = np.ascontiguousarray(data[:256])
    codes = np.zeros(count * (dim // chunk), np.uint8)
    flag = C.c_int32(0)
    timer = threading.Timer(0.010, lambda: setattr(flag, "value", 1))
    timer.start()
    st = L.qz_pq_assign(ctx.handle, C.c_void_p(data.ctypes.data), count, dim, chunk, C.c_void_p(cen.ctypes.data),
                        C.c_void_p(C.addressof(flag)), C.c_void_p(codes.ctypes.data))
    timer.join()
    assert st == 3, f"qz_pq_assign returned {st}, expected QZ_STOPPED"
    # through the trait mirror: Stopped surfaces as the reference's error value
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    cond, _ = _timer_stop(0.010)
    _expect_stopped(lambda: qz.EncodedVectorsPQ.encode(data, qz.VecStorageBuilder(), vp, chunk, 1, cond, ctx=ctx, centroids=cen))


# ------------------------------------------------------------------------------------------ save / load
def test_pq_save_load_roundtrip(ctx, tmp_path):
    count, dim, chunk = 600, 18, 4  # 5 chunks, the last one short (pq.rs:116-121)
    data = O.uniform(5, count * dim, -1.0, 1.0).reshape(count, dim)
    query = O.uniform(6, dim, -1.0, 1.0)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.L2, False)
    g = qz.EncodedVectorsPQ.encode(data, qz.VecStorageBuilder(), vp, chunk, 1, ctx=ctx)
    dp, mp = str(tmp_path / "pq" / "data.bin"), str(tmp_path / "pq" / "meta.json")
    g.save(dp, mp)
    meta = json.load(open(mp))
    assert set(meta) == {"centroids", "vector_division", "vector_parameters"}  # Metadata, pq.rs:39-44
    assert meta["vector_division"] == [{"start": a, "end": min(a + chunk, dim)} for a in range(0, dim, chunk)]  # Range<usize>
    assert len(meta["centroids"]) == 256 and all(len(c) == dim for c in meta["centroids"])
    assert meta["vector_parameters"] == {"dim": dim, "count": count, "distance_type": "L2", "invert": False}
    assert np.array_equal(np.asarray(meta["centroids"], np.float32), g.centroids)  # shortest round-trip floats
    h = qz.EncodedVectorsPQ.load(dp, mp, vp, ctx=ctx)
    lut = g.encode_query(query)
    assert np.array_equal(lut, h.encode_query(query))
    assert np.array_equal(bits(g.score_all(lut)), bits(h.score_all(lut)))
    ia, ib = np.arange(50, dtype=np.uint32), np.arange(50, 100, dtype=np.uint32)
    assert np.array_equal(bits(g.score_pairs(ia, ib)), bits(h.score_pairs(ia, ib)))
    # the oracle scores the loaded bytes the same way (x86 summation order, bit for bit)
    o = O.OraclePQ(np.fromfile(dp, np.uint8), count, dim, O.pq_division(dim, chunk), g.centroids, 2, False)
    assert np.array_equal(bits(h.score_points(lut, ia)), bits(o.scan(lut.view(np.float32))[:50]))
    # wrong size / foreign metadata are rejected like encoded_storage.rs:40-50
    with pytest.raises(OSError):
        qz.EncodedVectorsPQ.load(dp, mp, qz.VectorParameters(dim, count + 1, qz.DistanceType.L2, False), ctx=ctx)


@pytest.mark.parametrize("store", [1, 16])
def test_binary_save_load_roundtrip(ctx, tmp_path, store):
    count, dim = 300, 387
    data = O.signs(8, count * dim).reshape(count, dim)
    query = O.signs(9, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, True)
    g = qz.EncodedVectorsBin.encode(data, qz.VecStorageBuilder(), vp, bits_store_type=store, ctx=ctx)
    dp, mp = str(tmp_path / "bin" / "data.bin"), str(tmp_path / "bin" / "meta.json")
    g.save(dp, mp)
    meta = json.load(open(mp))
    assert meta == {"vector_parameters": {"dim": dim, "count": count, "distance_type": "Dot", "invert": True}}  # binary.rs:21-24
    h = qz.EncodedVectorsBin.load(dp, mp, vp, bits_store_type=store, ctx=ctx)
    q = g.encode_query(query)
    assert np.array_equal(q, h.encode_query(query))
    o = O.OracleBin(np.fromfile(dp, np.uint8), count, dim, store, 0, True)
    assert np.array_equal(bits(h.score_all(q)), bits(o.scan(o.encode_query(query))))
    with pytest.raises(OSError):
        qz.EncodedVectorsBin.load(dp, mp, qz.VectorParameters(dim + 200, count, qz.DistanceType.Dot, True), bits_store_type=store, ctx=ctx)


def test_save_of_a_device_encoded_index(ctx, tmp_path):
    """an index encoded on the device (no host copy of the storage) saves the bytes a host encode would"""
    torch = pytest.importorskip("torch")
    count, dim = 5000, 96
    data = O.uniform(15, count * dim).reshape(count, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    d = torch.from_numpy(data).to("cuda:0")
    torch.cuda.synchronize()
    g = qz.EncodedVectorsU8.encode_dev(d.data_ptr(), vp, None, ctx=ctx)
    dp, mp = str(tmp_path / "u8d" / "data.bin"), str(tmp_path / "u8d" / "meta.json")
    g.save(dp, mp)
    o = O.OracleU8.encode(data, 0, False)
    assert np.array_equal(np.fromfile(dp, np.uint8), o.storage)
    h = qz.EncodedVectorsU8.load(dp, mp, vp, ctx=ctx)
    assert (h.alpha, h.offset, h.multiplier) == (o.alpha, o.offset, o.multiplier)
    meta = json.load(open(mp))
    meta["vector_parameters"]["count"] = count * 2  # a stale meta file must not make the index read past the bytes
    json.dump(meta, open(mp, "w"))
    with pytest.raises(OSError):
        qz.EncodedVectorsU8.load(dp, mp, vp, ctx=ctx)


# ------------------------------------------------------------------------------------------ id lists on the device
def test_score_points_and_pairs_dev(ctx):
    """qz_score_points_dev / qz_score_pairs_dev (graph-search callers keep candidate lists on the device:
    demos/benches/encode.rs:91-102, binary.rs:62-72, pq.rs:41-46) against the oracle's score_point / score_internal"""
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    count, dim, n = 20_000, 200, 7001
    rng = np.random.default_rng(4)
    ia = rng.integers(0, count, n).astype(np.uint32)
    ib = rng.integers(0, count, n).astype(np.uint32)
    d_a = torch.from_numpy(ia.view(np.int32)).to(dev)
    d_b = torch.from_numpy(ib.view(np.int32)).to(dev)
    d_out = torch.zeros(n, dtype=torch.float32, device=dev)
    data = O.uniform(31, count * dim, -1.0, 1.0).reshape(count, dim)
    query = O.uniform(32, dim, -1.0, 1.0)

    def run_points(ix, blob):
        d_q = torch.from_numpy(np.frombuffer(bytes(blob), np.uint8).copy()).to(dev)
        torch.cuda.synchronize()
        ix.score_points_dev(d_q.data_ptr(), d_b.data_ptr(), n, d_out.data_ptr())
        ctx.synchronize()
        return d_out.cpu().numpy().copy()

    def run_pairs(ix):
        torch.cuda.synchronize()
        ix.score_pairs_dev(d_a.data_ptr(), d_b.data_ptr(), n, d_out.data_ptr())
        ctx.synchronize()
        return d_out.cpu().numpy().copy()

    for dist in (qz.DistanceType.Dot, qz.DistanceType.L1):
        vp = qz.VectorParameters(dim, count, dist, False)
        g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, ctx=ctx)
        o = O.OracleU8.encode(data, int(dist), False)
        enc = g.encode_query(query)
        assert np.array_equal(bits(run_points(g, enc)), bits(o.scan(o.encode_query(query))[ib]))
        assert np.array_equal(bits(run_pairs(g)), bits(o.score_pairs(ia, ib)))
        g.close()
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    b = qz.EncodedVectorsBin.encode(data, qz.VecStorageBuilder(), vp, ctx=ctx)
    ob = O.OracleBin.encode(data, 0, False, 16)
    assert np.array_equal(bits(run_points(b, b.encode_query(query))), bits(ob.scan(ob.encode_query(query))[ib]))
    assert np.array_equal(bits(run_pairs(b)), bits(ob.score_pairs(ia, ib)))
    b.close()
    cen = O.uniform(33, 256 * dim, -1.0, 1.0).reshape(256, dim)
    p = qz.EncodedVectorsPQ.encode(data, qz.VecStorageBuilder(), vp, 2, 1, ctx=ctx, centroids=cen)
    op = O.OraclePQ.encode(data, 0, False, 2, centroids=cen)
    lut = p.encode_query(query)
    assert np.array_equal(bits(run_points(p, lut)), bits(op.scan(lut.view(np.float32))[ib]))  # exact x86 order
    assert np.array_equal(bits(run_pairs(p)), bits(op.score_pairs(ia, ib)))
    p.close()


# ------------------------------------------------------------------------------------------ threads and devices
def test_two_threads_share_one_context(ctx):
    """the scoring entry points take &self in the reference and are called from any thread: two threads
    hammer one context (two indexes, different codecs) through the host-facing calls and every answer must be right"""
    count, dim, k = 30_011, 256, 40
    data = O.uniform(51, count * dim, -1.0, 1.0).reshape(count, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, ctx=ctx)
    b = qz.EncodedVectorsBin.encode(data, qz.VecStorageBuilder(), vp, ctx=ctx)
    o, ob = O.OracleU8.encode(data, 0, False), O.OracleBin.encode(data, 0, False, 16)
    queries = O.uniform(52, 24 * dim, -1.0, 1.0).reshape(24, dim)
    want_u8 = [o.search(o.encode_query(q), k, True) for q in queries]
    want_bin = [ob.search(ob.encode_query(q), k, True) for q in queries]
    errors = []

    def worker(ix, want, reps):
        try:
            for r in range(reps):
                for q, (ws, wi) in zip(queries, want):
                    s, i = ix.search(q, k, True)
                    if not (np.array_equal(i, wi) and np.array_equal(bits(s), bits(ws))):
                        errors.append((type(ix).__name__, r))
                        return
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    ts = [threading.Thread(target=worker, args=(g, want_u8, 6)), threading.Thread(target=worker, args=(b, want_bin, 6)),
          threading.Thread(target=worker, args=(g, want_u8, 6))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors
    g.close()
    b.close()


def test_search_reads_the_query_of_this_call(ctx):
    """qz_search_topk launches the scan as the programmatic dependent of the query-encode kernel: the scan's CTAs are
    resident, streaming tiles, before the encoded query exists, and must not score a row until it does.  A stream of
    calls with a DIFFERENT query each (same staging buffers every time) must give, call by call, what encode_query +
    search_top_k give for that query (two separate synchronised calls: no overlap) -- a scan that picked up the
    previous call's encoded query would return the previous call's answer.  Sized so that every SM has a CTA and the
    tile counter is on (the C2 kernel variant: 1536-d)."""
    count, dim, k = 300_000, 1536, 100
    rng = np.random.default_rng(77)
    data = rng.random((count, dim), dtype=np.float32)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, ctx=ctx)
    queries = rng.random((12, dim), dtype=np.float32)
    queries[5] = -queries[4]           # an answer far away from its predecessor's
    want = [g.search_top_k(g.encode_query(q), k, True) for q in queries]
    assert len({tuple(w[1][:5]) for w in want}) > 6  # the queries really have different answers
    for rep in range(5):
        for q, (ws, wi) in zip(queries, want):
            s, i = g.search(q, k, True)
            assert np.array_equal(i, wi) and np.array_equal(bits(s), bits(ws)), rep
    g.close()


def test_two_contexts_on_two_devices_in_one_process():
    """qz_ctx_create(device, ...) twice in one process (the natural single-process multi-GPU host): every kernel's
    shared-memory opt-in is per device, so the second device must get it too (round-1 defect: process-wide statics)"""
    torch = pytest.importorskip("torch")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    count, dim, k = 50_000, 1536, 100
    data = O.uniform(61, count * dim).reshape(count, dim)
    query = O.uniform(62, dim)
    o = O.OracleU8.encode(data, 0, False)
    ws, wi = o.search(o.encode_query(query), k, True, threads=8)
    ob = O.OracleBin.encode(data - 0.5, 0, False, 16)
    bws, bwi = ob.search(ob.encode_query(query - 0.5), k, True, threads=8)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    for device in (1, 0, 1):
        c = qz.Context(device)
        g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, ctx=c)
        s, i = g.search(query, k, True)
        assert np.array_equal(i, wi) and np.array_equal(bits(s), bits(ws)), f"u8 scan on device {device}"
        bs, bi = g.search_batch(np.stack([query, query]), k, True)
        assert np.array_equal(bi[1], wi) and np.array_equal(bits(bs[1]), bits(ws)), f"batched search on device {device}"
        b = qz.EncodedVectorsBin.encode(data - 0.5, qz.VecStorageBuilder(), vp, ctx=c)
        s, i = b.search(query - 0.5, k, True)
        assert np.array_equal(i, bwi) and np.array_equal(bits(s), bits(bws)), f"binary scan on device {device}"
        cen = data[:256].copy()
        p = qz.EncodedVectorsPQ.encode(data[:20_000], qz.VecStorageBuilder(), qz.VectorParameters(dim, 20_000, qz.DistanceType.Dot, False),
                                       2, 1, ctx=c, centroids=cen)
        lut = p.encode_query(query)
        op = O.OraclePQ.encode(data[:20_000], 0, False, 2, centroids=cen)
        assert np.allclose(p.score_all(lut), op.scan(lut.view(np.float32)), rtol=1e-5, atol=0), f"pq scan on device {device}"
        for ix in (g, b, p):
            ix.close()
        c.close()


# ------------------------------------------------------------------------------------------ edge cases
def test_minmax_with_zeros_of_mixed_sign(ctx):
    """find_min_max_from_iter (quantile.rs:5-19) folds with strict < / >, so of two zeros of different sign it keeps
    the one it saw first — and the sign of a zero offset reaches the per-vector headers (u8.rs:96-111) and the query
    offset.  The device fold finds the first zero's position and takes its sign: Metadata, storage bytes and
    encoded queries stay bit-identical in both orders, for a zero minimum and for a zero maximum."""
    rng = np.random.default_rng(5)
    for sign in (1.0, -1.0):  # zero minimum (non-negative data) / zero maximum (non-positive data)
        for first, second in ((0.0, -0.0), (-0.0, 0.0)):
            data = (rng.random((64, 24), dtype=np.float32) * np.float32(sign)).astype(np.float32)
            data[3, 5], data[40, 7] = first, second  # the extreme is a zero, seen in this order
            vp = qz.VectorParameters(24, 64, qz.DistanceType.Dot, False)
            g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, ctx=ctx)
            o = O.OracleU8.encode(data, 0, False)
            if sign > 0:
                assert np.signbit(np.float32(o.offset)) == np.signbit(np.float32(first))  # the oracle follows the reference
            assert bits([g.alpha, g.offset, g.multiplier]).tolist() == bits([o.alpha, o.offset, o.multiplier]).tolist()
            assert np.array_equal(np.asarray(g.encoded_vectors), o.storage)
            q = rng.random(24, dtype=np.float32)
            gq, oq = g.encode_query(q), o.encode_query(q)
            assert np.array_equal(gq, oq)
            assert np.array_equal(bits(g.score_all(gq)), bits(o.scan(oq)))
            # the device-resident encode applies the same rule
            torch = pytest.importorskip("torch")
            d = torch.from_numpy(data).to("cuda:0")
            torch.cuda.synchronize()
            gd = qz.EncodedVectorsU8.encode_dev(d.data_ptr(), vp, None, ctx=ctx)
            assert bits([gd.offset]).tolist() == bits([o.offset]).tolist() and np.array_equal(gd.download(), o.storage)
            g.close()
            gd.close()


def test_batch_search_empty_index_and_param_update(ctx):
    torch = pytest.importorskip("torch")
    dim, k = 64, 10
    empty = qz.EncodedVectorsU8(ctx, np.zeros(0, np.uint8), qz.VectorParameters(dim, 0, qz.DistanceType.Dot, False), 0.1, 0.0, 0.01)
    s, i = empty.search_batch(O.uniform(1, 3 * dim).reshape(3, dim), k, True)
    assert np.isnan(s).all() and (i == 0xFFFFFFFF).all()
    empty.close()
    # the device-side mirror of the batched path is rebuilt when the Metadata scalars change under it
    count = 4000
    data = O.uniform(71, count * dim, -1.0, 1.0).reshape(count, dim)
    queries = O.uniform(72, 5 * dim, -1.0, 1.0).reshape(5, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, ctx=ctx)
    o = O.OracleU8.encode(data, 0, False)
    g.search_batch(queries, k, True)
    ctx.check(ctx._L.qz_u8_index_set_params(g.handle, float(g.alpha), float(g.offset), float(-g.multiplier)))
    o2 = O.OracleU8(o.storage, count, dim, 0, False, o.alpha, o.offset, -o.multiplier)
    s, i = g.search_batch(queries, k, True)
    for qi in range(5):
        enc = o.encode_query(queries[qi])
        ws, wi = o2.search(enc, k, True)
        assert np.array_equal(i[qi], wi) and np.array_equal(bits(s[qi]), bits(ws))
    g.close()


# ------------------------------------------------------------------------------------------ chained scans
def test_chained_scans_give_the_results_of_separate_scans(ctx):
    """consecutive *_dev scans on a context that owns its stream are launched as programmatic dependents (the
    streaming of scan i+1 overlaps the final merge of scan i, tile counters alternate): 300 back-to-back scans over
    two indexes of different size (one CTA per SM / a few CTAs) with 8 different queries and k values, none
    synchronised, must equal the same scans made one at a time through the host-buffer call"""
    import torch
    rng = np.random.default_rng(21)
    big_n, small_n, dim = 400_000, 3_000, 256
    big = rng.random((big_n, dim), dtype=np.float32)
    g_big = qz.EncodedVectorsU8.encode(big, qz.VecStorageBuilder(), qz.VectorParameters(dim, big_n, qz.DistanceType.Dot, False), ctx=ctx)
    g_small = qz.EncodedVectorsU8.encode(big[:small_n], qz.VecStorageBuilder(),
                                         qz.VectorParameters(dim, small_n, qz.DistanceType.Dot, False), ctx=ctx)
    b_big = qz.EncodedVectorsBin.encode(big - 0.5, qz.VecStorageBuilder(), qz.VectorParameters(dim, big_n, qz.DistanceType.Dot, False), ctx=ctx)
    queries = [rng.random(dim, dtype=np.float32) for _ in range(8)]
    plan = []  # (index, encoded query on the device, k, largest, expected)
    for n in range(300):
        ix = (g_big, g_small, b_big)[n % 3]
        q = queries[n % 8] - (0.5 if ix is b_big else 0.0)
        k, largest = (100, 17, 128, 1)[n % 4], n % 5 != 0
        plan.append((ix, q, k, largest))
    expected = {}
    enc_dev = {}
    for ix, q, k, largest in plan:
        key = (id(ix), q.tobytes(), k, largest)
        if key not in expected:
            enc = ix.encode_query(q)
            expected[key] = ix.search_top_k(enc, k, largest)
            enc_dev[key] = torch.from_numpy(np.frombuffer(bytes(enc), np.uint8).copy()).cuda()
    torch.cuda.synchronize()
    out_s = torch.zeros((len(plan), 128), dtype=torch.float32, device="cuda")
    out_i = torch.zeros((len(plan), 128), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    for n, (ix, q, k, largest) in enumerate(plan):  # nothing but scans on the context's stream
        ix.search_top_k_dev(enc_dev[(id(ix), q.tobytes(), k, largest)].data_ptr(), k, largest, 0, out_s[n].data_ptr(), out_i[n].data_ptr())
    ctx.synchronize()
    hs, hi = out_s.cpu().numpy(), out_i.cpu().numpy().view(np.uint32)
    for n, (ix, q, k, largest) in enumerate(plan):
        es, ei = expected[(id(ix), q.tobytes(), k, largest)]
        assert np.array_equal(hi[n, :ei.size], ei), n
        assert np.array_equal(bits(hs[n, :es.size]), bits(es)), n
    for g in (g_big, g_small, b_big):
        g.close()
