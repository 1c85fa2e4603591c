"""tools/recall.py (demos/src/ann_benchmark_data.rs:123-220) on a tiny on-disk triplet: the file
readers, the same_10/20/30 bookkeeping and the three codecs' search path, end to end."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

pytestmark = pytest.mark.gpu


def test_recall_on_fvecs_triplet(tmp_path):
    import recall as R
    rng = np.random.default_rng(3)
    centers = rng.standard_normal((20, 64)).astype(np.float32)
    train = (centers[rng.integers(0, 20, 5000)] + 0.5 * rng.standard_normal((5000, 64))).astype(np.float32)
    test = (centers[rng.integers(0, 20, 16)] + 0.5 * rng.standard_normal((16, 64))).astype(np.float32)
    tn = train / np.linalg.norm(train, axis=1, keepdims=True)
    qn = test / np.linalg.norm(test, axis=1, keepdims=True)
    gt = np.argsort(-(qn @ tn.T), axis=1, kind="stable")[:, :100].astype(np.int32)
    paths = [str(tmp_path / n) for n in ("base.fvecs", "query.fvecs", "gt.ivecs")]
    for p, a in zip(paths, (train, test, gt)):
        R.write_vecs(p, a)
        assert np.array_equal(R.read_vecs(p), a)
    lines = []
    res = R.run(paths[0], paths[1], paths[2], metric="cosine", queries=16, codecs=("u8", "bin", "pq"), out=lines.append)
    assert [r["name"].split()[0] for r in res] == ["u8", "binary", "pq"]
    u8, b, pq = res
    # 8-bit codes reproduce almost all exact neighbours; 1 bit per dimension far fewer, but far above chance
    assert u8["same_10"] >= 9.0 and u8["same_30"] >= 27.0, u8
    assert 1.0 <= b["same_30"] <= u8["same_30"], b
    assert pq["same_30"] >= 15.0, pq
    for r in res:
        assert r["same_10"] <= 10 and r["same_20"] <= 20 and r["same_30"] <= 30
        assert set(r["timings"]) == {"min", "avg", "p95", "p99", "max"} and r["timings"]["min"] <= r["timings"]["max"]
    text = "\n".join(lines)
    assert "same_10:" in text and "p99 search time:" in text and "gt.ivecs" in text
