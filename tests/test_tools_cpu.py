"""CPU-side checks of the tooling that carries no GPU work: the TexMex / npy readers of tools/recall.py, its timing
summary (ann_benchmark_data.rs:203-220 index arithmetic), and the C -> Rust type mapping of tools/gen_rust_ffi.py."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_vecs_readers_roundtrip(tmp_path):
    import recall as R
    rng = np.random.default_rng(0)
    for name, a in (("a.fvecs", rng.standard_normal((7, 5)).astype(np.float32)),
                    ("b.ivecs", rng.integers(0, 1 << 30, (4, 100)).astype(np.int32)),
                    ("c.bvecs", rng.integers(0, 256, (3, 16)).astype(np.uint8)),
                    ("d.npy", rng.standard_normal((2, 3)).astype(np.float32))):
        p = str(tmp_path / name)
        R.write_vecs(p, a)
        b = R.read_vecs(p)
        assert b.dtype == a.dtype and np.array_equal(a, b), name
    # the record header really is the dimension as int32, little endian, in front of every row
    raw = np.fromfile(str(tmp_path / "a.fvecs"), np.int32)
    assert raw[0] == 5 and raw[6] == 5
    with open(str(tmp_path / "bad.fvecs"), "wb") as f:
        f.write(b"\x05\x00\x00\x00" + b"\x00" * 7)
    with pytest.raises(ValueError):
        R.read_vecs(str(tmp_path / "bad.fvecs"))


def test_print_timings_uses_the_reference_percentile_indices():
    import recall as R
    t = list(range(1, 201))  # 200 timings: p95 = sorted[(200 as f32 * 0.95) as usize] = sorted[190], p99 = sorted[198]
    lines = []
    r = R.print_timings(t[::-1], lines.append)
    assert (r["min"], r["max"], r["p95"], r["p99"]) == (1, 200, 191, 199) and r["avg"] == 100.5
    assert "p95 search time" in lines[0]
    assert R.print_timings([], lines.append) == {}


def test_rust_type_mapping():
    import gen_rust_ffi as G
    assert G.rust_type("const volatile int32_t *") == "*const i32"
    assert G.rust_type("qz_ctx **") == "*mut *mut qz_ctx"
    assert G.rust_type("const qz_ctx *") == "*const qz_ctx"
    assert G.rust_type("const void *") == "*const c_void"
    assert G.rust_type("uint64_t") == "u64" and G.rust_type("float *") == "*mut f32"
    protos = {n: (r, p) for n, r, p in G.prototypes()}
    assert protos["qz_last_error"] == ("*const c_char", [("ctx", "*const qz_ctx")])
    assert protos["qz_ctx_destroy"][0] == "" and protos["qz_u8_row_bytes"] == ("u32", [("dim", "u32")])
    stop = dict(protos["qz_u8_encode"][1])["stop"]
    assert stop == "*const i32"
