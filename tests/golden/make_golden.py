#!/usr/bin/env python
"""Generates tests/golden/ref_kernels.npz: outputs of the REFERENCE ITSELF run in the build container.

The reference's tests hold no literal known-answer vectors, so the golden vectors are
made by executing the reference's own native scorers — quantization/cpp/sse.c and avx2.c compiled
unmodified into oracle/_ref/libsimd_utils_ref.so by oracle/Makefile (flags of
quantization/build.rs:28-34) — on seeded inputs.  Needs /root/reference (or a prebuilt oracle/_ref);
the fixture itself travels to the GPU box.   python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402

# (rows, actual_dim): reference test shape 80 (= dim 65 padded; hits the 16-byte AVX tail), bench shapes
U8_SHAPES = [(64, 16), (64, 48), (129, 80), (64, 128), (32, 1024), (32, 1536), (16, 2048)]
# (rows, row_bytes): u128 rows of dims 1, 387, 1536 and the u8-store word sizes of dims 33, 65
BIN_SHAPES = [(128, 16), (128, 64), (64, 192), (128, 8), (128, 4)]


def codes(seed, n, worst_rows=0, width=1):
    """7-bit codes from the SplitMix64 stream; the last `worst_rows` rows are all 127 / all 0 alternately"""
    c = (O.uniform(seed, n * width) * 128.0).astype(np.uint8).reshape(n, width)
    c = np.minimum(c, 127).astype(np.uint8)
    for r in range(worst_rows):
        c[n - 1 - r] = 127 if r % 2 == 0 else 0
    return c


def main():
    assert O.ref.available, "oracle/_ref/libsimd_utils_ref.so missing: run `make -C oracle` with /root/reference present"
    out = {}
    for si, (rows, ad) in enumerate(U8_SHAPES):
        v = codes(1000 + si, rows, worst_rows=2, width=ad)
        q = codes(2000 + si, 2, worst_rows=1, width=ad)  # q[0] random, q[1] all 127
        for qi in range(2):
            for name in ("impl_score_dot_avx", "impl_score_dot_sse", "impl_score_l1_avx", "impl_score_l1_sse"):
                r = np.array([O.ref.call_u8(name, q[qi], v[i]) for i in range(rows)], np.float32)
                out[f"u8_{ad}_q{qi}_{name}"] = r
        out[f"u8_{ad}_v"] = v
        out[f"u8_{ad}_q"] = q
    for si, (rows, rb) in enumerate(BIN_SHAPES):
        raw = (O.uniform(3000 + si, rows * rb) * 256.0).astype(np.uint8).reshape(rows, rb)
        q = (O.uniform(4000 + si, rb) * 256.0).astype(np.uint8)
        name, word = {16: ("impl_xor_popcnt_sse_uint128", 16), 64: ("impl_xor_popcnt_sse_uint128", 16),
                      192: ("impl_xor_popcnt_sse_uint128", 16), 8: ("impl_xor_popcnt_sse_uint64", 8),
                      4: ("impl_xor_popcnt_sse_uint32", 4)}[rb]
        r = np.array([O.ref.call_popcnt(name, q, raw[i], rb // word) for i in range(rows)], np.uint32)
        out[f"bin_{rb}_v"] = raw
        out[f"bin_{rb}_q"] = q
        out[f"bin_{rb}_{name}"] = r
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_kernels.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
