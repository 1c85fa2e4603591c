"""small PQ scan (3 LUT slices, several tiles) for sanitizer runs"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import quantization_b200 as qz
ctx = qz.Context(0)
dim, chunk, count = 640, 2, 20_000
n_chunks = dim // chunk
rng = np.random.default_rng(1)
codes = rng.integers(0, 256, size=(count, n_chunks), dtype=np.uint8)
cen = rng.random((256, dim), dtype=np.float32)
q = rng.random(dim, dtype=np.float32)
g = qz.EncodedVectorsPQ(ctx, codes.reshape(-1), qz.VectorParameters(dim, count, qz.DistanceType.Dot, False), chunk, cen)
lut = g.encode_query(q)
s = g.score_all(lut)
L = lut.view(np.float32).reshape(n_chunks, 256)
ref = L[np.arange(n_chunks)[None, :], codes].astype(np.float64).sum(1)
print("max err", np.abs(s - ref).max())
ks, ki = g.search_top_k(lut, 100, True)
print("topk ok", np.array_equal(ki, np.argsort(-s, kind="stable")[:100].astype(np.uint32)))
