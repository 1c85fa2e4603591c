"""debug: bit-level repeatability of the PQ streaming scan (experiment only)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle as O
import quantization_b200 as qz
ctx = qz.Context(0)
def bits(a): return np.ascontiguousarray(a, np.float32).view(np.uint32)
def case(dim, chunk, count):
    n_chunks = (dim + chunk - 1) // chunk
    rng = np.random.default_rng(dim + count)
    codes = rng.integers(0, 256, size=(count, n_chunks), dtype=np.uint8)
    cen = O.uniform(77, 256 * dim).reshape(256, dim)
    query = O.uniform(78, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsPQ(ctx, codes.reshape(-1), vp, chunk, cen)
    glut = g.encode_query(query)
    lut = glut.view(np.float32).reshape(n_chunks, 256)
    ref = lut[np.arange(n_chunks)[None, :], codes].astype(np.float64).sum(1)
    runs = [g.score_all(glut) for _ in range(3)]
    for n, r in enumerate(runs[:-1]):
        bad = np.nonzero(bits(r) != bits(runs[-1]))[0]
        print(dim, chunk, count, "run", n, "rows differing from the last run:", bad.size)
        if bad.size:
            per_cta = -(-count // 148); per_cta = -(-per_cta // 64) * 64
            print("   block_rows", per_cta, "CTAs:", sorted(set((bad // per_cta).tolist())))
            for row in bad[:40]:
                print("   row", int(row), "in-block", int(row % per_cta), "this %.7f last %.7f ref %.7f" % (r[row], runs[-1][row], ref[row]))
    for largest in (True, False):
        for k in (1, 100, 128, 200):
            g.search_top_k(glut, k, largest)
case(32, 2, 1_300_003)
case(2048, 8, 60_001)
case(320, 1, 9_000)
case(2048, 8, 60_001)
