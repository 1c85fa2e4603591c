"""Recall harness in the shape of the reference's test_knn (demos/src/ann_benchmark_data.rs:123-185): per query
encode_query + scan of every point + the 30 best, then same_10 / same_20 / same_30 against the exact f32 neighbours
and the timing percentiles of print_timings (:203-220).  The ann-benchmarks HDF5 files need the network, so the
data is synthetic: a Gaussian mixture on the unit sphere (cosine = dot after normalisation, like the reference's
cosine_preprocess).  The exact neighbours come from an f32 matmul (torch; harness only, not the product path).

    python tools/recall.py [rows] [dim] [queries]
"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import quantization_b200 as qz

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 256
nq = int(sys.argv[3]) if len(sys.argv) > 3 else 200
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(42)
centers = torch.randn((1000, dim), generator=g, device=dev)
assign = torch.randint(0, 1000, (rows,), generator=g, device=dev)
data = centers[assign] + 0.6 * torch.randn((rows, dim), generator=g, device=dev)
data = data / data.norm(dim=1, keepdim=True)
qa = torch.randint(0, 1000, (nq,), generator=g, device=dev)
queries = centers[qa] + 0.6 * torch.randn((nq, dim), generator=g, device=dev)
queries = queries / queries.norm(dim=1, keepdim=True)
exact = torch.topk(queries @ data.T, 30, dim=1).indices.cpu().numpy()  # neighbours of the dataset file
q_host = queries.cpu().numpy()
ctx = qz.Context(0)
vp = qz.VectorParameters(dim, rows, qz.DistanceType.Dot, False)


def print_timings(t):
    t = np.sort(np.asarray(t))
    print(f"  search time ms: min {t[0]:.3f}  avg {t.mean():.3f}  p95 {t[int(len(t) * 0.95)]:.3f}  "
          f"p99 {t[int(len(t) * 0.99)]:.3f}  max {t[-1]:.3f}")


def test_knn(name, enc):
    same = np.zeros(3)
    timings = []
    for j in range(nq):
        t0 = time.perf_counter()
        s, ids = enc.search(q_host[j], 30, True)  # encode_query + scan + top-30, one call
        timings.append((time.perf_counter() - t0) * 1e3)
        for n, kk in enumerate((10, 20, 30)):
            same[n] += len(set(ids[:kk].tolist()) & set(exact[j].tolist()))  # same_count(&knn[0..kk], neighbors)
    same /= nq
    print(f"{name}: queries {nq}  same_10 {same[0]:.2f}  same_20 {same[1]:.2f}  same_30 {same[2]:.2f}")
    print_timings(timings[5:])


t0 = time.perf_counter()
u8 = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, None, ctx=ctx)
torch.cuda.synchronize(); print(f"u8 encode {rows}x{dim}: {(time.perf_counter() - t0) * 1e3:.1f} ms")
test_knn("u8 (no quantile)", u8)
u8q = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, 0.99, ctx=ctx)
test_knn("u8 (quantile 0.99)", u8q)
del u8, u8q
b = qz.EncodedVectorsBin.empty(vp, ctx=ctx)
b.encode_rows_dev(data.data_ptr(), 0, rows)
test_knn("binary", b)
del b
t0 = time.perf_counter()
pq = qz.EncodedVectorsPQ.encode_dev(data.data_ptr(), vp, 2, ctx=ctx)
torch.cuda.synchronize(); print(f"pq train + assign (chunk 2): {(time.perf_counter() - t0) * 1e3:.1f} ms")
test_knn("pq chunk 2", pq)
