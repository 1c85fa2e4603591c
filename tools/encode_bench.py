"""Encode-side throughput (u8 encode, binary encode, PQ k-means / assign / LUT build reported separately).
Device-resident f32 input -> device-resident encoded storage, whole C-ABI call timed with CUDA events on the
context stream (+ wall clock for calls that synchronise inside, e.g. the k-means convergence loop).
   python tools/encode_bench.py [u8] [u8q] [bin] [pqassign] [pqtrain] [lut]      (default: all)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import quantization_b200 as qz

what = set(sys.argv[1:]) or {"u8", "u8q", "bin", "pqassign", "pqtrain", "lut"}
rows = int(os.environ.get("ROWS", 1_000_000))
dim = int(os.environ.get("DIM", 1536))
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = qz.Context(0, stream=stream.cuda_stream)
g = torch.Generator(device=dev); g.manual_seed(42)
data = torch.rand((rows, dim), generator=g, device=dev)
vp = qz.VectorParameters(dim, rows, qz.DistanceType.Dot, False)
in_bytes = rows * dim * 4


def timed(name, f, reps, alg_bytes=None, flops=None, note=""):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(stream)
    for _ in range(reps): r = f()
    e1.record(stream); torch.cuda.synchronize(); wall = (time.perf_counter() - t0) / reps * 1e3
    ms = e0.elapsed_time(e1) / reps
    s = f"{name}: {ms:.3f} ms (wall {wall:.3f})  {rows / ms / 1e6:.2f} G vectors/s"
    if alg_bytes: s += f"  {alg_bytes / ms / 1e6:.0f} GB/s algorithmic"
    if flops: s += f"  {flops / ms / 1e9:.2f} TFLOP/s f32"
    print(s + ("  " + note if note else ""), flush=True)
    return r


import ctypes as C
from quantization_b200 import _lib
from quantization_b200.encoded_vectors import ptr
L = ctx._L


def u8_call(ix, quantile):  # the C-ABI call alone, into an index allocated once (no cudaMalloc in the timed region)
    a, o, m = C.c_float(), C.c_float(), C.c_float()
    qv = C.c_float(quantile if quantile is not None else 1.0)  # kept alive for the call
    q = C.cast(C.pointer(qv), _lib.f32p) if quantile is not None else None
    ctx.check(L.qz_u8_encode_dev(ctx.handle, C.c_void_p(data.data_ptr()), rows, dim, 0, 0, q, None, 0, 1, None,
                                 C.c_void_p(ix.storage_dev), C.byref(a), C.byref(o), C.byref(m)))


if "u8" in what or "u8q" in what:
    ix8 = qz.EncodedVectorsU8(ctx, None, vp, 0.0, 0.0, 0.0)
if "u8" in what:  # minmax pass + encode pass: 2*4*dim + actual_dim + 4 bytes per vector
    timed("u8 encode (min/max + codes + headers)", lambda: u8_call(ix8, None), 10, alg_bytes=rows * (8 * dim + dim + 4))
if "u8q" in what:  # + sample gather and two radix selects over <= 100k sampled vectors
    timed("u8 encode, quantile 0.99", lambda: u8_call(ix8, 0.99), 5, alg_bytes=rows * (8 * dim + dim + 4))
if "bin" in what:
    ix = qz.EncodedVectorsBin.empty(vp, ctx=ctx)
    timed("binary encode", lambda: ix.encode_rows_dev(data.data_ptr(), 0, rows), 10, alg_bytes=rows * (4 * dim + dim // 8))
    del ix
cent = None
if "pqtrain" in what:  # 768 k-means problems (10k-vector sample, 256 centroids, <= 100 iterations, 1e-5) side by side
    def train():
        c = np.zeros((256, dim), np.float32); it = np.zeros(dim // 2, np.uint32)
        ctx.check(ctx._L.qz_pq_train_dev(ctx.handle, C.c_void_p(data.data_ptr()), rows, dim, 2, None, 0, 7, 100, 1e-5, None, ptr(c), ptr(it)))
        return c, it
    cent, it = timed("pq train (k-means, chunk 2)", train, 1)
    n_s = min(rows, 10_000)
    print(f"   iterations per chunk: min {it.min()} median {int(np.median(it))} max {it.max()}; "
          f"assignment flops per iteration = {n_s} x 256 x {dim} x 3", flush=True)
if "pqassign" in what:  # argmin over 256 centroids per chunk: 3 flops per (vector, centroid, dim)
    if cent is None:
        cent = np.random.default_rng(3).random((256, dim), dtype=np.float32)
    ixp = qz.EncodedVectorsPQ(ctx, None, vp, 2, cent)
    timed("pq assign (chunk 2)", lambda: ctx.check(L.qz_pq_assign_dev(ctx.handle, C.c_void_p(data.data_ptr()), rows, dim, 2,
                                                                       ptr(ixp.centroids), None, C.c_void_p(ixp.storage_dev))),
          3, alg_bytes=rows * (4 * dim + dim // 2), flops=rows * 256 * dim * 3)
if "lut" in what:
    small = qz.VectorParameters(dim, 1000, qz.DistanceType.Dot, False)
    if cent is None:
        cent = np.random.default_rng(3).random((256, dim), dtype=np.float32)
    p = qz.EncodedVectorsPQ.encode_dev(data.data_ptr(), small, 2, ctx=ctx, centroids=cent)
    for nq in (1, 64, 4096):
        qs = torch.rand((nq, dim), generator=g, device=dev)
        out = torch.empty((nq, p.query_bytes), dtype=torch.uint8, device=dev)
        f = lambda: p.encode_query_dev(qs.data_ptr(), nq, out.data_ptr())
        f(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(20): f()
        e1.record(stream); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(f"pq LUT build nq={nq}: {ms * 1e3:.1f} us  ({nq * p.query_bytes / ms / 1e6:.0f} GB/s of LUT written)", flush=True)
