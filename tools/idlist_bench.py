"""id-list scoring throughput: score_point on n random ids, device-resident ids.  WL=u8|bin|pq"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import quantization_b200 as qz

WL = os.environ.get("WL", "u8")
rows = {"u8": 1_000_000, "bin": 10_000_000, "pq": 1_000_000}[WL]
row_bytes = {"u8": 1540, "bin": 192, "pq": 768}[WL]
dim = 1536
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = qz.Context(0, stream=stream.cuda_stream)
g = torch.Generator(device=dev); g.manual_seed(42)
vp = qz.VectorParameters(dim, rows, qz.DistanceType.Dot, False)
if WL == "u8":
    data = torch.rand((rows, dim), generator=g, device=dev)
    ix = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, None, ctx=ctx); del data
elif WL == "pq":
    data = torch.rand((rows, dim), generator=g, device=dev)
    ix = qz.EncodedVectorsPQ.encode_dev(data.data_ptr(), vp, 2, ctx=ctx); del data
else:
    ix = qz.EncodedVectorsBin.empty(vp, ctx=ctx)
    for r in range(0, rows, 1_000_000):
        chunk = torch.rand((1_000_000, dim), generator=g, device=dev) - 0.5
        ix.encode_rows_dev(chunk.data_ptr(), r, 1_000_000); del chunk
q = np.random.default_rng(43).random(dim, dtype=np.float32) - (0.5 if WL == "bin" else 0.0)
enc = ix.encode_query(q)
d_q = torch.from_numpy(np.frombuffer(bytes(enc), np.uint8).copy()).to(dev)
for n in (1_000, 32_768, 1_000_000):
    ids = torch.randint(0, rows, (n,), device=dev, dtype=torch.int32, generator=g)
    out = torch.zeros(n, dtype=torch.float32, device=dev)
    f = lambda: ix.score_points_dev(d_q.data_ptr(), ids.data_ptr(), n, out.data_ptr())
    for _ in range(5): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 200 if n < 100_000 else 30
    e0.record(stream)
    for _ in range(reps): f()
    e1.record(stream); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    # parity of a few against the host path
    h = ix.score_points(enc, ids[:64].cpu().numpy().astype(np.uint32))
    ok = np.array_equal(h.view(np.uint32), out[:64].cpu().numpy().view(np.uint32))
    print(f"{WL} n={n}: {ms*1000:.1f} us  {n/ms/1e3:.1f} M ids/s  {n*row_bytes/ms/1e6:.0f} GB/s of rows  parity {ok}", flush=True)
