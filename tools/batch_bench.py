"""Timing of the batched tcgen05 path (C5 shape per GPU: 4096 queries x rows x 768-d). Experiment / profiling driver."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import quantization_b200 as qz

rows = int(os.environ.get("ROWS", 1_250_000)); dim = int(os.environ.get("DIM", 768)); nq = int(os.environ.get("NQ", 4096)); k = 100
reps = int(os.environ.get("REPS", 5))
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = qz.Context(0, stream=stream.cuda_stream)
g = torch.Generator(device=dev); g.manual_seed(42)
ix = None
vp = qz.VectorParameters(dim, rows, qz.DistanceType.Dot, False)
data = torch.rand((rows, dim), generator=g, device=dev)
ix = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, None, ctx=ctx)
del data
q = torch.rand((nq, dim), generator=g, device=dev)
d_s = torch.zeros((nq, k), dtype=torch.float32, device=dev); d_i = torch.zeros((nq, k), dtype=torch.int32, device=dev)
for _ in range(2): ix.search_batch_dev(q.data_ptr(), nq, k, True, d_s.data_ptr(), d_i.data_ptr())
torch.cuda.synchronize()
ctx.set_timing(True)
t0 = time.perf_counter()
for _ in range(reps): ix.search_batch_dev(q.data_ptr(), nq, k, True, d_s.data_ptr(), d_i.data_ptr())
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / reps
kms, kn = ctx.get_timing(); ctx.set_timing(False)
ops = 2.0 * nq * rows * ix.actual_dim
print(f"batch {nq} x {rows} x {dim}: {dt*1e3:.3f} ms/call wall, gemm kernels {kms/reps:.3f} ms/call ({kn//reps} launches) "
      f"-> {ops/dt/1e12:.1f} TOP/s wall, {ops/(kms/reps*1e-3)/1e12:.1f} TOP/s in-kernel; {nq*rows/dt/1e9:.2f} G scored pairs/s")
# spot parity vs the single-query scan
qs = q[:3].cpu().numpy()
for i in range(3):
    ws, wi = ix.search_top_k(ix.encode_query(qs[i]), k, True)
    assert np.array_equal(d_i[i].cpu().numpy().view(np.uint32), wi), "batched ids differ from the single-query scan"
print("parity ok")
