"""out-of-core scan throughput: 1M x 1536-d u8 storage in pinned host memory, streamed in chunks"""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import quantization_b200 as qz

rows, dim, k = 1_000_000, 1536, 100
dev = torch.device("cuda", 0)
ctx = qz.Context(0)
g = torch.Generator(device=dev); g.manual_seed(42)
data = torch.rand((rows, dim), generator=g, device=dev)
vp = qz.VectorParameters(dim, rows, qz.DistanceType.Dot, False)
ix = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, None, ctx=ctx); del data
host = np.empty(rows * 1540, np.uint8)
ctx.check(ctx._L.qz_index_download(ix.handle, C.c_void_p(host.ctypes.data)))
q = ix.encode_query(np.random.default_rng(43).random(dim, dtype=np.float32))
rs, ri = ix.search_top_k(q, k, True)
for pinned in (False, True):
    if pinned: ctx.check(ctx._L.qz_host_pin(ctx.handle, C.c_void_p(host.ctypes.data), host.nbytes))
    for chunk in (32_768, 131_072, 500_000):
        vpc = qz.VectorParameters(dim, chunk, qz.DistanceType.Dot, False)
        a = qz.EncodedVectorsU8(ctx, None, vpc, ix.alpha, ix.offset, ix.multiplier)
        b = qz.EncodedVectorsU8(ctx, None, vpc, ix.alpha, ix.offset, ix.multiplier)
        a.search_top_k_chunked(b, host, rows, q, k, True)
        t0 = time.perf_counter(); n = 3
        for _ in range(n): s, i = a.search_top_k_chunked(b, host, rows, q, k, True)
        dt = (time.perf_counter() - t0) / n
        ok = np.array_equal(i, ri) and np.array_equal(s.view(np.uint32), rs.view(np.uint32))
        print(f"pinned={pinned} chunk={chunk}: {dt*1e3:.1f} ms  {host.nbytes/dt/1e9:.1f} GB/s  {rows/dt/1e6:.1f} M vec/s  parity {ok}", flush=True)
        del a, b
