"""A/B timing of the scan kernel: full-score mode (no top-k) vs fused top-k.  WL=u8|bin.  Experiment only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import quantization_b200 as qz

WL = os.environ.get("WL", "u8")
rows, dim, k = (1_000_000 if WL == "u8" else 10_000_000), 1536, 100
row_bytes = 1540 if WL == "u8" else 192
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = qz.Context(0, stream=stream.cuda_stream)
g = torch.Generator(device=dev); g.manual_seed(42)
vp = qz.VectorParameters(dim, rows, qz.DistanceType.Dot, False)
if WL == "u8":
    data = torch.rand((rows, dim), generator=g, device=dev)
    ix = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, None, ctx=ctx)
    del data
else:
    ix = qz.EncodedVectorsBin.empty(vp, ctx=ctx)
    for r in range(0, rows, 1_000_000):
        chunk = torch.rand((1_000_000, dim), generator=g, device=dev) - 0.5
        ix.encode_rows_dev(chunk.data_ptr(), r, 1_000_000)
        del chunk
q = np.random.default_rng(43).random(dim, dtype=np.float32) - (0.0 if WL == "u8" else 0.5)
enc = torch.from_numpy(ix.encode_query(q).copy()).to(dev)
d_scores = torch.zeros(rows, dtype=torch.float32, device=dev)
d_s = torch.zeros(k, dtype=torch.float32, device=dev); d_i = torch.zeros(k, dtype=torch.int32, device=dev)
def run(fn, n=200):
    for _ in range(10): fn()
    ctx.set_timing(True)
    for _ in range(n): fn()
    torch.cuda.synchronize()
    ms, cnt = ctx.get_timing(); ctx.set_timing(False)
    return ms / cnt
a = run(lambda: ix.score_all_dev(enc.data_ptr(), d_scores.data_ptr()))
b = run(lambda: ix.search_top_k_dev(enc.data_ptr(), k, True, 0, d_s.data_ptr(), d_i.data_ptr()))
print(f"{WL} cfg {os.environ.get('QZ_SCAN_TILE_BYTES')} {os.environ.get('QZ_SCAN_SLOTS')}: scores-only {a:.4f} ms ({rows*row_bytes/a/1e6:.0f} GB/s)   top-k {b:.4f} ms ({rows*row_bytes/b/1e6:.0f} GB/s)")
