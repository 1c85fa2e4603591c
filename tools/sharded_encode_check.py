"""Sharded u8 encode over NCCL, run with torchrun --nproc-per-node G: every rank holds its row range of a
seeded corpus, ShardedEncodeU8 all-reduces min/max (and all-gathers the quantile sample), and the rank's shard must be
byte-identical to its slice of a single-GPU encode of the whole corpus done on the same GPU.  Then a sharded search
over the encoded shards must equal the single-GPU search."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import quantization_b200 as qz
from quantization_b200 import sharded as S

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)  # library calls and NCCL share one stream
ctx = qz.Context(local, stream=stream.cuda_stream)
count, dim = int(os.environ.get("ROWS", 400_000)), 1536
ok_all = True
for quantile in (None, 0.99):
    g = torch.Generator(device=dev); g.manual_seed(42)
    data = torch.rand((count, dim), generator=g, device=dev)  # every rank draws the same corpus, keeps its rows
    torch.cuda.synchronize()
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    whole = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, quantile, ctx=ctx, seed=0)
    lo, hi = S.shard_range(count, rank, world)
    enc = S.ShardedEncodeU8(ctx, device=dev)
    mine = data[lo:hi].contiguous(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    shard = enc.encode(mine, vp, quantile, seed=0)
    torch.cuda.synchronize(); ms = (time.perf_counter() - t0) * 1e3
    rb = whole.row_bytes
    same = np.array_equal(shard.download(), whole.download()[lo * rb: hi * rb]) and shard.alpha == whole.alpha and shard.offset == whole.offset
    q = np.random.default_rng(43).random(dim, dtype=np.float32)
    ss = S.ShardedSearch(shard, device=dev)
    d_q = torch.from_numpy(np.frombuffer(bytes(shard.encode_query(q)), np.uint8).copy()).to(dev)
    s, i = ss.search_top_k(d_q.data_ptr(), 100, True)
    ws, wi = whole.search(q, 100, True)
    same_search = np.array_equal(i, wi) and np.array_equal(s.view(np.uint32), ws.view(np.uint32))
    flag = torch.tensor([int(same and same_search)], device=dev)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"world {world} quantile {quantile}: shards byte-identical to the single-GPU encode and sharded top-100 == single-GPU: "
              f"{bool(flag.item())}; sharded encode of {hi - lo} rows/rank {ms:.2f} ms wall", flush=True)
    ok_all &= bool(flag.item())
    del data, whole, shard, mine
# PQ: all-gather of the k-means sample, identical training on every rank, local assignment
pcount, pdim = 100_000, 256
g = torch.Generator(device=dev); g.manual_seed(7)
data = torch.rand((pcount, pdim), generator=g, device=dev); torch.cuda.synchronize()
vp = qz.VectorParameters(pdim, pcount, qz.DistanceType.Dot, False)
whole = qz.EncodedVectorsPQ.encode_dev(data.data_ptr(), vp, 2, ctx=ctx, seed=3)
lo, hi = S.shard_range(pcount, rank, world)
mine = data[lo:hi].contiguous(); torch.cuda.synchronize()
t0 = time.perf_counter()
shard = S.ShardedEncodePQ(ctx, device=dev).encode(mine, vp, 2, seed=3)
torch.cuda.synchronize(); ms = (time.perf_counter() - t0) * 1e3
nch = pdim // 2
same = np.array_equal(shard.download(), whole.download()[lo * nch: hi * nch]) and \
    np.array_equal(shard.centroids.view(np.uint32), whole.centroids.view(np.uint32))
flag = torch.tensor([int(same)], device=dev)
if world > 1:
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"world {world} PQ {pcount} x {pdim} chunk 2: centroids bit-identical on every rank and code shards are slices of the "
          f"single-GPU codes: {bool(flag.item())}; {ms:.1f} ms wall", flush=True)
ok_all &= bool(flag.item())
if world > 1:
    dist.destroy_process_group()
sys.exit(0 if ok_all else 1)
