"""debug driver for the rooted gather: a programmable sequence of exchange steps on 2+ GPUs (torchrun)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import quantization_b200 as qz
from quantization_b200.sharded import ShardedSearch, shard_storage

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = qz.Context(local, stream=stream.cuda_stream)
count, dim = 100_003, 128
data = np.random.default_rng(5).random((count, dim), dtype=np.float32)
vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
full = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, ctx=ctx)
row = full.get_quantized_vector_size(vp)
part, n, base = shard_storage(np.asarray(full.encoded_vectors), row, count, rank, world)
shard = qz.EncodedVectorsU8(ctx, part, qz.VectorParameters(dim, n, qz.DistanceType.Dot, False), full.alpha, full.offset, full.multiplier, id_base=base)
s = ShardedSearch(shard, None, dev, peer=True)
enc = full.encode_query(np.random.default_rng(1).random(dim, dtype=np.float32))
d_q = torch.from_numpy(np.frombuffer(bytes(enc), np.uint8).copy()).to(dev)
# sequence: tokens "a" (all-to-all), "0"/"1" (rooted at that rank), "kN" (switch k), "f" (plain full-index scan), "n" (nccl)
seq = os.environ.get("SEQ", "a 0 0 0 0 0 a 0 1 1 1 1 1 a 1 k1 a 0").split()
k = 100
for i, tok in enumerate(seq):
    if tok[0] == "k":
        k = int(tok[1:])
    elif tok == "f":
        full.search_top_k(enc, k, True)
    else:
        s.root = None if tok == "a" else int(tok)
        s.search_top_k_dev(d_q.data_ptr(), k, True)
    try:
        torch.cuda.synchronize()
    except Exception as e:
        print(f"rank {rank}: FAILED at token {i} '{tok}' (k={k}): {e}", flush=True)
        os._exit(1)
    print(f"rank {rank}: ok {i} '{tok}'", flush=True)
print(f"rank {rank}: sequence done", flush=True)
s.root = None
dist.barrier(); s.close(); dist.destroy_process_group()
