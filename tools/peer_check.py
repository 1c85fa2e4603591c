"""Multi-GPU check of the peer-memory top-k exchange (run under torchrun, one rank per GPU).

Every rank encodes the same seeded corpus, keeps its own point-id range, and answers the same queries three ways:
the peer-memory exchange fused into the scan launch (qz_scan_topk_peer_dev), its rooted form (qz_scan_topk_gather_dev:
only the root waits and merges; rooted and all-to-all steps are interleaved on ONE peer group, and the non-root
ranks are allowed to run ahead), the all-gather + merge path, and a single-GPU scan of the whole corpus.  All must
agree bit for bit (the rooted form on its root).
Prints per-step device times of both exchange paths.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import quantization_b200 as qz
from quantization_b200.sharded import ShardedSearch, shard_storage


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx = qz.Context(local, stream=stream.cuda_stream)
    ok = True
    for codec, count, dim in (("u8", 300_007, 256), ("bin", 500_003, 1536), ("u8", 1000, 64), ("pq", 200_003, 512)):
        rng = np.random.default_rng(5)
        data = rng.random((count, dim), dtype=np.float32) - (0.5 if codec == "bin" else 0.0)
        vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
        if codec == "u8":
            full = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, ctx=ctx)
            row = full.get_quantized_vector_size(vp)
            part, n, base = shard_storage(np.asarray(full.encoded_vectors), row, count, rank, world)
            vps = qz.VectorParameters(dim, n, qz.DistanceType.Dot, False)
            shard = qz.EncodedVectorsU8(ctx, part, vps, full.alpha, full.offset, full.multiplier, id_base=base)
        elif codec == "pq":
            n_chunks = dim // 2
            codes = rng.integers(0, 256, size=(count, n_chunks), dtype=np.uint8)
            cen = rng.random((256, dim), dtype=np.float32)
            full = qz.EncodedVectorsPQ(ctx, codes.reshape(-1), vp, 2, cen)
            part, n, base = shard_storage(codes.reshape(-1), n_chunks, count, rank, world)
            vps = qz.VectorParameters(dim, n, qz.DistanceType.Dot, False)
            shard = qz.EncodedVectorsPQ(ctx, part, vps, 2, cen, id_base=base)
        else:
            full = qz.EncodedVectorsBin.encode(data, qz.VecStorageBuilder(), vp, ctx=ctx)
            row = full.row_bytes
            part, n, base = shard_storage(np.asarray(full.encoded_vectors), row, count, rank, world)
            vps = qz.VectorParameters(dim, n, qz.DistanceType.Dot, False)
            shard = qz.EncodedVectorsBin(ctx, part, vps, id_base=base)
        s_peer = ShardedSearch(shard, None, dev, peer=True)
        s_nccl = ShardedSearch(shard, None, dev)
        for qi in range(4):
            q = np.random.default_rng(100 + qi).random(dim, dtype=np.float32) - (0.5 if codec == "bin" else 0.0)
            enc = full.encode_query(q)
            d_q = torch.from_numpy(np.frombuffer(bytes(enc), np.uint8).copy()).to(dev)
            for k, largest in [tuple(map(int, t.split(":"))) for t in os.environ.get("PEER_CHECK_KS", "100:1,1:1,128:0,17:1").split(",")]:
                a_s, a_i = s_peer.search_top_k(d_q.data_ptr(), k, largest, count)
                b_s, b_i = s_nccl.search_top_k(d_q.data_ptr(), k, largest, count)
                r_s, r_i = full.search_top_k(enc, k, largest)
                same = (np.array_equal(a_i, b_i) and np.array_equal(a_i, r_i)
                        and np.array_equal(a_s.view(np.uint32), r_s.view(np.uint32))
                        and np.array_equal(b_s.view(np.uint32), r_s.view(np.uint32)))
                # rooted gather on the same peer group, several steps back to back without a host synchronisation
                # in between (non-root ranks run ahead; the root's acknowledgement is the only flow control), with
                # a changing root and an all-to-all step in the middle
                for root in ((world - 1, 0) if os.environ.get("PEER_CHECK_REV") else (0, world - 1)):
                    if os.environ.get("PEER_CHECK_TRACE"):
                        torch.cuda.synchronize()
                        print(f"[trace] rank {rank} {codec} {count} q{qi} k={k} root={root}", file=sys.stderr, flush=True)
                    s_peer.root = root
                    for _ in range(5):
                        g_s, g_i = s_peer.search_top_k_dev(d_q.data_ptr(), k, largest)
                        if os.environ.get("PEER_CHECK_TRACE"):
                            torch.cuda.synchronize()
                            print(f"[trace] rank {rank} rooted step done", file=sys.stderr, flush=True)
                    s_peer.root = None
                    s_peer.search_top_k_dev(d_q.data_ptr(), k, largest)
                    s_peer.root = root
                    g_s, g_i = s_peer.search_top_k_dev(d_q.data_ptr(), k, largest)
                    torch.cuda.synchronize()
                    if rank == root:
                        n = min(k, count)
                        gs, gi = g_s.cpu().numpy()[:n], g_i.cpu().numpy().view(np.uint32)[:n]
                        same = same and np.array_equal(gi, r_i) and np.array_equal(gs.view(np.uint32), r_s.view(np.uint32))
                s_peer.root = None
                if not same:
                    ok = False
                    print(f"rank {rank}: MISMATCH {codec} {count}x{dim} q{qi} k={k} largest={largest}", flush=True)
        # timing: device time per step, max over ranks
        if count >= 100_000:
            for name, s in (("peer", s_peer), ("gather", s_peer), ("nccl", s_nccl)):
                s_peer.root = 0 if name == "gather" else None
                for _ in range(20):
                    s.search_top_k_dev(d_q.data_ptr(), 100, True)
                dist.barrier()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(500):
                    s.search_top_k_dev(d_q.data_ptr(), 100, True)
                e1.record()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) / 500], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                if rank == 0:
                    print(f"{codec} {count}x{dim} world {world} {name}: {t.item() * 1000:.1f} us/step", flush=True)
            s_peer.root = None
        s_peer.close()
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    if rank == 0:
        print("PEER_CHECK", "OK" if flag.item() == 0 else "FAILED", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 0 else 1)


if __name__ == "__main__":
    main()
