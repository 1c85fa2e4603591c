"""Multi-GPU parity of the peer-memory top-k exchange (csrc/peer.cu, row_scan.cuh PeerExchange).

With >= 2 GPUs: tools/peer_check.py under torchrun (one rank per GPU) — peer exchange, NCCL all-gather + merge and
a single-GPU scan of the whole corpus must agree bit for bit (u8 and binary, several k, both orders, a shard
smaller than one tile).  With 1 GPU: the world-size-1 forms of the same entry points against the plain calls."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle as O
import quantization_b200 as qz

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_peer_exchange_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29733", os.path.join(ROOT, "tools", "peer_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "PEER_CHECK OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_peer_world1_matches_plain(ctx):
    """world = 1: qz_scan_topk_peer_dev / qz_search_topk_peer are the plain scan (no exchange, no second merge)"""
    import torch
    count, dim, k = 20_011, 96, 50
    data = O.uniform(3, count * dim).reshape(count, dim)
    query = O.uniform(4, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, ctx=ctx)
    L = ctx._L
    handle, peer = np.zeros(64, np.uint8), C.c_void_p()
    ctx.check(L.qz_peer_create(ctx.handle, 1, 0, C.c_void_p(handle.ctypes.data), C.byref(peer)))
    try:
        ctx.check(L.qz_peer_connect(peer, C.c_void_p(handle.ctypes.data)))
        s, i, n = np.empty(k, np.float32), np.empty(k, np.uint32), C.c_uint32(0)
        q = np.ascontiguousarray(query, np.float32)
        ctx.check(L.qz_search_topk_peer(g.handle, peer, C.c_void_p(q.ctypes.data), k, 1, C.c_void_p(s.ctypes.data),
                                        C.c_void_p(i.ctypes.data), C.byref(n)))
        rs, ri = g.search_top_k(g.encode_query(query), k, True)
        assert n.value == k and np.array_equal(i, ri) and np.array_equal(s.view(np.uint32), rs.view(np.uint32))
        # argument errors
        assert L.qz_search_topk_peer(g.handle, peer, C.c_void_p(q.ctypes.data), 129, 1, C.c_void_p(s.ctypes.data),
                                     C.c_void_p(i.ctypes.data), C.byref(n)) == 1
        bad = C.c_void_p()
        assert L.qz_peer_create(ctx.handle, 9, 0, C.c_void_p(handle.ctypes.data), C.byref(bad)) == 1
        assert L.qz_peer_create(ctx.handle, 2, 2, C.c_void_p(handle.ctypes.data), C.byref(bad)) == 1
    finally:
        L.qz_peer_destroy(peer)
