"""qz_merge_topk_dev (the step after the NCCL all-gather, and the reducer of per-CTA lists): pruned
single-CTA merge, its exact tree-merge fallback, and the multi-launch tree for many lists."""
import ctypes as C

import numpy as np
import pytest

import oracle as O
import quantization_b200 as qz

pytestmark = pytest.mark.gpu


def keys_of(scores, ids, largest):
    return np.array([O.make_key(float(s), int(i), largest) for s, i in zip(scores, ids)], np.uint64)


def run_merge(ctx, lists, k, largest):
    import torch
    n_lists = len(lists)
    flat = np.zeros((n_lists, k), np.uint64)
    for g, l in enumerate(lists):
        flat[g, :len(l)] = l
    d_in = torch.from_numpy(flat.view(np.int64)).cuda()
    d_s = torch.zeros(k, dtype=torch.float32, device="cuda")
    d_i = torch.zeros(k, dtype=torch.int32, device="cuda")
    d_k = torch.zeros(k, dtype=torch.int64, device="cuda")
    ctx.check(ctx._L.qz_merge_topk_dev(ctx.handle, C.c_void_p(d_in.data_ptr()), n_lists, k, int(largest),
                                       C.c_void_p(d_k.data_ptr()), C.c_void_p(d_s.data_ptr()), C.c_void_p(d_i.data_ptr())))
    ctx.synchronize()
    return d_k.cpu().numpy().view(np.uint64), d_s.cpu().numpy(), d_i.cpu().numpy().view(np.uint32)


def make_lists(rng, n_lists, k, n_per, largest, mode):
    lists, all_s, all_i = [], [], []
    for g in range(n_lists):
        ids = np.arange(g * n_per, (g + 1) * n_per, dtype=np.uint32)
        if mode == "random":
            sc = rng.standard_normal(n_per).astype(np.float32)
        elif mode == "stacked":      # list g entirely above list g+1: defeats head-based pruning
            sc = (rng.random(n_per).astype(np.float32) + np.float32(n_lists - g))
        elif mode == "ties":         # a handful of distinct values: ids decide
            sc = rng.integers(0, 3, n_per).astype(np.float32)
        else:                        # "flat": everything equal
            sc = np.zeros(n_per, np.float32)
        ks = np.sort(keys_of(sc, ids, largest))[::-1][:k]
        lists.append(ks)
        all_s.append(sc)
        all_i.append(ids)
    return lists, np.concatenate(all_s), np.concatenate(all_i)


@pytest.mark.parametrize("mode", ["random", "stacked", "ties", "flat"])
@pytest.mark.parametrize("n_lists,k,n_per", [(1, 100, 500), (2, 100, 60), (8, 100, 1000), (8, 1024, 3000),
                                              (148, 100, 300), (148, 128, 128), (148, 1000, 1200), (300, 64, 100),
                                              (700, 10, 40)])
@pytest.mark.parametrize("largest", [True, False])
def test_merge(ctx, mode, n_lists, k, n_per, largest):
    rng = np.random.default_rng(n_lists * 1000 + k)
    lists, s, i = make_lists(rng, n_lists, k, n_per, largest, mode)
    gk, gs, gi = run_merge(ctx, lists, k, largest)
    es, ei = O.merge_topk(s, i, k, largest)
    m = es.size
    assert np.array_equal(gi[:m], ei)
    assert np.array_equal(gs[:m].view(np.uint32), es.view(np.uint32))
    assert not gk[m:].any()  # unfilled slots are the empty key
