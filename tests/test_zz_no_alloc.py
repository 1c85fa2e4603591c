"""qz_merge_topk_static_dev: the allocation-free variant of qz_merge_topk_dev (include/qz_cuda.h; no counterpart in
the reference).  Its intermediate lists live in a file-scope __device__
array of the library (csrc/no_alloc.cu) instead of cudaMalloc'ed scratch, so the test is: same inputs through both
entry points, results identical bit for bit (integer / index work: no tolerance), on calls made back to back so that
the second one runs over the first one's leftovers in the shared array.

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest

import oracle as O
import quantization_b200 as qz

pytestmark = pytest.mark.gpu

MAX_LISTS = 4096  # QZ_MERGE_STATIC_MAX_LISTS


def make_keys(scores, ids, largest):
    """u64 keys of topk.cuh: (order-preserving bits of the score ^ flip) << 32 | (0xFFFFFFFF - id)"""
    b = np.ascontiguousarray(scores, np.float32).view(np.uint32)
    o = np.where(b >> 31, ~b, b | np.uint32(0x80000000)).astype(np.uint32)
    if not largest:
        o = ~o
    return (o.astype(np.uint64) << np.uint64(32)) | (np.uint32(0xFFFFFFFF) - ids.astype(np.uint32)).astype(np.uint64)


def sorted_lists(seed, n_lists, k, n_per, largest, mode="random"):
    """[n_lists][k] best-first keys: the k best of n_per scored points per list, ids unique over all lists"""
    rng = np.random.default_rng(seed)
    if mode == "ties":
        sc = rng.integers(0, 3, (n_lists, n_per)).astype(np.float32)
    else:
        sc = rng.standard_normal((n_lists, n_per)).astype(np.float32)
    ids = np.arange(n_lists * n_per, dtype=np.uint32).reshape(n_lists, n_per)
    keys = make_keys(sc, ids, largest)
    keys = np.sort(keys, axis=1)[:, ::-1][:, :k]
    out = np.zeros((n_lists, k), np.uint64)  # lists shorter than k are padded with the empty key 0
    out[:, :keys.shape[1]] = keys
    return np.ascontiguousarray(out), sc.ravel(), ids.ravel()


def call(ctx, fn, d_in, n_lists, k, largest, check_mem=False):
    """one call through the C-ABI; with check_mem the device's free memory must be the same before and after it"""
    import torch
    d_k = torch.full((k,), -1, dtype=torch.int64, device="cuda")
    d_s = torch.full((k,), -1.0, dtype=torch.float32, device="cuda")
    d_i = torch.full((k,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    st = fn(ctx.handle, C.c_void_p(d_in.data_ptr()), n_lists, k, int(largest), C.c_void_p(d_k.data_ptr()),
            C.c_void_p(d_s.data_ptr()), C.c_void_p(d_i.data_ptr()))
    ctx.check(st)
    ctx.synchronize()
    if check_mem:
        assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the device's free memory"
    return d_k.cpu().numpy().view(np.uint64), d_s.cpu().numpy().view(np.uint32), d_i.cpu().numpy().view(np.uint32)


def test_key_helper_matches_oracle():
    rng = np.random.default_rng(0)
    s = np.concatenate([rng.standard_normal(40).astype(np.float32), np.float32([0.0, -0.0, np.inf, -np.inf])])
    i = rng.integers(0, 2**32, s.size, dtype=np.uint64).astype(np.uint32)
    for largest in (True, False):
        assert [int(x) for x in make_keys(s, i, largest)] == [O.make_key(float(a), int(b), largest) for a, b in zip(s, i)]


# every case needs more than 32768 / Kp lists, so at least one level of the merge goes through the scratch array;
# the sequence goes big -> small -> big so each call finds another call's keys in the array
CASES = [
    (1000, 100, 150, True, "random"),    # Kp 128: 256 lists per CTA, 4 intermediate lists
    (4096, 1024, 1024, False, "random"), # the stated maximum: 128 intermediate lists of 1024 keys, then 4, then 1
    (300, 64, 40, True, "ties"),         # input lists shorter than k, ids decide
    (4096, 1000, 1000, True, "random"),
    (700, 10, 40, False, "random"),
]


def test_static_merge_equals_merge_topk_dev_on_consecutive_calls(ctx):
    import torch
    L = ctx._L
    warm = False  # the very first launch loads the kernel (lazy module loading), which is set-up, not a call's cost
    for n, (n_lists, k, n_per, largest, mode) in enumerate(CASES):
        lists, s, i = sorted_lists(100 + n, n_lists, k, n_per, largest, mode)
        d_in = torch.from_numpy(lists.view(np.int64)).cuda()
        ek, es, ei = call(ctx, L.qz_merge_topk_dev, d_in, n_lists, k, largest)
        # two calls in a row on the same input, then the next case right behind them
        for rep in range(2):
            gk, gs, gi = call(ctx, L.qz_merge_topk_static_dev, d_in, n_lists, k, largest, check_mem=warm)
            warm = True
            assert np.array_equal(gk, ek), f"case {n} rep {rep}: keys differ"
            assert np.array_equal(gs, es) and np.array_equal(gi, ei), f"case {n} rep {rep}"
        if n == 0:  # and both agree with the oracle's merge of the underlying points
            os_, oi = O.merge_topk(s, i, k, largest)
            assert np.array_equal(gi[:oi.size], oi) and np.array_equal(gs[:os_.size], os_.view(np.uint32))


def test_static_merge_small_inputs_and_limits(ctx):
    """single-launch sizes (no scratch needed) still work; sizes beyond the array are refused with a clear error"""
    import torch
    L = ctx._L
    for n_lists, k in [(1, 100), (2, 7), (148, 128)]:
        lists, _, _ = sorted_lists(7, max(n_lists, 1), k, 2 * k, True)
        d_in = torch.from_numpy(lists.view(np.int64)).cuda()
        a = call(ctx, L.qz_merge_topk_dev, d_in, n_lists, k, True)
        b = call(ctx, L.qz_merge_topk_static_dev, d_in, n_lists, k, True)
        assert all(np.array_equal(x, y) for x, y in zip(a, b))
    d_in = torch.zeros(8, dtype=torch.int64, device="cuda")
    st = L.qz_merge_topk_static_dev(ctx.handle, C.c_void_p(d_in.data_ptr()), MAX_LISTS + 1, 1, 1, None, None, None)
    assert st == 1  # QZ_ARGUMENTS
    msg = L.qz_last_error(ctx.handle).decode()
    assert "QZ_MERGE_STATIC_MAX_LISTS" in msg and str(MAX_LISTS) in msg, msg
    assert L.qz_merge_topk_static_dev(ctx.handle, C.c_void_p(d_in.data_ptr()), 4, 0, 1, None, None, None) == 1
    assert L.qz_merge_topk_static_dev(ctx.handle, C.c_void_p(d_in.data_ptr()), 4, 1025, 1, None, None, None) == 1
    ctx.synchronize()
