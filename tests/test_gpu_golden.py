"""GPU parity against the committed golden vectors (tests/golden/ref_kernels.npz = outputs of the
reference's own cpp/sse.c + avx2.c scorers, see tests/golden/make_golden.py).  The CUDA scan kernels
are driven through the C-ABI on the fixture's raw rows with multiplier 1 and zero offsets, so the
returned f32 score IS the integer sum the reference kernel returned: bit-exact."""
import os

import numpy as np
import pytest

import quantization_b200 as qz

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_kernels.npz"))


@pytest.mark.parametrize("ad", [16, 48, 80, 128, 1024, 1536, 2048])
@pytest.mark.parametrize("dist,name", [(qz.DistanceType.Dot, "impl_score_dot_avx"), (qz.DistanceType.L2, "impl_score_dot_avx"),
                                       (qz.DistanceType.L1, "impl_score_l1_avx")])
def test_u8_kernels_match_reference_outputs(ctx, ad, dist, name):
    v, q = GOLD[f"u8_{ad}_v"], GOLD[f"u8_{ad}_q"]
    rows = v.shape[0]
    storage = np.zeros((rows, ad + 4), np.uint8)  # header f32 vector_offset = 0.0
    storage[:, 4:] = v
    vp = qz.VectorParameters(ad, rows, dist, False)
    ix = qz.EncodedVectorsU8(ctx, storage.reshape(-1), vp, 1.0, 0.0, 1.0)  # multiplier 1 -> score == S
    for qi in range(2):
        blob = np.zeros(ad + 4, np.uint8)  # query offset 0.0
        blob[4:] = q[qi]
        want = GOLD[f"u8_{ad}_q{qi}_{name}"]
        assert np.array_equal(ix.score_all(blob).view(np.uint32), want.view(np.uint32))
        assert np.array_equal(ix.score_points(blob, np.arange(rows)).view(np.uint32), want.view(np.uint32))
        s, ids = ix.search_top_k(blob, 10, True)
        order = np.lexsort((np.arange(rows), -want))[:10]
        assert np.array_equal(ids, order.astype(np.uint32)) and np.array_equal(s, want[order])
    ix.close()


@pytest.mark.parametrize("rb,dim,store,name", [(16, 128, 16, "impl_xor_popcnt_sse_uint128"),
                                               (64, 512, 16, "impl_xor_popcnt_sse_uint128"),
                                               (192, 1536, 16, "impl_xor_popcnt_sse_uint128"),
                                               (8, 64, 1, "impl_xor_popcnt_sse_uint64"),
                                               (4, 32, 1, "impl_xor_popcnt_sse_uint32")])
def test_binary_kernels_match_reference_outputs(ctx, rb, dim, store, name):
    v, q = GOLD[f"bin_{rb}_v"], GOLD[f"bin_{rb}_q"]
    rows = v.shape[0]
    vp = qz.VectorParameters(dim, rows, qz.DistanceType.Dot, True)  # (Dot, invert) -> score = xor - zeros
    ix = qz.EncodedVectorsBin(ctx, v.reshape(-1).copy(), vp, bits_store_type=store)
    assert ix.row_bytes == rb
    x = GOLD[f"bin_{rb}_{name}"].astype(np.float32)
    want = x - (np.float32(dim) - x)                                # calculate_metric, binary.rs:237-252
    assert np.array_equal(ix.score_all(q.copy()), want)
    ix.close()
