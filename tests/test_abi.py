"""The drop-in boundary: libqz_cuda.so loads and exports exactly what include/qz_cuda.h declares.
No compute here (no GPU on the CPU box)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "qz_cuda.h")).read()
    return sorted(set(re.findall(r"QZ_API\s+[\w\s\*]+?\b(qz_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from quantization_b200 import _lib, build
    build.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in qz_cuda.h but not exported"


def test_python_binding_covers_header():
    from quantization_b200 import _lib
    assert _lib.declared_symbols() == header_symbols()


def test_pure_host_entry_points():
    from quantization_b200 import _lib
    L = _lib.lib()
    assert L.qz_u8_actual_dim(65) == 80 and L.qz_u8_row_bytes(1536) == 1540   # u8.rs:252-259
    # binary.rs:99-116 (u8 store) / :152-159 (u128 store)
    assert [L.qz_bin_row_bytes(d, 1) for d in (0, 1, 8, 33, 65, 387)] == [0, 1, 1, 8, 16, 64]
    assert [L.qz_bin_row_bytes(d, 16) for d in (1, 387, 1536)] == [16, 64, 192]


def test_no_gpu_fails_loudly():
    """Without a device the context constructor must raise, never fall back."""
    import pytest
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("GPU present")
    import quantization_b200 as qz
    with pytest.raises(qz.EncodingError):
        qz.Context(0)
