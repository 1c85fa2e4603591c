"""The drop-in boundary: libqz_cuda.so loads and exports exactly what include/qz_cuda.h declares.
No compute here (no GPU on the CPU box)."""
import ctypes
import os
import re
import sys

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


def test_header_compiles_as_c99_and_links(tmp_path):
    """include/qz_cuda.h is a plain C header: tests/abi_smoke.c (C99, -pedantic -Werror) includes it, takes the address
    of every declared function, links -lqz_cuda and runs (the boundary a cgo / bindgen / C host sees)"""
    import subprocess
    from quantization_b200 import _lib, build
    build.build()
    src = os.path.join(ROOT, "tests", "abi_smoke.c")
    text = open(src).read()
    for sym in header_symbols():
        assert f"F({sym})" in text, f"{sym} is declared in qz_cuda.h but not referenced by tests/abi_smoke.c"
    exe = str(tmp_path / "abi_smoke")
    libdir = os.path.dirname(_lib.LIB_PATH)
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", src, "-I", os.path.join(ROOT, "include"),
                        "-L", libdir, "-lqz_cuda", f"-Wl,-rpath,{libdir}", "-Wl,-rpath,/usr/local/cuda/lib64", "-o", exe],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "abi_smoke:" in r.stdout, r.stdout + r.stderr


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


def test_rust_ffi_declarations_are_current():
    """rust/cuda_ffi.rs is generated from include/qz_cuda.h (tools/gen_rust_ffi.py) and must declare every symbol"""
    import subprocess
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_rust_ffi.py"), "--check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    rs = open(os.path.join(ROOT, "rust", "cuda_ffi.rs")).read()
    for sym in header_symbols():
        assert f"pub fn {sym}(" in rs, sym


def test_binding_pointer_helper_and_output_blocks():
    """ptr() must give the array's own address on the fast (buffer protocol) and the slow (read-only, strided) way;
    the reusable top-k output block is one per thread and per k"""
    import threading

    import numpy as np

    from quantization_b200._index import DeviceIndex
    from quantization_b200.encoded_vectors import ptr

    a = np.arange(64, dtype=np.float32)
    assert ptr(a).value == a.ctypes.data
    assert ptr(a[3:]).value == a[3:].ctypes.data
    assert ptr(a[::2]).value == a[::2].ctypes.data          # not contiguous: slow way, same address
    ro = np.arange(8, dtype=np.uint32)
    ro.setflags(write=False)
    assert ptr(ro).value == ro.ctypes.data                  # read-only: slow way
    assert ptr(None) is None and ptr(np.empty(0, np.uint8)) is None

    class Bare(DeviceIndex):
        def __init__(self):  # no context: only the per-thread block logic is exercised
            self._tls = threading.local()

    ix = Bare()
    mine = ix._topk_out(100)
    assert ix._topk_out(100) is mine and ix._topk_out(10) is not mine
    assert mine[0].dtype == np.float32 and mine[1].dtype == np.uint32 and mine[0].size == mine[1].size == 100
    assert mine[3].value == mine[0].ctypes.data and mine[4].value == mine[1].ctypes.data
    other = []
    t = threading.Thread(target=lambda: other.append(ix._topk_out(100)))
    t.start()
    t.join()
    assert other[0] is not mine and other[0][0] is not mine[0]
