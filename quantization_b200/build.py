"""In-tree build of libqz_cuda.so (sm_100a only).  `python -m quantization_b200.build [--force]`.

The library is the product: hand-written CUDA behind the C-ABI of include/qz_cuda.h.  It is
built with nvcc directly (no torch extension machinery) so that it has no Python or torch
dependency and can be linked from the reference's quantization/build.rs.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libqz_cuda.so")
SOURCES = ["qz_ctx.cu", "qz_api.cu", "row_scan.cu", "row_scan_dot.cu", "row_scan_l1.cu", "row_scan_xor.cu", "point_kernels.cu", "pq.cu", "pq_scan2.cu", "encode.cu", "encode_api.cu", "kmeans.cu", "pq_encode_api.cu", "u8_batch.cu", "peer.cu", "chunked.cu", "no_alloc.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-O2", "--fmad=false", "-Xptxas", "-v",
]
# QZ_TUNING=1: compile the experiment switches (QZ_SCAN_*, QZ_PQ_*, QZ_BATCH_* environment variables read by
# csrc/qz_internal.h tune_u32 / tune_flag) into the library.  The default (production) build has none of them.
TUNING = os.environ.get("QZ_TUNING", "") not in ("", "0")
if TUNING:
    FLAGS.append("-DQZ_TUNING")
STAMP = os.path.join(OBJ, "flags.stamp")


def _deps():
    out = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    out.append(os.path.join(HERE, "..", "include", "qz_cuda.h"))
    return out


def _stamp_matches():
    try:
        return open(STAMP).read() == " ".join(FLAGS)
    except OSError:
        return False


def needs_build():
    if not os.path.exists(LIB) or not _stamp_matches():
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in _deps())


def _headers_mtime():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if not f.endswith(".cu")]
    hs.append(os.path.join(HERE, "..", "include", "qz_cuda.h"))
    return max(os.path.getmtime(h) for h in hs)


def _compile(src, force=False):
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    # incremental: an object newer than its source and every header is reused
    if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(os.path.join(CSRC, src)), _headers_mtime()):
        return obj
    r = subprocess.run([NVCC, *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj], capture_output=True, text=True)
    with open(obj + ".log", "w") as f:
        f.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    if not _stamp_matches():  # objects built with other flags (e.g. a QZ_TUNING build) are not reused
        force = True
    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(lambda f: _compile(f, force), SOURCES))
    r = subprocess.run([NVCC, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
                        "-cudart", "shared", "-Xlinker", "-rpath,/usr/local/cuda/lib64"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(STAMP, "w") as f:
        f.write(" ".join(FLAGS))
    if verbose:
        for o in objs:
            print(open(o + ".log").read())
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
