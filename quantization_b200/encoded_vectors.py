"""API types of the reference crate, mirrored (quantization/src/encoded_vectors.rs:6-35,
quantization/src/lib.rs:18-24, quantization/src/encoded_storage.rs:7-70) + the device context."""
import atexit
import ctypes as C
import enum
import sys
import threading
import weakref
from dataclasses import dataclass

import numpy as np

from . import _lib


class DistanceType(enum.IntEnum):
    """encoded_vectors.rs:6-11"""
    Dot = 0
    L1 = 1
    L2 = 2


@dataclass
class VectorParameters:
    """encoded_vectors.rs:13-19"""
    dim: int
    count: int
    distance_type: DistanceType
    invert: bool

    def to_json(self):
        return {"dim": int(self.dim), "count": int(self.count), "distance_type": DistanceType(self.distance_type).name,
                "invert": bool(self.invert)}

    @classmethod
    def from_json(cls, d):
        return cls(int(d["dim"]), int(d["count"]), DistanceType[d["distance_type"]], bool(d["invert"]))


class EncodingError(Exception):
    """lib.rs:18-24: IOError / EncodingError / ArgumentsError / Stopped"""

    def __init__(self, kind, description=""):
        super().__init__(f"{kind}: {description}" if description else kind)
        self.kind = kind
        self.description = description

    def __eq__(self, other):
        return isinstance(other, EncodingError) and self.kind == other.kind

    __hash__ = Exception.__hash__


EncodingError.Stopped = EncodingError("Stopped")

_KIND = {_lib.QZ_ARGUMENTS: "ArgumentsError", _lib.QZ_ENCODING: "EncodingError", _lib.QZ_STOPPED: "Stopped",
         _lib.QZ_IO: "IOError"}


class VecStorageBuilder:
    """`impl EncodedStorageBuilder<Vec<u8>> for Vec<u8>` (encoded_storage.rs:62-70)"""

    def __init__(self):
        self.data = bytearray()

    def push_vector_data(self, other):
        self.data += bytes(other)

    def extend_rows(self, rows):  # bulk form of push_vector_data, same resulting bytes
        self.data += np.ascontiguousarray(rows, np.uint8).tobytes()

    def build(self):
        return np.frombuffer(bytes(self.data), np.uint8)


def get_vector_data(storage, index, vector_size):
    """`EncodedStorage::get_vector_data` for a flat byte array (encoded_storage.rs:28-30)"""
    return storage[vector_size * index: vector_size * (index + 1)]


# Native handles are released in a fixed order at interpreter exit — indexes, then contexts — before the CUDA
# runtime or a borrowed (e.g. torch) stream is torn down; __del__ does nothing once the interpreter is finalizing.
_live_indexes = weakref.WeakSet()
_live_contexts = weakref.WeakSet()


@atexit.register
def _close_all():
    for ix in list(_live_indexes):
        try:
            ix.close()
        except Exception:
            pass
    for c in list(_live_contexts):
        try:
            c.close()
        except Exception:
            pass


class Context:
    """One CUDA device + stream (qz_ctx)."""

    def __init__(self, device=0, stream=None):
        self._L = _lib.lib()
        h = C.c_void_p()
        st = self._L.qz_ctx_create(int(device), C.c_void_p(stream) if stream else None, C.byref(h))
        if st != _lib.QZ_OK:
            raise EncodingError(_KIND.get(st, "EncodingError"), self._L.qz_last_global_error().decode())
        self.handle = h
        self.device = device
        _live_contexts.add(self)

    def check(self, st):
        if st == _lib.QZ_OK:
            return
        if st == _lib.QZ_STOPPED:
            raise EncodingError("Stopped")
        raise EncodingError(_KIND.get(st, "EncodingError"), self._L.qz_last_error(self.handle).decode())

    def synchronize(self):
        self.check(self._L.qz_ctx_synchronize(self.handle))

    def set_timing(self, enabled=True):
        self.check(self._L.qz_ctx_set_timing(self.handle, int(enabled)))

    def get_timing(self):
        """(total ms, launches) of the dominant scan kernel since set_timing"""
        ms, n = C.c_double(0), C.c_uint64(0)
        self.check(self._L.qz_ctx_get_timing(self.handle, C.byref(ms), C.byref(n)))
        return ms.value, int(n.value)

    def set_scan_chaining(self, enabled=True):
        """promise that nothing writing a scan's inputs is enqueued between consecutive *_dev scans on this context's
        (caller-provided) stream, so that they may overlap at the edges (include/qz_cuda.h)"""
        self.check(self._L.qz_ctx_set_scan_chaining(self.handle, int(bool(enabled))))

    def set_phase_stamps(self, d_ptr):
        """device buffer of sm_count x 8 u64 the scan kernels write their phase stamps to (0 / None = off)"""
        self.check(self._L.qz_ctx_set_phase_stamps(self.handle, C.c_void_p(int(d_ptr) if d_ptr else None)))

    @property
    def stream_handle(self):
        """the cudaStream_t (as an int) the context's `*_dev` calls enqueue on"""
        return int(self._L.qz_ctx_stream(self.handle) or 0)

    @property
    def launch_count(self):
        return int(self._L.qz_ctx_launch_count(self.handle))

    @property
    def sm_count(self):
        return int(self._L.qz_ctx_sm_count(self.handle))

    def close(self):
        if self.handle:
            self._L.qz_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        if sys.is_finalizing():
            return
        try:
            self.close()
        except Exception:
            pass


_default_ctx = {}


def default_context(device=0):
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]


class StopFlag:
    """Bridges the reference's `stop_condition: impl Fn() -> bool` to the C-ABI's polled int flag:
    a watcher thread evaluates the callable while the (GIL-releasing) native call runs."""

    def __init__(self, stop_condition, period=0.0005):
        self.fn = stop_condition
        self.flag = C.c_int32(0)
        self.period = period
        self._done = threading.Event()
        self._t = None

    def __enter__(self):
        if self.fn is None:
            return None
        if self.fn():
            self.flag.value = 1
            return C.addressof(self.flag)

        def watch():
            while not self._done.is_set():
                if self.fn():
                    self.flag.value = 1
                    return
                self._done.wait(self.period)

        self._t = threading.Thread(target=watch, daemon=True)
        self._t.start()
        return C.addressof(self.flag)

    def __exit__(self, *a):
        self._done.set()
        if self._t:
            self._t.join()


def as_matrix(orig_data, vector_parameters):
    """Materialise the reference's `orig_data` iterator as a [count, dim] f32 matrix and apply
    validate_vector_parameters (encoded_vectors.rs:47-70)."""
    if isinstance(orig_data, np.ndarray) and orig_data.ndim == 2:
        m = np.ascontiguousarray(orig_data, np.float32)
    else:
        rows = [np.asarray(v, np.float32) for v in orig_data]
        for v in rows:
            if v.shape != (vector_parameters.dim,):
                raise EncodingError("ArgumentsError", f"Vector length {v.size} does not match vector parameters dim "
                                                      f"{vector_parameters.dim}")
        m = np.stack(rows) if rows else np.zeros((0, vector_parameters.dim), np.float32)
    if m.shape[0] != vector_parameters.count:
        raise EncodingError("ArgumentsError", f"Vector count {m.shape[0]} does not match vector parameters count "
                                              f"{vector_parameters.count}")
    if m.shape[0] and m.shape[1] != vector_parameters.dim:
        raise EncodingError("ArgumentsError", f"Vector length {m.shape[1]} does not match vector parameters dim "
                                              f"{vector_parameters.dim}")
    return m


def ptr(a):
    """address of a numpy array's first byte as a c_void_p (None for None / empty).  `a.ctypes.data` builds a helper
    object per call (~2.5 us, more than the rest of a binding call together); the buffer protocol gives the same
    address in a third of that, and only read-only or non-contiguous arrays take the slow way."""
    if a is None or not a.size:
        return None
    try:
        return C.c_void_p(C.addressof(C.c_char.from_buffer(a)))
    except (TypeError, ValueError, BufferError):
        return C.c_void_p(a.ctypes.data)
