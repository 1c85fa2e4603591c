"""EncodedVectorsU8 — host-side mirror of quantization/src/encoded_vectors_u8.rs over libqz_cuda.

Same names, argument meaning and error behaviour as the reference; every arithmetic step runs
in the CUDA library (no CPU fallback)."""
import ctypes as C
import json
import os

import numpy as np

from . import _lib
from ._index import DeviceIndex, pick_ctx
from .encoded_vectors import (DistanceType, EncodingError, StopFlag, VecStorageBuilder, VectorParameters, as_matrix,
                              ptr)

ALIGNMENT = 16  # u8.rs:12


def _f32_json(x):
    return float(str(np.float32(x)))  # shortest round-trip repr, like serde_json/ryu for f32


class EncodedVectorsU8(DeviceIndex):
    def __init__(self, ctx, encoded_vectors, vector_parameters, alpha, offset, multiplier, id_base=0):
        super().__init__(ctx)
        self.encoded_vectors = encoded_vectors  # the host TStorage (flat bytes), still owned by the caller
        self.vector_parameters = vector_parameters
        self.actual_dim = self.get_actual_dim(vector_parameters)
        self.alpha, self.offset, self.multiplier = np.float32(alpha), np.float32(offset), np.float32(multiplier)
        h = C.c_void_p()
        host = np.ascontiguousarray(encoded_vectors, np.uint8) if encoded_vectors is not None else None
        st = self._L.qz_u8_index_create(ctx.handle, ptr(host),
                                        vector_parameters.count, vector_parameters.dim,
                                        int(vector_parameters.distance_type), int(vector_parameters.invert),
                                        float(alpha), float(offset), float(multiplier), id_base, C.byref(h))
        ctx.check(st)
        self.handle = h

    # ---- encode, u8.rs:34-140 -------------------------------------------------------------------
    @classmethod
    def encode(cls, orig_data, storage_builder, vector_parameters, quantile=None, stop_condition=lambda: False, *,
               ctx=None, sample_ids=None, seed=0, id_base=0):
        ctx = pick_ctx(ctx)
        L = ctx._L
        data = as_matrix(orig_data, vector_parameters)
        row = cls.get_quantized_vector_size(vector_parameters)
        out = np.empty(vector_parameters.count * row, np.uint8)
        a, o, m = C.c_float(), C.c_float(), C.c_float()
        q = C.byref(C.c_float(quantile)) if quantile is not None else None
        sid = np.ascontiguousarray(sample_ids, np.uint64) if sample_ids is not None else None
        with StopFlag(stop_condition) as flag:
            st = L.qz_u8_encode(ctx.handle, ptr(data), vector_parameters.count, vector_parameters.dim,
                                int(vector_parameters.distance_type), int(vector_parameters.invert),
                                C.cast(q, _lib.f32p) if q is not None else None, ptr(sid) if sid is not None else None,
                                sid.size if sid is not None else 0, seed, C.c_void_p(flag) if flag else None, ptr(out),
                                C.byref(a), C.byref(o), C.byref(m))
        ctx.check(st)
        if hasattr(storage_builder, "extend_rows"):
            storage_builder.extend_rows(out)
        else:
            for i in range(vector_parameters.count):
                storage_builder.push_vector_data(out[i * row:(i + 1) * row])
        return cls(ctx, storage_builder.build(), vector_parameters, a.value, o.value, m.value, id_base)

    @classmethod
    def encode_dev(cls, d_data, vector_parameters, quantile=None, *, ctx=None, sample_ids=None, seed=0, id_base=0):
        """encode() for a [count, dim] f32 matrix already resident on the device (`d_data` = device
        pointer): the codes are written straight into the new index, nothing crosses PCIe."""
        ctx = pick_ctx(ctx)
        L = ctx._L
        self = cls(ctx, None, vector_parameters, 0.0, 0.0, 0.0, id_base)
        a, o, m = C.c_float(), C.c_float(), C.c_float()
        q = C.byref(C.c_float(quantile)) if quantile is not None else None
        sid = np.ascontiguousarray(sample_ids, np.uint64) if sample_ids is not None else None
        st = L.qz_u8_encode_dev(ctx.handle, C.c_void_p(d_data), vector_parameters.count, vector_parameters.dim,
                                int(vector_parameters.distance_type), int(vector_parameters.invert),
                                C.cast(q, _lib.f32p) if q is not None else None, ptr(sid) if sid is not None else None,
                                sid.size if sid is not None else 0, seed, None, C.c_void_p(self.storage_dev),
                                C.byref(a), C.byref(o), C.byref(m))
        ctx.check(st)
        ctx.check(L.qz_u8_index_set_params(self.handle, a.value, o.value, m.value))
        self.alpha, self.offset, self.multiplier = np.float32(a.value), np.float32(o.value), np.float32(m.value)
        return self

    # ---- batched multi-query search on the tensor cores (BASELINE config 5) ------------------------
    def search_batch(self, queries, k, largest=True):
        """f32 queries [nq, dim] -> (scores [nq, k], ids [nq, k]); per query identical to
        search_top_k(encode_query(q), k, largest) (slots past `count`: NaN / 0xFFFFFFFF)"""
        q = np.ascontiguousarray(queries, np.float32)
        if q.ndim != 2 or q.shape[1] != self.vector_parameters.dim:
            raise EncodingError("ArgumentsError", f"queries must be [nq, {self.vector_parameters.dim}]")
        nq = q.shape[0]
        scores = np.empty((nq, k), np.float32)
        ids = np.empty((nq, k), np.uint32)
        self.ctx.check(self._L.qz_u8_search_batch(self.handle, ptr(q), nq, k, int(bool(largest)), ptr(scores), ptr(ids)))
        return scores, ids

    def search_batch_dev(self, d_queries, nq, k, largest, d_scores, d_ids):
        self.ctx.check(self._L.qz_u8_search_batch_dev(self.handle, C.c_void_p(d_queries), nq, k, int(bool(largest)),
                                                      C.c_void_p(d_scores), C.c_void_p(d_ids)))

    @staticmethod
    def get_quantized_vector_size(vector_parameters):
        """u8.rs:252-255"""
        return EncodedVectorsU8.get_actual_dim(vector_parameters) + 4

    @staticmethod
    def get_actual_dim(vector_parameters):
        """u8.rs:257-259"""
        return vector_parameters.dim + (ALIGNMENT - vector_parameters.dim % ALIGNMENT) % ALIGNMENT

    # ---- save / load, u8.rs:263-288 (same JSON field names as the serde Metadata) -----------------
    def save(self, data_path, meta_path):
        meta = {"actual_dim": int(self.actual_dim), "alpha": _f32_json(self.alpha), "offset": _f32_json(self.offset),
                "multiplier": _f32_json(self.multiplier), "vector_parameters": self.vector_parameters.to_json()}
        for p in (meta_path, data_path):
            d = os.path.dirname(p)
            if d:
                os.makedirs(d, exist_ok=True)
        with open(meta_path, "w") as f:
            json.dump(meta, f, separators=(",", ":"))
        self.host_storage().tofile(data_path)

    @classmethod
    def load(cls, data_path, meta_path, vector_parameters, *, ctx=None, id_base=0):
        ctx = pick_ctx(ctx)
        with open(meta_path) as f:
            meta = json.load(f)
        size = cls.get_quantized_vector_size(vector_parameters)  # sized from the CALLER's parameters, u8.rs:280-282
        buf = np.fromfile(data_path, np.uint8)
        expected = size * vector_parameters.count
        if buf.size != expected:  # encoded_storage.rs:40-50
            raise OSError(f"Loaded storage size {buf.size} is not equal to expected size {expected}")
        vp = VectorParameters.from_json(meta["vector_parameters"])
        if (vp.dim, vp.count) != (vector_parameters.dim, vector_parameters.count):
            # the storage was size-checked against the CALLER's parameters; an index built from metadata that
            # disagrees with them would read past the loaded bytes (a stale or foreign meta file)
            raise OSError(f"EncodedVectorsU8 metadata describes {vp.count} x {vp.dim} vectors, the caller asked for "
                          f"{vector_parameters.count} x {vector_parameters.dim}")
        return cls(ctx, buf, vp, meta["alpha"], meta["offset"], meta["multiplier"], id_base)
