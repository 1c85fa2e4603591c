"""EncodedVectorsBin — host-side mirror of quantization/src/encoded_vectors_binary.rs over libqz_cuda."""
import ctypes as C
import json
import os

import numpy as np

from ._index import DeviceIndex, pick_ctx
from .encoded_vectors import StopFlag, VectorParameters, as_matrix, ptr

# BitsStoreType (binary.rs:26-42): the reference is generic over u8 and u128 words
BITS_STORE_U8 = 1
BITS_STORE_U128 = 16


class EncodedVectorsBin(DeviceIndex):
    def __init__(self, ctx, encoded_vectors, vector_parameters, bits_store_type=BITS_STORE_U128, id_base=0):
        super().__init__(ctx)
        self.encoded_vectors = encoded_vectors
        self.vector_parameters = vector_parameters
        self.bits_store_type = bits_store_type
        h = C.c_void_p()
        host = np.ascontiguousarray(encoded_vectors, np.uint8) if encoded_vectors is not None else None
        st = self._L.qz_bin_index_create(ctx.handle, ptr(host),
                                         vector_parameters.count, vector_parameters.dim, bits_store_type,
                                         int(vector_parameters.distance_type), int(vector_parameters.invert), id_base,
                                         C.byref(h))
        ctx.check(st)
        self.handle = h

    # ---- encode, binary.rs:165-191 ------------------------------------------------------------------
    @classmethod
    def encode(cls, orig_data, storage_builder, vector_parameters, stop_condition=lambda: False, *,
               bits_store_type=BITS_STORE_U128, ctx=None, id_base=0):
        ctx = pick_ctx(ctx)
        L = ctx._L
        data = as_matrix(orig_data, vector_parameters)
        row = cls.get_quantized_vector_size_from_params(vector_parameters, bits_store_type, ctx)
        out = np.zeros(vector_parameters.count * row, np.uint8)
        with StopFlag(stop_condition) as flag:
            st = L.qz_bin_encode(ctx.handle, ptr(data), vector_parameters.count, vector_parameters.dim,
                                 bits_store_type, C.c_void_p(flag) if flag else None, ptr(out))
        ctx.check(st)
        if hasattr(storage_builder, "extend_rows"):
            storage_builder.extend_rows(out)
        else:
            for i in range(vector_parameters.count):
                storage_builder.push_vector_data(out[i * row:(i + 1) * row])
        return cls(ctx, storage_builder.build(), vector_parameters, bits_store_type, id_base)

    @classmethod
    def empty(cls, vector_parameters, *, bits_store_type=BITS_STORE_U128, ctx=None, id_base=0):
        """an index whose (zeroed) device storage is to be filled with encode_rows_dev"""
        return cls(pick_ctx(ctx), None, vector_parameters, bits_store_type, id_base)

    def encode_rows_dev(self, d_data, first_row, n_rows):
        """encode n_rows device-resident f32 rows into rows [first_row, first_row + n_rows) of this index"""
        dst = self.storage_dev + first_row * self.row_bytes
        self.ctx.check(self._L.qz_bin_encode_dev(self.ctx.handle, C.c_void_p(d_data), n_rows,
                                                 self.vector_parameters.dim, self.bits_store_type, None,
                                                 C.c_void_p(dst)))

    @staticmethod
    def get_quantized_vector_size_from_params(vector_parameters, bits_store_type=BITS_STORE_U128, ctx=None):
        """binary.rs:210-213"""
        return int(pick_ctx(ctx)._L.qz_bin_row_bytes(vector_parameters.dim, bits_store_type))

    # ---- save / load, binary.rs:260-286 ---------------------------------------------------------------
    def save(self, data_path, meta_path):
        for p in (meta_path, data_path):
            d = os.path.dirname(p)
            if d:
                os.makedirs(d, exist_ok=True)
        with open(meta_path, "w") as f:
            json.dump({"vector_parameters": self.vector_parameters.to_json()}, f, separators=(",", ":"))
        self.host_storage().tofile(data_path)

    @classmethod
    def load(cls, data_path, meta_path, vector_parameters, *, bits_store_type=BITS_STORE_U128, ctx=None, id_base=0):
        ctx = pick_ctx(ctx)
        with open(meta_path) as f:
            meta = json.load(f)
        size = cls.get_quantized_vector_size_from_params(vector_parameters, bits_store_type, ctx)
        buf = np.fromfile(data_path, np.uint8)
        expected = size * vector_parameters.count
        if buf.size != expected:
            raise OSError(f"Loaded storage size {buf.size} is not equal to expected size {expected}")
        vp = VectorParameters.from_json(meta["vector_parameters"])
        if (vp.dim, vp.count) != (vector_parameters.dim, vector_parameters.count):
            raise OSError(f"EncodedVectorsBin metadata describes {vp.count} x {vp.dim} vectors, the caller asked for "
                          f"{vector_parameters.count} x {vector_parameters.dim}")
        return cls(ctx, buf, vp, bits_store_type, id_base)
