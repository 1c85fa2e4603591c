"""EncodedVectorsPQ — host-side mirror of quantization/src/encoded_vectors_pq.rs over libqz_cuda."""
import ctypes as C
import json
import os

import numpy as np

from ._index import DeviceIndex, pick_ctx
from .encoded_vectors import StopFlag, VectorParameters, as_matrix, ptr

KMEANS_SAMPLE_SIZE = 10_000   # pq.rs:22
KMEANS_MAX_ITERATIONS = 100   # pq.rs:23
KMEANS_ACCURACY = 1e-5        # pq.rs:24
CENTROIDS_COUNT = 256         # pq.rs:25


class EncodedVectorsPQ(DeviceIndex):
    def __init__(self, ctx, encoded_vectors, vector_parameters, chunk_size, centroids, id_base=0):
        super().__init__(ctx)
        self.encoded_vectors = encoded_vectors
        self.vector_parameters = vector_parameters
        self.chunk_size = int(chunk_size)
        self.centroids = np.ascontiguousarray(centroids, np.float32).reshape(CENTROIDS_COUNT, vector_parameters.dim)
        self.vector_division = self.get_vector_division(vector_parameters.dim, chunk_size)
        h = C.c_void_p()
        host = np.ascontiguousarray(encoded_vectors, np.uint8) if encoded_vectors is not None else None
        st = self._L.qz_pq_index_create(ctx.handle, ptr(host),
                                        vector_parameters.count, vector_parameters.dim, self.chunk_size,
                                        ptr(self.centroids), int(vector_parameters.distance_type),
                                        int(vector_parameters.invert), id_base, C.byref(h))
        ctx.check(st)
        self.handle = h

    # ---- encode, pq.rs:56-107 ------------------------------------------------------------------------
    @classmethod
    def encode(cls, data, storage_builder, vector_parameters, chunk_size, max_kmeans_threads=1,
               stop_condition=lambda: False, *, ctx=None, sample_ids=None, seed=0, centroids=None, id_base=0):
        """`max_kmeans_threads` is accepted for signature parity; the device trains all chunks at once and
        sums in point order (the reference's result for max_kmeans_threads = 1)."""
        ctx = pick_ctx(ctx)
        L = ctx._L
        m = as_matrix(data, vector_parameters)
        vp = vector_parameters
        n_chunks = cls.get_quantized_vector_size(vp, chunk_size)
        sid = np.ascontiguousarray(sample_ids, np.uint64) if sample_ids is not None else None
        with StopFlag(stop_condition) as flag:
            fp = C.c_void_p(flag) if flag else None
            if centroids is None:
                centroids = np.zeros((CENTROIDS_COUNT, vp.dim), np.float32)
                ctx.check(L.qz_pq_train(ctx.handle, ptr(m), vp.count, vp.dim, chunk_size, ptr(sid) if sid is not None else None,
                                        sid.size if sid is not None else 0, seed, KMEANS_MAX_ITERATIONS,
                                        KMEANS_ACCURACY, fp, ptr(centroids), None))
            centroids = np.ascontiguousarray(centroids, np.float32)
            codes = np.zeros(vp.count * n_chunks, np.uint8)
            ctx.check(L.qz_pq_assign(ctx.handle, ptr(m), vp.count, vp.dim, chunk_size, ptr(centroids), fp, ptr(codes)))
        if hasattr(storage_builder, "extend_rows"):
            storage_builder.extend_rows(codes)
        else:
            for i in range(vp.count):
                storage_builder.push_vector_data(codes[i * n_chunks:(i + 1) * n_chunks])
        return cls(ctx, storage_builder.build(), vp, chunk_size, centroids, id_base)

    @classmethod
    def encode_dev(cls, d_data, vector_parameters, chunk_size, *, ctx=None, sample_ids=None, seed=0, centroids=None,
                   id_base=0):
        """encode() for a [count, dim] f32 matrix resident on the device: k-means and assignment run
        there and the codes are written straight into the new index."""
        ctx = pick_ctx(ctx)
        L = ctx._L
        vp = vector_parameters
        sid = np.ascontiguousarray(sample_ids, np.uint64) if sample_ids is not None else None
        iters = np.zeros(max(cls.get_quantized_vector_size(vp, chunk_size), 1), np.uint32)
        if centroids is None:
            centroids = np.zeros((CENTROIDS_COUNT, vp.dim), np.float32)
            ctx.check(L.qz_pq_train_dev(ctx.handle, C.c_void_p(d_data), vp.count, vp.dim, chunk_size,
                                        ptr(sid) if sid is not None else None, sid.size if sid is not None else 0, seed,
                                        KMEANS_MAX_ITERATIONS, KMEANS_ACCURACY, None, ptr(centroids), ptr(iters)))
        self = cls(ctx, None, vp, chunk_size, centroids, id_base)
        self.kmeans_iterations = iters
        ctx.check(L.qz_pq_assign_dev(ctx.handle, C.c_void_p(d_data), vp.count, vp.dim, chunk_size, ptr(self.centroids),
                                     None, C.c_void_p(self.storage_dev)))
        return self

    @staticmethod
    def get_quantized_vector_size(vector_parameters, chunk_size):
        """pq.rs:109-114"""
        return len(range(0, vector_parameters.dim, chunk_size))

    @staticmethod
    def get_vector_division(dim, chunk_size):
        """pq.rs:116-121"""
        return [(i, min(i + chunk_size, dim)) for i in range(0, dim, chunk_size)]

    # ---- save / load, pq.rs:498-523 ---------------------------------------------------------------
    def save(self, data_path, meta_path):
        meta = {"centroids": [[float(str(x)) for x in row] for row in self.centroids],
                "vector_division": [{"start": a, "end": b} for a, b in self.vector_division],
                "vector_parameters": self.vector_parameters.to_json()}
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
        div = meta["vector_division"]
        size = len(div)  # pq.rs:515
        buf = np.fromfile(data_path, np.uint8)
        expected = size * vector_parameters.count
        if buf.size != expected:
            raise OSError(f"Loaded storage size {buf.size} is not equal to expected size {expected}")
        vp = VectorParameters.from_json(meta["vector_parameters"])
        if (vp.dim, vp.count) != (vector_parameters.dim, vector_parameters.count):
            # the storage was size-checked against the CALLER's parameters; an index built from metadata that
            # disagrees with them would read past the loaded bytes (a stale or foreign meta file)
            raise OSError(f"EncodedVectorsPQ metadata describes {vp.count} x {vp.dim} vectors, the caller asked for "
                          f"{vector_parameters.count} x {vector_parameters.dim}")
        chunk_size = (div[0]["end"] - div[0]["start"]) if div else 1
        cen = np.asarray(meta["centroids"], np.float32).reshape(CENTROIDS_COUNT, vp.dim)
        return cls(ctx, buf, vp, chunk_size, cen, id_base)
