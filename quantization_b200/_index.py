"""Shared behaviour of the three device-resident codecs: the `EncodedVectors` trait methods
(encoded_vectors.rs:21-35) that are codec-independent once an index handle exists, plus the
fused scan entry points that only make sense on the GPU (score_all / search_top_k)."""
import ctypes as C
import sys
import threading

import numpy as np

from . import _lib
from .encoded_vectors import Context, EncodingError, _live_indexes, default_context, ptr


class DeviceIndex:
    """Owns a qz_index handle; subclasses create it."""

    def __init__(self, ctx: Context):
        self.ctx = ctx
        self._L = ctx._L
        self.handle = None
        self._tls = threading.local()
        _live_indexes.add(self)

    def _topk_out(self, k):
        """this thread's reusable output block of a host-facing top-k call, (scores, ids, n, &scores, &ids, &n), with
        the addresses resolved once: two np.empty + two pointer look-ups per call cost more than the library's own
        host code.  Callers return copies.  Per thread: the scoring entry points may be called from any thread."""
        try:
            return self._tls.out[k]
        except AttributeError:
            self._tls.out = {}
        except KeyError:
            pass
        scores, ids, n = np.empty(k, np.float32), np.empty(k, np.uint32), C.c_uint32(0)
        blk = (scores, ids, n, ptr(scores), ptr(ids), C.byref(n))
        self._tls.out[k] = blk
        return blk

    # ---- trait: encode_query / score_point / score_internal -------------------------------
    def encode_query(self, query):
        q = np.ascontiguousarray(query, np.float32).reshape(-1)
        if q.size != self.vector_parameters.dim:
            raise EncodingError("ArgumentsError", f"query length {q.size} != dim {self.vector_parameters.dim}")
        out = np.zeros(max(int(self._L.qz_query_bytes(self.handle)), 1), np.uint8)
        self.ctx.check(self._L.qz_encode_query(self.handle, ptr(q), ptr(out)))
        return out[: int(self._L.qz_query_bytes(self.handle))]

    def score_point(self, query, i):
        return float(self.score_points(query, [i])[0])

    def score_internal(self, i, j):
        return float(self.score_pairs([i], [j])[0])

    # ---- batched forms of the same ---------------------------------------------------------
    def score_points(self, query, rows):
        rows = np.ascontiguousarray(rows, np.uint32)
        out = np.empty(rows.size, np.float32)
        q = np.ascontiguousarray(query, np.uint8)
        self.ctx.check(self._L.qz_score_points(self.handle, ptr(q), ptr(rows), rows.size, ptr(out)))
        return out

    def score_pairs(self, rows_a, rows_b):
        a = np.ascontiguousarray(rows_a, np.uint32)
        b = np.ascontiguousarray(rows_b, np.uint32)
        if a.size != b.size:
            raise EncodingError("ArgumentsError", "pair lists differ in length")
        out = np.empty(a.size, np.float32)
        self.ctx.check(self._L.qz_score_pairs(self.handle, ptr(a), ptr(b), a.size, ptr(out)))
        return out

    # ---- the fused scans ---------------------------------------------------------------------
    def score_all(self, query):
        """score_point(query, i) for every i"""
        out = np.empty(self.count, np.float32)
        q = np.ascontiguousarray(query, np.uint8)
        self.ctx.check(self._L.qz_scan_scores(self.handle, ptr(q), ptr(out)))
        return out

    def search_top_k(self, query, k, largest=True):
        """(scores, ids) of the k best points, best first; ties go to the lower id"""
        q = np.ascontiguousarray(query, np.uint8)
        scores = np.empty(k, np.float32)
        ids = np.empty(k, np.uint32)
        n = C.c_uint32(0)
        self.ctx.check(self._L.qz_scan_topk(self.handle, ptr(q), k, int(bool(largest)), ptr(scores), ptr(ids),
                                            C.byref(n)))
        return scores[: n.value].copy(), ids[: n.value].copy()

    def search(self, query_f32, k, largest=True):
        """encode_query + search_top_k in one library call (one synchronisation): raw f32 query in, (scores, ids) out"""
        q = np.ascontiguousarray(query_f32, np.float32).reshape(-1)
        if q.size != self.vector_parameters.dim:
            raise EncodingError("ArgumentsError", f"query length {q.size} != dim {self.vector_parameters.dim}")
        scores, ids, n, p_scores, p_ids, p_n = self._topk_out(k)
        self.ctx.check(self._L.qz_search_topk(self.handle, ptr(q), k, int(bool(largest)), p_scores, p_ids, p_n))
        return scores[: n.value].copy(), ids[: n.value].copy()

    def search_top_k_chunked(self, other, host_storage, count, query, k, largest=True):
        """out-of-core scan (qz_scan_topk_chunked): `self` and `other` are two buffer indexes of the same shape
        (storage None, count = chunk capacity); host_storage = the flat encoded bytes of `count` rows"""
        st = np.ascontiguousarray(host_storage, np.uint8)
        q = np.ascontiguousarray(query, np.uint8)
        scores, ids, n = np.empty(k, np.float32), np.empty(k, np.uint32), C.c_uint32(0)
        self.ctx.check(self._L.qz_scan_topk_chunked(self.handle, other.handle, ptr(st), count, ptr(q), k,
                                                    int(bool(largest)), ptr(scores), ptr(ids), C.byref(n)))
        return scores[: n.value].copy(), ids[: n.value].copy()

    # ---- device-pointer forms (asynchronous on the context stream) -------------------------
    def search_top_k_dev(self, d_query, k, largest, d_keys=0, d_scores=0, d_ids=0):
        self.ctx.check(self._L.qz_scan_topk_dev(self.handle, C.c_void_p(d_query), k, int(bool(largest)),
                                                C.c_void_p(d_keys), C.c_void_p(d_scores), C.c_void_p(d_ids)))

    def score_points_dev(self, d_query, d_rows, n, d_scores):
        self.ctx.check(self._L.qz_score_points_dev(self.handle, C.c_void_p(d_query), C.c_void_p(d_rows), n,
                                                   C.c_void_p(d_scores)))

    def score_pairs_dev(self, d_rows_a, d_rows_b, n, d_scores):
        self.ctx.check(self._L.qz_score_pairs_dev(self.handle, C.c_void_p(d_rows_a), C.c_void_p(d_rows_b), n,
                                                  C.c_void_p(d_scores)))

    def score_all_dev(self, d_query, d_scores):
        self.ctx.check(self._L.qz_scan_scores_dev(self.handle, C.c_void_p(d_query), C.c_void_p(d_scores)))

    def encode_query_dev(self, d_queries, nq, d_out):
        self.ctx.check(self._L.qz_encode_query_dev(self.handle, C.c_void_p(d_queries), nq, C.c_void_p(d_out)))

    @property
    def count(self):
        return int(self._L.qz_index_count(self.handle))

    @property
    def query_bytes(self):
        return int(self._L.qz_query_bytes(self.handle))

    @property
    def row_bytes(self):
        return int(self._L.qz_index_row_bytes(self.handle))

    @property
    def storage_dev(self):
        return int(self._L.qz_index_storage_dev(self.handle) or 0)

    def host_storage(self):
        """the flat encoded bytes on the host (EncodedStorage): the caller's array when the index was made from one,
        else (device-side encode, sharded encode) a copy fetched from the device once and kept"""
        if self.encoded_vectors is None:
            self.encoded_vectors = self.download()
        return np.ascontiguousarray(self.encoded_vectors, np.uint8).reshape(-1)

    def download(self):
        out = np.empty(self.count * self.row_bytes, np.uint8)
        if out.size:
            self.ctx.check(self._L.qz_index_download(self.handle, ptr(out)))
        return out

    def close(self):
        if self.handle:
            self._L.qz_index_destroy(self.handle)
            self.handle = None

    def __del__(self):
        if sys.is_finalizing():
            return
        try:
            self.close()
        except Exception:
            pass


def pick_ctx(ctx):
    return ctx if ctx is not None else default_context(0)
