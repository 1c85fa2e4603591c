"""ctypes wrapper over oracle/liboracle.so and oracle/_ref/libsimd_utils_ref.so.

TEST INFRASTRUCTURE ONLY — never imported by quantization_b200/.
Function names follow the reference (encode, encode_query, score_point, score_internal);
each wrapper's docstring cites the reference lines the C function restates.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
DOT, L1, L2 = 0, 1, 2
DIST = {"Dot": DOT, "L1": L1, "L2": L2, DOT: DOT, L1: L1, L2: L2}

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)
_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)


def build(force=False):
    """(Re)build liboracle.so and, when /root/reference is present, _ref/."""
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "qz_oracle.c")
    ref_so = os.path.join(_HERE, "_ref", "libsimd_utils_ref.so")
    need = force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src)
    need = need or (not os.path.exists(ref_so) and os.path.exists("/root/reference/quantization/cpp/avx2.c"))
    if need:
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return so


def _p(a, t):
    return a.ctypes.data_as(t)


def _load():
    lib = C.CDLL(build())
    lib.qzo_u8_actual_dim.restype = C.c_uint32
    lib.qzo_u8_multiplier.restype = C.c_float
    lib.qzo_u8_multiplier.argtypes = [C.c_float, C.c_int, C.c_int]
    lib.qzo_u8_code.restype = C.c_uint8
    lib.qzo_u8_code.argtypes = [C.c_float, C.c_float, C.c_float]
    lib.qzo_score_dot.restype = C.c_float
    lib.qzo_score_l1.restype = C.c_float
    lib.qzo_bin_row_bytes.restype = C.c_uint32
    lib.qzo_bin_xor_popcnt.restype = C.c_uint32
    lib.qzo_bin_metric.restype = C.c_float
    lib.qzo_pq_division.restype = C.c_uint32
    lib.qzo_topk.restype = C.c_uint32
    lib.qzo_merge_topk.restype = C.c_uint32
    lib.qzo_kmeans.restype = C.c_uint32
    lib.qzo_make_key.restype = C.c_uint64
    lib.qzo_make_key.argtypes = [C.c_float, C.c_uint32, C.c_int]
    for n in ("qzo_u8_scan_topk_mt", "qzo_bin_scan_topk_mt", "qzo_pq_scan_topk_mt"):
        getattr(lib, n).restype = C.c_uint32
    return lib


_lib = _load()


class _Ref:
    """The reference's own C scorers (cpp/sse.c, cpp/avx2.c), compiled unmodified."""

    def __init__(self):
        path = os.path.join(_HERE, "_ref", "libsimd_utils_ref.so")
        self.available = os.path.exists(path)
        self.lib = C.CDLL(path) if self.available else None
        if self.available:
            for n in ("impl_score_dot_avx", "impl_score_l1_avx", "impl_score_dot_sse", "impl_score_l1_sse"):
                f = getattr(self.lib, n)
                f.restype = C.c_float
                f.argtypes = [_u8p, _u8p, C.c_uint32]
            for n in ("impl_xor_popcnt_sse_uint128", "impl_xor_popcnt_sse_uint64", "impl_xor_popcnt_sse_uint32"):
                f = getattr(self.lib, n)
                f.restype = C.c_uint32
                f.argtypes = [_u8p, _u8p, C.c_uint32]

    def fn(self, name):
        """raw function pointer (void*) usable as the `kernel` argument of the scan loops"""
        return C.cast(getattr(self.lib, name), C.c_void_p) if self.available else None

    def call_u8(self, name, q, v):
        q = np.ascontiguousarray(q, np.uint8)
        v = np.ascontiguousarray(v, np.uint8)
        return float(getattr(self.lib, name)(_p(q, _u8p), _p(v, _u8p), C.c_uint32(q.size)))

    def call_popcnt(self, name, a, b, count):
        a = np.ascontiguousarray(a, np.uint8)
        b = np.ascontiguousarray(b, np.uint8)
        return int(getattr(self.lib, name)(_p(a, _u8p), _p(b, _u8p), C.c_uint32(count)))


ref = _Ref()


def u8_kernel(distance, which="avx"):
    """integer scorer used by x86 score_point (encoded_vectors_u8.rs:335-349 -> cpp/avx2.c);
    None (= scalar restatement of u8.rs:456-474) when oracle/_ref is unavailable."""
    if which == "scalar" or not ref.available:
        return None
    return ref.fn(("impl_score_l1_" if DIST[distance] == L1 else "impl_score_dot_") + which)


# ---------------------------------------------------------------- synthetic data
def uniform(seed, n, lo=0.0, hi=1.0, first=0):
    out = np.empty(n, np.float32)
    _lib.qzo_fill_uniform(C.c_uint64(seed), C.c_uint64(first), C.c_uint64(n), C.c_float(lo), C.c_float(hi), _p(out, _f32p))
    return out


def signs(seed, n, first=0):
    out = np.empty(n, np.float32)
    _lib.qzo_fill_signs(C.c_uint64(seed), C.c_uint64(first), C.c_uint64(n), _p(out, _f32p))
    return out


# ---------------------------------------------------------------- ordering / top-k
def make_key(score, idx, largest):
    return int(_lib.qzo_make_key(C.c_float(score), C.c_uint32(idx), C.c_int(int(largest))))


def topk(scores, k, largest, id_base=0):
    scores = np.ascontiguousarray(scores, np.float32)
    os_, oi = np.empty(k, np.float32), np.empty(k, np.uint32)
    m = _lib.qzo_topk(_p(scores, _f32p), C.c_uint64(scores.size), C.c_uint32(k), C.c_int(int(largest)),
                      C.c_uint32(id_base), _p(os_, _f32p), _p(oi, _u32p))
    return os_[:m].copy(), oi[:m].copy()


def merge_topk(scores, ids, k, largest):
    scores = np.ascontiguousarray(scores, np.float32)
    ids = np.ascontiguousarray(ids, np.uint32)
    os_, oi = np.empty(k, np.float32), np.empty(k, np.uint32)
    m = _lib.qzo_merge_topk(_p(scores, _f32p), _p(ids, _u32p), C.c_uint64(scores.size), C.c_uint32(k),
                            C.c_int(int(largest)), _p(os_, _f32p), _p(oi, _u32p))
    return os_[:m].copy(), oi[:m].copy()


# ---------------------------------------------------------------- u8 scalar quantizer
def u8_actual_dim(dim):
    return int(_lib.qzo_u8_actual_dim(C.c_uint32(dim)))


def minmax(data):
    """quantile.rs:5-19"""
    data = np.ascontiguousarray(data, np.float32).reshape(-1)
    mn, mx = C.c_float(), C.c_float()
    _lib.qzo_minmax(_p(data, _f32p), C.c_uint64(data.size), C.byref(mn), C.byref(mx))
    return mn.value, mx.value


def quantile_interval(sample_values, count, slice_size, quantile):
    """quantile.rs:21-71 on an explicit sample; returns (lo, hi) or None"""
    v = np.array(sample_values, np.float32).reshape(-1)
    lo, hi = C.c_float(), C.c_float()
    ok = _lib.qzo_quantile_interval(_p(v, _f32p), C.c_uint64(v.size), C.c_uint64(count), C.c_uint64(slice_size),
                                    C.c_float(quantile), C.byref(lo), C.byref(hi))
    return (lo.value, hi.value) if ok else None


class OracleU8:
    """EncodedVectorsU8 (encoded_vectors_u8.rs:14-453) restated on the CPU."""

    def __init__(self, storage, count, dim, distance, invert, alpha, offset, multiplier):
        self.storage = np.ascontiguousarray(storage, np.uint8).reshape(-1)
        self.count, self.dim = int(count), int(dim)
        self.distance, self.invert = DIST[distance], bool(invert)
        self.alpha, self.offset, self.multiplier = np.float32(alpha), np.float32(offset), np.float32(multiplier)
        self.actual_dim = u8_actual_dim(dim)
        self.stride = self.actual_dim + 4

    @classmethod
    def encode(cls, data, distance, invert, quantile=None, sample_ids=None):
        """encode, u8.rs:34-140.  `sample_ids` fixes the quantile sample (random in the reference,
        quantile.rs:32-34); default = all ids when count <= 100_000, else the first 100_000."""
        data = np.ascontiguousarray(data, np.float32)
        count, dim = data.shape
        d = DIST[distance]
        ad = u8_actual_dim(dim)
        if count == 0:  # u8.rs:43-54
            return cls(np.zeros(0, np.uint8), 0, dim, d, invert, 0.0, 0.0, 0.0)
        mn, mx = minmax(data)
        if quantile is not None:
            slice_size = min(count, 100_000)
            ids = np.arange(slice_size) if sample_ids is None else np.sort(np.asarray(sample_ids))
            iv = quantile_interval(data[ids], count, slice_size, quantile)
            if iv is not None:
                mn, mx = iv
        alpha, offset = C.c_float(), C.c_float()
        _lib.qzo_u8_alpha_offset(C.c_float(mn), C.c_float(mx), C.byref(alpha), C.byref(offset))
        storage = np.empty(count * (ad + 4), np.uint8)
        _lib.qzo_u8_encode(_p(data, _f32p), C.c_uint64(count), C.c_uint32(dim), C.c_int(d), C.c_int(int(invert)),
                           alpha, offset, _p(storage, _u8p))
        mult = _lib.qzo_u8_multiplier(alpha, C.c_int(d), C.c_int(int(invert)))
        return cls(storage, count, dim, d, invert, alpha.value, offset.value, mult)

    def encode_query(self, q):
        """u8.rs:290-329 -> bytes [f32 offset][actual_dim codes]"""
        q = np.ascontiguousarray(q, np.float32)
        out = np.empty(self.stride, np.uint8)
        _lib.qzo_u8_encode_query(_p(q, _f32p), C.c_uint32(self.dim), C.c_int(self.distance), C.c_int(int(self.invert)),
                                 C.c_float(self.alpha), C.c_float(self.offset), _p(out, _u8p))
        return out

    def scan(self, enc_query, first=0, n=None, kernel="avx"):
        """score_point(q, i) for i in [first, first+n), u8.rs:331-384"""
        n = self.count - first if n is None else n
        out = np.empty(n, np.float32)
        _lib.qzo_u8_scan(_p(self.storage, _u8p), C.c_uint64(first), C.c_uint64(n), C.c_uint32(self.actual_dim),
                         C.c_int(self.distance), C.c_float(self.multiplier), _p(enc_query, _u8p),
                         u8_kernel(self.distance, kernel), _p(out, _f32p))
        return out

    def score_point(self, enc_query, i, kernel="avx"):
        return float(self.scan(enc_query, i, 1, kernel)[0])

    def score_pairs(self, ia, ib, kernel="avx"):
        """score_internal, u8.rs:386-453"""
        ia = np.ascontiguousarray(ia, np.uint32)
        ib = np.ascontiguousarray(ib, np.uint32)
        out = np.empty(ia.size, np.float32)
        _lib.qzo_u8_score_pairs(_p(self.storage, _u8p), C.c_uint32(self.actual_dim), C.c_int(self.distance),
                                C.c_int(int(self.invert)), C.c_float(self.multiplier), C.c_float(self.offset),
                                _p(ia, _u32p), _p(ib, _u32p), C.c_uint64(ia.size), u8_kernel(self.distance, kernel),
                                _p(out, _f32p))
        return out

    def score_internal(self, i, j, kernel="avx"):
        return float(self.score_pairs([i], [j], kernel)[0])

    def search(self, enc_query, k, largest, id_base=0, threads=1, kernel="avx"):
        """the caller-side scan + top-k (demos/src/ann_benchmark_data.rs:150-167)"""
        os_, oi = np.empty(k, np.float32), np.empty(k, np.uint32)
        m = _lib.qzo_u8_scan_topk_mt(_p(self.storage, _u8p), C.c_uint64(self.count), C.c_uint32(self.actual_dim),
                                     C.c_int(self.distance), C.c_float(self.multiplier), _p(enc_query, _u8p),
                                     u8_kernel(self.distance, kernel), C.c_uint32(k), C.c_int(int(largest)),
                                     C.c_uint32(id_base), C.c_int(threads), _p(os_, _f32p), _p(oi, _u32p))
        return os_[:m].copy(), oi[:m].copy()


def score_dot_scalar(q, v):
    q = np.ascontiguousarray(q, np.uint8); v = np.ascontiguousarray(v, np.uint8)
    return float(_lib.qzo_score_dot(_p(q, _u8p), _p(v, _u8p), C.c_uint32(q.size)))


def score_l1_scalar(q, v):
    q = np.ascontiguousarray(q, np.uint8); v = np.ascontiguousarray(v, np.uint8)
    return float(_lib.qzo_score_l1(_p(q, _u8p), _p(v, _u8p), C.c_uint32(q.size)))


# ---------------------------------------------------------------- binary quantizer
def bin_row_bytes(dim, store=16):
    """binary.rs:99-116 (u8 store -> store=1), :152-159 (u128 store -> store=16)"""
    return int(_lib.qzo_bin_row_bytes(C.c_uint32(dim), C.c_int(store)))


def bin_xor_popcnt(a, b):
    """xor_popcnt over two equal-length byte rows (binary.rs:44-97, 119-150), scalar restatement"""
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return int(_lib.qzo_bin_xor_popcnt(_p(a, _u8p), _p(b, _u8p), C.c_uint32(a.size)))


class OracleBin:
    """EncodedVectorsBin (encoded_vectors_binary.rs:11-315) restated on the CPU."""

    def __init__(self, storage, count, dim, store, distance, invert):
        self.storage = np.ascontiguousarray(storage, np.uint8).reshape(-1)
        self.count, self.dim, self.store = int(count), int(dim), int(store)
        self.distance, self.invert = DIST[distance], bool(invert)
        self.row_bytes = bin_row_bytes(dim, store)

    @classmethod
    def encode(cls, data, distance, invert, store=16):
        """binary.rs:165-208"""
        data = np.ascontiguousarray(data, np.float32)
        count, dim = data.shape
        rb = bin_row_bytes(dim, store)
        storage = np.zeros(count * rb, np.uint8)
        _lib.qzo_bin_encode(_p(data, _f32p), C.c_uint64(count), C.c_uint32(dim), C.c_int(store), _p(storage, _u8p))
        return cls(storage, count, dim, store, distance, invert)

    def encode_query(self, q):
        """binary.rs:288-291"""
        q = np.ascontiguousarray(q, np.float32).reshape(1, -1)
        out = np.zeros(max(self.row_bytes, 1), np.uint8)
        _lib.qzo_bin_encode(_p(q, _f32p), C.c_uint64(1), C.c_uint32(self.dim), C.c_int(self.store), _p(out, _u8p))
        return out[: self.row_bytes]

    def _kernels(self, use_ref):
        if use_ref and ref.available:
            return (ref.fn("impl_xor_popcnt_sse_uint128"), ref.fn("impl_xor_popcnt_sse_uint64"),
                    ref.fn("impl_xor_popcnt_sse_uint32"))
        return (None, None, None)

    def scan(self, enc_query, first=0, n=None, use_ref=True):
        """score_point, binary.rs:293-300 with calculate_metric :237-252"""
        n = self.count - first if n is None else n
        out = np.empty(n, np.float32)
        eq = np.ascontiguousarray(np.concatenate([enc_query, np.zeros(16, np.uint8)]))
        k = self._kernels(use_ref)
        _lib.qzo_bin_scan(_p(self.storage, _u8p), C.c_uint64(first), C.c_uint64(n), C.c_uint32(self.dim),
                          C.c_int(self.store), C.c_int(self.distance), C.c_int(int(self.invert)), _p(eq, _u8p),
                          k[0], k[1], k[2], _p(out, _f32p))
        return out

    def score_pairs(self, ia, ib):
        """score_internal, binary.rs:302-314"""
        ia = np.ascontiguousarray(ia, np.uint32)
        ib = np.ascontiguousarray(ib, np.uint32)
        out = np.empty(ia.size, np.float32)
        _lib.qzo_bin_score_pairs(_p(self.storage, _u8p), C.c_uint32(self.dim), C.c_int(self.store),
                                 C.c_int(self.distance), C.c_int(int(self.invert)), _p(ia, _u32p), _p(ib, _u32p),
                                 C.c_uint64(ia.size), _p(out, _f32p))
        return out

    def search(self, enc_query, k, largest, id_base=0, threads=1, use_ref=True):
        os_, oi = np.empty(k, np.float32), np.empty(k, np.uint32)
        eq = np.ascontiguousarray(np.concatenate([enc_query, np.zeros(16, np.uint8)]))
        kk = self._kernels(use_ref)
        m = _lib.qzo_bin_scan_topk_mt(_p(self.storage, _u8p), C.c_uint64(self.count), C.c_uint32(self.dim),
                                      C.c_int(self.store), C.c_int(self.distance), C.c_int(int(self.invert)),
                                      _p(eq, _u8p), kk[0], kk[1], kk[2], C.c_uint32(k), C.c_int(int(largest)),
                                      C.c_uint32(id_base), C.c_int(threads), _p(os_, _f32p), _p(oi, _u32p))
        return os_[:m].copy(), oi[:m].copy()


# ---------------------------------------------------------------- product quantizer
def pq_division(dim, chunk_size):
    """get_vector_division, pq.rs:116-121 -> bounds array of n_chunks+1"""
    n = int(_lib.qzo_pq_division(C.c_uint32(dim), C.c_uint32(chunk_size), None))
    b = np.empty(n + 1, np.uint32)
    _lib.qzo_pq_division(C.c_uint32(dim), C.c_uint32(chunk_size), _p(b, _u32p))
    return b


def kmeans(data, n_centroids=256, max_iter=100, max_threads=1, accuracy=1e-5, rng_seed=7):
    """kmeans.rs:7-47; returns (centroids [n_centroids, dim], iterations)"""
    data = np.ascontiguousarray(data, np.float32)
    n, dim = data.shape
    cen = np.empty((n_centroids, dim), np.float32)
    it = _lib.qzo_kmeans(_p(data, _f32p), C.c_uint64(n), C.c_uint32(dim), C.c_uint32(n_centroids), C.c_uint32(max_iter),
                         C.c_uint32(max_threads), C.c_float(accuracy), C.c_uint64(rng_seed), _p(cen, _f32p), None)
    return cen, int(it)


class OraclePQ:
    """EncodedVectorsPQ (encoded_vectors_pq.rs:27-594) restated on the CPU."""

    N_CENTROIDS = 256  # pq.rs:25

    def __init__(self, codes, count, dim, bounds, centroids, distance, invert):
        self.codes = np.ascontiguousarray(codes, np.uint8).reshape(-1)
        self.count, self.dim = int(count), int(dim)
        self.bounds = np.ascontiguousarray(bounds, np.uint32)
        self.n_chunks = self.bounds.size - 1
        self.centroids = np.ascontiguousarray(centroids, np.float32).reshape(self.N_CENTROIDS, dim)
        self.distance, self.invert = DIST[distance], bool(invert)

    @staticmethod
    def find_centroids(data, bounds, sample_ids=None, max_threads=1, rng_seed=7, max_iter=100, accuracy=1e-5):
        """pq.rs:278-342; sample ids explicit (random Permutor in the reference, :300-307)"""
        data = np.ascontiguousarray(data, np.float32)
        count, dim = data.shape
        if sample_ids is None:
            sample_ids = np.arange(min(count, 10_000))
        sample_ids = np.ascontiguousarray(np.sort(sample_ids), np.uint64)
        cen = np.empty((256, dim), np.float32)
        _lib.qzo_pq_find_centroids(_p(data, _f32p), C.c_uint64(count), C.c_uint32(dim), _p(bounds, _u32p),
                                   C.c_uint32(bounds.size - 1), _p(sample_ids, _u64p), C.c_uint64(sample_ids.size),
                                   C.c_uint32(256), C.c_uint32(max_iter), C.c_uint32(max_threads), C.c_float(accuracy),
                                   C.c_uint64(rng_seed), _p(cen, _f32p))
        return cen

    @staticmethod
    def assign(data, bounds, centroids):
        """encode_vector for every row, pq.rs:237-265"""
        data = np.ascontiguousarray(data, np.float32)
        count, dim = data.shape
        nch = bounds.size - 1
        codes = np.empty(count * nch, np.uint8)
        cen = np.ascontiguousarray(centroids, np.float32)
        _lib.qzo_pq_assign(_p(data, _f32p), C.c_uint64(count), C.c_uint32(dim), _p(bounds, _u32p), C.c_uint32(nch),
                           _p(cen, _f32p), C.c_uint32(256), _p(codes, _u8p))
        return codes

    @classmethod
    def encode(cls, data, distance, invert, chunk_size, sample_ids=None, max_threads=1, centroids=None):
        """pq.rs:56-107"""
        data = np.ascontiguousarray(data, np.float32)
        count, dim = data.shape
        bounds = pq_division(dim, chunk_size)
        if centroids is None:
            centroids = cls.find_centroids(data, bounds, sample_ids, max_threads)
        codes = cls.assign(data, bounds, centroids) if count else np.zeros(0, np.uint8)
        return cls(codes, count, dim, bounds, centroids, distance, invert)

    def encode_query(self, q):
        """pq.rs:525-547 -> lut [n_chunks*256] f32"""
        q = np.ascontiguousarray(q, np.float32)
        lut = np.empty(self.n_chunks * 256, np.float32)
        _lib.qzo_pq_lut(_p(q, _f32p), C.c_uint32(self.dim), _p(self.bounds, _u32p), C.c_uint32(self.n_chunks),
                        _p(self.centroids, _f32p), C.c_uint32(256), C.c_int(self.distance), C.c_int(int(self.invert)),
                        _p(lut, _f32p))
        return lut

    def scan(self, lut, first=0, n=None, order="sse"):
        """score_point, pq.rs:549-561 (x86 -> score_point_sse :405-440; 'simple' -> :476-494)"""
        n = self.count - first if n is None else n
        out = np.empty(n, np.float32)
        _lib.qzo_pq_scan(_p(self.codes, _u8p), C.c_uint64(first), C.c_uint64(n), C.c_uint32(self.n_chunks),
                         _p(lut, _f32p), C.c_uint32(256), C.c_int(0 if order == "sse" else 1), _p(out, _f32p))
        return out

    def score_pairs(self, ia, ib):
        """score_internal, pq.rs:566-593"""
        ia = np.ascontiguousarray(ia, np.uint32)
        ib = np.ascontiguousarray(ib, np.uint32)
        out = np.empty(ia.size, np.float32)
        _lib.qzo_pq_score_pairs(_p(self.codes, _u8p), C.c_uint32(self.dim), _p(self.bounds, _u32p),
                                C.c_uint32(self.n_chunks), _p(self.centroids, _f32p), C.c_int(self.distance),
                                C.c_int(int(self.invert)), _p(ia, _u32p), _p(ib, _u32p), C.c_uint64(ia.size),
                                _p(out, _f32p))
        return out

    def search(self, lut, k, largest, id_base=0, threads=1, order="sse"):
        os_, oi = np.empty(k, np.float32), np.empty(k, np.uint32)
        m = _lib.qzo_pq_scan_topk_mt(_p(self.codes, _u8p), C.c_uint64(self.count), C.c_uint32(self.n_chunks),
                                     _p(lut, _f32p), C.c_uint32(256), C.c_int(0 if order == "sse" else 1),
                                     C.c_uint32(k), C.c_int(int(largest)), C.c_uint32(id_base), C.c_int(threads),
                                     _p(os_, _f32p), _p(oi, _u32p))
        return os_[:m].copy(), oi[:m].copy()
