"""ctypes loader + faiss-python-wrapper-shaped front end for oracle/flat_ref.c.

TEST INFRASTRUCTURE (see oracle/__init__.py).  ``IndexFlatRef`` mirrors the part of
``faiss.IndexFlatL2`` / ``faiss.IndexFlatIP`` that the reference calls
(faiss/python/class_wrappers.py ``replacement_add`` / ``replacement_search`` /
``replacement_remove_ids``; call sites: RAG/src/chain_server/utils.py:306,
community/chat-llama-nemotron/backend-rag/src/rag_service.py:88,194,197,292,301).
"""
from __future__ import annotations

import ctypes
import glob
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1

_f32p = ctypes.POINTER(ctypes.c_float)
_i64p = ctypes.POINTER(ctypes.c_int64)


def _ptr(a, t):
    return a.ctypes.data_as(t)


def _find_openblas_sgemm():
    """numpy bundles an ILP64 OpenBLAS (scipy_cblas_sgemm64_); borrow its sgemm so
    the nq>=20 path is timed with a real BLAS, as faiss would use."""
    try:
        libdir = os.path.join(os.path.dirname(os.path.dirname(np.__file__)), "numpy.libs")
        for p in glob.glob(os.path.join(libdir, "libscipy_openblas64_*.so")):
            lib = ctypes.CDLL(p, mode=ctypes.RTLD_GLOBAL)
            fn = getattr(lib, "scipy_cblas_sgemm64_", None)
            if fn is not None:
                return lib, ctypes.cast(fn, ctypes.c_void_p)
    except OSError:
        pass
    return None, None


def load_flat_ref(rebuild: bool = False):
    global _LIB
    if _LIB is not None and not rebuild:
        return _LIB
    so = os.path.join(_HERE, "libflat_ref.so")
    if rebuild or not os.path.exists(so):
        subprocess.check_call(["make", "-C", _HERE, "clean", "all"], stdout=subprocess.DEVNULL)
    try:
        lib = ctypes.CDLL(so)
    except OSError:
        subprocess.check_call(["make", "-C", _HERE, "clean", "all"], stdout=subprocess.DEVNULL)
        lib = ctypes.CDLL(so)
    lib.oracle_knn.argtypes = [ctypes.c_int, _f32p, ctypes.c_int64, _f32p, ctypes.c_int64,
                               ctypes.c_int64, ctypes.c_int64, _f32p, _i64p, ctypes.c_int]
    lib.oracle_knn.restype = ctypes.c_int
    lib.oracle_knn_allcores.argtypes = lib.oracle_knn.argtypes[:-1]
    lib.oracle_knn_allcores.restype = ctypes.c_int
    lib.oracle_renorm_L2.argtypes = [ctypes.c_int64, ctypes.c_int64, _f32p]
    lib.oracle_renorm_L2.restype = None
    lib.oracle_remove_ids.argtypes = [_f32p, _i64p, ctypes.c_int64, _i64p, ctypes.c_int64]
    lib.oracle_remove_ids.restype = ctypes.c_int64
    lib.oracle_set_sgemm.argtypes = [ctypes.c_void_p]
    lib.oracle_has_blas.restype = ctypes.c_int
    lib.oracle_max_threads.restype = ctypes.c_int
    lib.oracle_set_threads.argtypes = [ctypes.c_int]
    lib.oracle_set_threads.restype = None
    for name in ("oracle_fvec_inner_product", "oracle_fvec_L2sqr"):
        f = getattr(lib, name)
        f.argtypes = [_f32p, _f32p, ctypes.c_int64]
        f.restype = ctypes.c_float
    blas, fn = _find_openblas_sgemm()
    if fn is not None:
        lib.oracle_set_sgemm(fn)
        lib._blas_keepalive = blas
    _LIB = lib
    return lib


class FlatRef:
    """Stateless functions."""

    @staticmethod
    def knn(metric, q, x, k, path=0):
        lib = load_flat_ref()
        q = np.ascontiguousarray(q, dtype=np.float32)
        x = np.ascontiguousarray(x, dtype=np.float32)
        nq, d = q.shape
        D = np.empty((nq, k), dtype=np.float32)
        I = np.empty((nq, k), dtype=np.int64)
        rc = lib.oracle_knn(metric, _ptr(q, _f32p), nq, _ptr(x, _f32p), x.shape[0], d, k,
                            _ptr(D, _f32p), _ptr(I, _i64p), path)
        if rc != 0:
            raise RuntimeError("oracle_knn: bad arguments")
        return D, I

    @staticmethod
    def knn_allcores(metric, q, x, k):
        lib = load_flat_ref()
        q = np.ascontiguousarray(q, dtype=np.float32)
        x = np.ascontiguousarray(x, dtype=np.float32)
        nq, d = q.shape
        D = np.empty((nq, k), dtype=np.float32)
        I = np.empty((nq, k), dtype=np.int64)
        rc = lib.oracle_knn_allcores(metric, _ptr(q, _f32p), nq, _ptr(x, _f32p), x.shape[0], d, k,
                                     _ptr(D, _f32p), _ptr(I, _i64p))
        if rc != 0:
            raise RuntimeError("oracle_knn_allcores: bad arguments")
        return D, I

    @staticmethod
    def normalize_L2(x):
        """faiss.normalize_L2: in place, rows of zero norm untouched."""
        assert x.dtype == np.float32 and x.flags.c_contiguous and x.ndim == 2
        load_flat_ref().oracle_renorm_L2(x.shape[1], x.shape[0], _ptr(x, _f32p))


class IndexFlatRef:
    """The slice of faiss.IndexFlat{L2,IP} the reference uses."""

    def __init__(self, d: int, metric: int = METRIC_L2):
        self.d = int(d)
        self.metric_type = metric
        self.is_trained = True
        self._codes = np.empty((0, self.d), dtype=np.float32)

    @property
    def ntotal(self):
        return self._codes.shape[0]

    def add(self, x):
        n, d = x.shape
        assert d == self.d
        x = np.ascontiguousarray(x, dtype="float32")
        self._codes = np.concatenate([self._codes, x], axis=0)

    def search(self, x, k):
        n, d = x.shape
        assert d == self.d
        assert k > 0
        return FlatRef.knn(self.metric_type, x, self._codes, k)

    def remove_ids(self, x):
        x = np.ascontiguousarray(x, dtype=np.int64)
        assert x.ndim == 1
        nt = ctypes.c_int64(self.ntotal)
        codes = np.ascontiguousarray(self._codes)
        nrem = load_flat_ref().oracle_remove_ids(_ptr(codes, _f32p), ctypes.byref(nt), self.d,
                                                 _ptr(x, _i64p), x.size)
        self._codes = codes[: nt.value].copy()
        return int(nrem)

    def reconstruct(self, key):
        if not 0 <= key < self.ntotal:
            raise RuntimeError("reconstruct: key out of range")
        return self._codes[key].copy()

    def reset(self):
        self._codes = np.empty((0, self.d), dtype=np.float32)

    def merge_from(self, other, add_id=0):
        """faiss 1.7.4 IndexFlatCodes::merge_from (check_compatible_for_merge: same d, metric, type): the other
        index's codes are appended and the other index is emptied; add_id is unused by flat indexes."""
        if not isinstance(other, IndexFlatRef) or other.d != self.d or other.metric_type != self.metric_type:
            raise RuntimeError("merge_from: indexes are not compatible")
        self._codes = np.concatenate([self._codes, other._codes], axis=0)
        other.reset()


def IndexFlatL2Ref(d):
    return IndexFlatRef(d, METRIC_L2)


def IndexFlatIPRef(d):
    return IndexFlatRef(d, METRIC_INNER_PRODUCT)
