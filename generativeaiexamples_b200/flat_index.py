"""``B200IndexFlat``: the faiss ``IndexFlatL2`` / ``IndexFlatIP`` python surface on libb200r.

Mirrors what ``faiss/python/class_wrappers.py`` (faiss 1.7.4) gives the reference:
``replacement_add`` / ``replacement_search`` / ``replacement_remove_ids`` /
``replacement_reconstruct`` plus ``faiss.normalize_L2`` -- same names, same argument
meaning, same error behaviour (AssertionError on a dimension mismatch or ``k <= 0``,
RuntimeError from the engine).  Reference call sites:
``RAG/src/chain_server/utils.py:306`` (``IndexFlatL2(dim)``),
``community/chat-llama-nemotron/backend-rag/src/rag_service.py:88,194,197,292,301,450``.

numpy arrays and torch tensors (CPU or CUDA) are accepted; torch is used only to read
``data_ptr()`` / the current stream and to allocate result tensors for device callers.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import numpy as np

from . import _lib

METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1

try:  # torch is plumbing only; the index works without it for numpy callers
    import torch
except Exception:  # pragma: no cover
    torch = None


def _is_torch(x):
    return torch is not None and isinstance(x, torch.Tensor)


def _torch_stream(device_index):
    return ctypes.c_void_p(torch.cuda.current_stream(device_index).cuda_stream)


class B200IndexFlat:
    """Exact (brute-force) index resident in one B200's HBM."""

    def __init__(self, d: int, metric: int = METRIC_L2, *, storage: str = "f32", device: int = 0,
                 normalize_on_add: bool = False, normalize_queries: bool = False, reserve_rows: int = 0,
                 id_base: int = 0, search_path: int = _lib.PATH_AUTO, shadow_bf16: bool = False,
                 devices=None):
        if storage not in ("f32", "bf16"):
            raise ValueError("storage must be 'f32' or 'bf16'")
        if shadow_bf16 and storage != "f32":
            raise ValueError("shadow_bf16 needs storage='f32'")
        self._h = None
        self._L = _lib.load()
        self.d = int(d)
        self.metric_type = int(metric)
        self.is_trained = True
        self.storage = storage
        self.normalize_on_add = bool(normalize_on_add)
        # devices=[0, 1, ...]: one index spread over several GPUs of this process (rows keep faiss's
        # "id == position" contract; see csrc/multi.cu).  self.device is then the primary GPU, the one
        # that CUDA tensors handed to add()/search() are expected on.
        self.devices = [int(x) for x in devices] if devices else [int(device)]
        self.device = self.devices[0]
        cfg = _lib.make_config(dim=self.d, metric=self.metric_type,
                               storage=_lib.BF16 if storage == "bf16" else _lib.F32,
                               normalize_on_add=normalize_on_add, normalize_queries=normalize_queries,
                               device=self.device, search_path=search_path, shadow_bf16=shadow_bf16,
                               reserve_rows=reserve_rows, id_base=id_base, devices=self.devices)
        h = ctypes.c_void_p()
        _lib.check(self._L.b200r_create(ctypes.byref(cfg), ctypes.byref(h)))
        self._h = h

    # -- lifetime -----------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._L.b200r_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def ntotal(self) -> int:
        out = ctypes.c_int64()
        _lib.check(self._L.b200r_ntotal(self._h, ctypes.byref(out)))
        return out.value

    # -- add ------------------------------------------------------------------
    def add(self, x) -> None:
        """faiss ``replacement_add``: x is (n, d); converted to C-contiguous float32."""
        if _is_torch(x) and x.is_cuda:
            n, d = x.shape
            assert d == self.d
            assert x.device.index in self.devices, "tensor and index live on different GPUs"
            x = x.contiguous().to(torch.float32)
            with torch.cuda.device(x.device):  # the stream handed over belongs to the tensor's GPU
                _lib.check(self._L.b200r_add(self._h, ctypes.c_void_p(x.data_ptr()), n, _lib.F32, _lib.DEVICE,
                                             _torch_stream(x.device.index)))
                # the ingest kernel reads x asynchronously on the caller's stream: keep it
                # alive until that stream has passed this point
                x.record_stream(torch.cuda.current_stream(x.device.index))
            return
        if _is_torch(x):
            x = x.numpy()
        n, d = x.shape
        assert d == self.d
        x = np.ascontiguousarray(x, dtype="float32")
        _lib.check(self._L.b200r_add(self._h, _lib.addr(x), n, _lib.F32, _lib.HOST, None))

    # -- search ---------------------------------------------------------------
    def search(self, x, k: int, D=None, I=None):
        """faiss ``replacement_search``: returns (D (n,k) float32, I (n,k) int64).

        numpy / CPU input -> numpy output.  CUDA tensor input -> CUDA tensors, filled
        asynchronously on the current torch stream (no host round trip)."""
        if _is_torch(x) and x.is_cuda:
            n, d = x.shape
            assert d == self.d
            assert k > 0
            assert x.device.index in self.devices, "tensor and index live on different GPUs"
            x = x.contiguous().to(torch.float32)
            if D is None:
                D = torch.empty((n, k), dtype=torch.float32, device=x.device)
            if I is None:
                I = torch.empty((n, k), dtype=torch.int64, device=x.device)
            assert D.shape == (n, k) and I.shape == (n, k) and D.is_contiguous() and I.is_contiguous()
            with torch.cuda.device(x.device):
                _lib.check(self._L.b200r_search(self._h, ctypes.c_void_p(x.data_ptr()), n, k, _lib.DEVICE,
                                                ctypes.c_void_p(D.data_ptr()), ctypes.c_void_p(I.data_ptr()),
                                                _lib.DEVICE, _torch_stream(x.device.index)))
                x.record_stream(torch.cuda.current_stream(x.device.index))
            return D, I
        if _is_torch(x):
            x = x.numpy()
        n, d = x.shape
        assert d == self.d
        x = np.ascontiguousarray(x, dtype="float32")
        assert k > 0
        if D is None and I is None:
            # one allocation for both results (two views of it): numpy's per-array overhead is what is left of a
            # small search's host time
            ids_off = (n * k * 4 + 7) // 8 * 8
            buf = np.empty(ids_off + n * k * 8, dtype=np.uint8)
            D = np.frombuffer(buf, np.float32, n * k, 0).reshape(n, k)
            I = np.frombuffer(buf, np.int64, n * k, ids_off).reshape(n, k)
        if D is None:
            D = np.empty((n, k), dtype=np.float32)
        else:
            assert D.shape == (n, k)
        if I is None:
            I = np.empty((n, k), dtype=np.int64)
        else:
            assert I.shape == (n, k)
        _lib.check(self._L.b200r_search(self._h, _lib.addr(x), n, k, _lib.HOST,
                                        _lib.addr(D), _lib.addr(I),
                                        _lib.HOST, None))
        return D, I

    # -- search + relevance threshold / MMR selection on the device (no faiss counterpart) -----------
    RELEVANCE_KINDS = {"EUCLIDEAN_DISTANCE": 0, "COSINE": 1, "MAX_INNER_PRODUCT": 2}

    def search_threshold(self, x, k: int, relevance: str, score_threshold: Optional[float]):
        """``search`` followed on the GPU by LangChain's relevance function and ``score >= score_threshold``
        (``VectorStore.similarity_search_with_relevance_scores``).  Returns ``(R, D, I, counts, out_of_range)``:
        for query ``i`` the first ``counts[i]`` entries are the kept hits (relevance float64, raw score float32,
        id), the dropped hits follow, then padding with id -1."""
        x = np.ascontiguousarray(x, dtype="float32")
        n, d = x.shape
        assert d == self.d
        assert k > 0
        R = np.empty((n, k), dtype=np.float64)
        D = np.empty((n, k), dtype=np.float32)
        I = np.empty((n, k), dtype=np.int64)
        counts = np.empty(n, dtype=np.int32)
        oor = np.empty(n, dtype=np.int32)
        _lib.check(self._L.b200r_search_threshold(
            self._h, _lib.addr(x), n, k, self.RELEVANCE_KINDS[relevance],
            0.0 if score_threshold is None else float(score_threshold), int(score_threshold is not None),
            _lib.addr(R), _lib.addr(D), _lib.addr(I),
            _lib.addr(counts), _lib.addr(oor)))
        return R, D, I, counts, oor

    def mmr_select(self, query, keys, k: int, lambda_mult: float = 0.5) -> np.ndarray:
        """LangChain's ``maximal_marginal_relevance`` over the stored rows ``keys``: positions into ``keys`` of the
        ``min(k, len(keys))`` picks, in picking order.  The rows are gathered and compared on the GPU."""
        query = np.ascontiguousarray(query, dtype="float32").reshape(-1)
        assert query.size == self.d
        keys = np.ascontiguousarray(keys, dtype="int64")
        assert keys.ndim == 1 and k > 0
        sel = np.empty(k, dtype=np.int32)
        n_sel = ctypes.c_int32(0)
        _lib.check(self._L.b200r_mmr(self._h, _lib.addr(query), _lib.addr(keys),
                                     keys.size, k, float(lambda_mult), _lib.addr(sel),
                                     ctypes.byref(n_sel)))
        return sel[: n_sel.value].copy()

    # -- remove / reconstruct / reset ----------------------------------------
    def remove_ids(self, x) -> int:
        """faiss ``replacement_remove_ids`` with an id array (IDSelectorBatch)."""
        x = np.ascontiguousarray(x, dtype="int64")
        assert x.ndim == 1
        out = ctypes.c_int64()
        _lib.check(self._L.b200r_remove_ids(self._h, _lib.addr(x), x.size, ctypes.byref(out)))
        return out.value

    def reconstruct(self, key: int, x: Optional[np.ndarray] = None) -> np.ndarray:
        if x is None:
            x = np.empty(self.d, dtype=np.float32)
        else:
            assert x.shape == (self.d,)
        _lib.check(self._L.b200r_reconstruct(self._h, int(key), _lib.addr(x)))
        return x

    def reconstruct_batch(self, keys) -> np.ndarray:
        """faiss ``reconstruct_batch``: (n, d) float32 for the given positions, one launch."""
        keys = np.ascontiguousarray(keys, dtype="int64")
        assert keys.ndim == 1
        out = np.empty((keys.size, self.d), dtype=np.float32)
        _lib.check(self._L.b200r_reconstruct_batch(self._h, _lib.addr(keys), keys.size,
                                                   _lib.addr(out)))
        return out

    def reset(self) -> None:
        _lib.check(self._L.b200r_reset(self._h))

    # -- persistence (faiss.write_index / read_index) ---------------------------
    def write(self, path: str) -> None:
        _lib.check(self._L.b200r_save(self._h, str(path).encode()))

    @classmethod
    def read(cls, path: str, *, storage: str = "f32", device: int = 0, devices=None, **kw) -> "B200IndexFlat":
        L = _lib.load()
        self = cls.__new__(cls)
        self._L = L
        self._h = None
        self.devices = [int(x) for x in devices] if devices else [int(device)]
        cfg = _lib.make_config(dim=1, metric=METRIC_L2, storage=_lib.BF16 if storage == "bf16" else _lib.F32,
                               normalize_on_add=kw.get("normalize_on_add", False),
                               normalize_queries=kw.get("normalize_queries", False), device=self.devices[0],
                               search_path=kw.get("search_path", _lib.PATH_AUTO),
                               shadow_bf16=kw.get("shadow_bf16", False), reserve_rows=0,
                               id_base=kw.get("id_base", 0), devices=self.devices)
        h = ctypes.c_void_p()
        _lib.check(L.b200r_load(str(path).encode(), ctypes.byref(cfg), ctypes.byref(h)))
        self._h = h
        self.storage, self.device, self.is_trained = storage, self.devices[0], True
        self.normalize_on_add = bool(kw.get("normalize_on_add", False))
        # d and metric come from the file: recover them through a zero-row probe
        st = self.stats()
        self.d, self.metric_type = _read_header(path)
        assert st["ntotal"] >= 0
        return self

    def merge_from(self, other: "B200IndexFlat", add_id: int = 0) -> None:
        """faiss ``index.merge_from(other)``: append ``other``'s vectors and empty ``other`` (``add_id`` is
        unused by flat indexes, as in faiss)."""
        if not isinstance(other, B200IndexFlat):
            raise TypeError("merge_from needs another B200IndexFlat")
        _lib.check(self._L.b200r_merge_from(self._h, other._h))

    # -- engine extras --------------------------------------------------------
    def set_id_base(self, base: int) -> None:
        _lib.check(self._L.b200r_set_id_base(self._h, int(base)))

    def set_shadow_filter(self, enabled: bool) -> None:
        """Index built with ``shadow_bf16=True``: filter batches on the bf16 shadow (default) or on the fp32 rows."""
        _lib.check(self._L.b200r_set_shadow_filter(self._h, int(enabled)))

    def set_timing(self, enabled: bool) -> None:
        _lib.check(self._L.b200r_set_timing(self._h, int(enabled)))

    def stats(self) -> dict:
        st = _lib.Stats()
        _lib.check(self._L.b200r_stats_get(self._h, ctypes.byref(st)))
        return st.as_dict()


def _read_header(path) -> Tuple[int, int]:
    import struct

    with open(path, "rb") as f:
        head = f.read(4 + 4 + 8 + 8 + 8 + 1 + 4)
    _, d, _, _, _, _, metric = struct.unpack("<IiqqqBi", head)
    return d, metric


def B200IndexFlatL2(d: int, **kw) -> B200IndexFlat:
    """``faiss.IndexFlatL2(d)`` (utils.py:306)."""
    return B200IndexFlat(d, METRIC_L2, **kw)


def B200IndexFlatIP(d: int, **kw) -> B200IndexFlat:
    """``faiss.IndexFlatIP(d)``."""
    return B200IndexFlat(d, METRIC_INNER_PRODUCT, **kw)


def normalize_L2(x, device: int = 0) -> None:
    """``faiss.normalize_L2(x)``: in place; rows with zero norm are left untouched."""
    L = _lib.load()
    if _is_torch(x) and x.is_cuda:
        assert x.dtype == torch.float32 and x.is_contiguous() and x.dim() == 2
        _lib.check(L.b200r_normalize_l2(ctypes.c_void_p(x.data_ptr()), x.shape[0], x.shape[1], _lib.DEVICE,
                                        x.device.index, _torch_stream(x.device.index)))
        return
    if _is_torch(x):
        x = x.numpy()
    assert x.dtype == np.float32 and x.flags.c_contiguous and x.ndim == 2
    _lib.check(L.b200r_normalize_l2(_lib.addr(x), x.shape[0], x.shape[1], _lib.HOST, device, None))
