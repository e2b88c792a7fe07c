"""ctypes binding of libb200r.so (the C ABI in include/b200r.h).

The library is built in-tree (``generativeaiexamples_b200/libb200r.so``) by
``__graft_entry__.build()`` or ``make -C generativeaiexamples_b200/csrc``.  It is the
only compute path: if it is missing, or no sm_100 GPU is visible when an index is
created, the caller gets an exception -- never a silent CPU substitute.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# (B200R_LIB: a developer override used by tools/ to compare two builds of the library; the product path is the in-tree file)
LIB_PATH = os.environ.get("B200R_LIB") or os.path.join(_HERE, "libb200r.so")

OK, E_INVALID, E_CUDA, E_UNSUPPORTED, E_NOMEM, E_IO = range(6)
F32, BF16 = 0, 1
HOST, DEVICE = 0, 1
PATH_AUTO, PATH_SCAN, PATH_GEMM = 0, 1, 2
MAX_DEVICES = 8


class Config(ctypes.Structure):
    _fields_ = [
        ("dim", ctypes.c_int32), ("metric", ctypes.c_int32), ("storage", ctypes.c_int32),
        ("normalize_on_add", ctypes.c_int32), ("normalize_queries", ctypes.c_int32),
        ("device", ctypes.c_int32), ("search_path", ctypes.c_int32), ("shadow_bf16", ctypes.c_int32),
        ("reserve_rows", ctypes.c_int64), ("id_base", ctypes.c_int64),
        ("n_devices", ctypes.c_int32), ("device_ids", ctypes.c_int32 * MAX_DEVICES),
    ]


class Stats(ctypes.Structure):
    _fields_ = [
        ("ntotal", ctypes.c_int64), ("capacity_rows", ctypes.c_int64), ("n_search_calls", ctypes.c_int64),
        ("n_queries", ctypes.c_int64), ("n_kernel_launches", ctypes.c_int64), ("bytes_scanned", ctypes.c_int64),
        ("n_rows_added", ctypes.c_int64), ("n_rows_removed", ctypes.c_int64), ("last_grid", ctypes.c_int32),
        ("last_path", ctypes.c_int32), ("num_sms", ctypes.c_int32), ("reserved0", ctypes.c_int32),
        ("last_search_ms", ctypes.c_double), ("n_gemm_fallbacks", ctypes.c_int64),
        ("last_kernel_us", ctypes.c_double), ("last_exchange_us", ctypes.c_double),
        ("n_devices", ctypes.c_int32), ("n_segments", ctypes.c_int32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if not k.startswith("reserved")}


# name -> (restype, argtypes); every symbol include/b200r.h declares
_vp, _i32, _i64, _int = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_int
SIGNATURES = {
    "b200r_create": (_int, [ctypes.POINTER(Config), ctypes.POINTER(_vp)]),
    "b200r_destroy": (_int, [_vp]),
    "b200r_add": (_int, [_vp, _vp, _i64, _int, _int, _vp]),
    "b200r_search": (_int, [_vp, _vp, _i64, _int, _int, _vp, _vp, _int, _vp]),
    "b200r_search_threshold": (_int, [_vp, _vp, _i64, _int, _i32, ctypes.c_double, _i32, _vp, _vp, _vp, _vp, _vp]),
    "b200r_mmr": (_int, [_vp, _vp, _vp, _i64, _int, ctypes.c_double, _vp, ctypes.POINTER(_i32)]),
    "b200r_remove_ids": (_int, [_vp, _vp, _i64, ctypes.POINTER(_i64)]),
    "b200r_ntotal": (_int, [_vp, ctypes.POINTER(_i64)]),
    "b200r_reconstruct": (_int, [_vp, _i64, _vp]),
    "b200r_reconstruct_batch": (_int, [_vp, _vp, _i64, _vp]),
    "b200r_reset": (_int, [_vp]),
    "b200r_normalize_l2": (_int, [_vp, _i64, _i32, _int, _i32, _vp]),
    "b200r_save": (_int, [_vp, ctypes.c_char_p]),
    "b200r_load": (_int, [ctypes.c_char_p, ctypes.POINTER(Config), ctypes.POINTER(_vp)]),
    "b200r_merge_topk": (_int, [_i32, _vp, _vp, _i64, _i32, _i32, _i64, _i64, _i32, _vp, _vp, _i32, _vp]),
    "b200r_search_noalloc": (_int, [_vp, _i64, _i32, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _i64, _i32, _vp]),
    "b200r_exchange_buffer_bytes": (_i64, []),
    "b200r_search_exchange": (_int, [_vp, _vp, _i64, _int, _int, ctypes.POINTER(_vp), _i32, _i32, ctypes.c_uint32, _vp,
                                     _vp, _int, _vp]),
    "b200r_exchange_status": (_int, [_vp, _i32, _vp, ctypes.POINTER(_i32)]),
    "b200r_reload_env": (_int, []),
    "b200r_auto_route": (_int, [_vp, _i64, _i64, ctypes.POINTER(_i32)]),
    "b200r_set_id_base": (_int, [_vp, _i64]),
    "b200r_stats_get": (_int, [_vp, ctypes.POINTER(Stats)]),
    "b200r_merge_from": (_int, [_vp, _vp]),
    "b200r_set_shadow_filter": (_int, [_vp, _i32]),
    "b200r_set_timing": (_int, [_vp, _i32]),
    "b200r_last_error": (ctypes.c_char_p, []),
    "b200r_version": (_int, []),
}

_lib = None


class B200RError(RuntimeError):
    """Non-zero status from libb200r (faiss raises RuntimeError for C++ exceptions)."""

    def __init__(self, code, msg):
        super().__init__(f"libb200r error {code}: {msg}")
        self.code = code


def load():
    """dlopen the library and bind every declared symbol; raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is not built. Run `python -c 'import __graft_entry__ as g; g.build()'` or "
            "`make -C generativeaiexamples_b200/csrc`. There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


_c_char, _addressof = ctypes.c_char, ctypes.addressof


def addr(a) -> int:
    """Address of a C-contiguous numpy array's data.  ``a.ctypes.data_as(c_void_p)`` costs ~3.9 us per array and a
    host search passes three of them; going through the buffer protocol costs ~0.5 us.  Read-only and empty arrays
    (which the buffer route refuses) take numpy's own accessor."""
    try:
        return _addressof(_c_char.from_buffer(a))
    except (TypeError, ValueError):
        return a.ctypes.data


def check(rc):
    if rc != OK:
        msg = load().b200r_last_error()
        msg = msg.decode("utf-8", "replace") if msg else ""
        if rc == E_NOMEM:
            raise MemoryError(f"libb200r: {msg}")
        raise B200RError(rc, msg)


def make_config(*, dim, metric, storage=F32, normalize_on_add=False, normalize_queries=False, device=0,
                search_path=PATH_AUTO, shadow_bf16=False, reserve_rows=0, id_base=0, devices=None) -> Config:
    """``b200r_config`` from keyword arguments; ``devices`` (a list of 2..8 CUDA ordinals) makes a
    multi-GPU index inside this process."""
    cfg = Config(dim=int(dim), metric=int(metric), storage=int(storage), normalize_on_add=int(bool(normalize_on_add)),
                 normalize_queries=int(bool(normalize_queries)), device=int(device), search_path=int(search_path),
                 shadow_bf16=int(bool(shadow_bf16)), reserve_rows=int(reserve_rows), id_base=int(id_base))
    devs = [int(d) for d in devices] if devices else []
    if len(devs) > MAX_DEVICES:
        raise ValueError(f"at most {MAX_DEVICES} devices, got {len(devs)}")
    if len(devs) == 1:
        cfg.device = devs[0]
        devs = []
    cfg.n_devices = len(devs)
    for i, d in enumerate(devs):
        cfg.device_ids[i] = d
    if devs:
        cfg.device = devs[0]
    return cfg


def reload_env() -> None:
    """Make libb200r read its ``B200R_*`` environment switches again (it reads them once per process)."""
    check(load().b200r_reload_env())
