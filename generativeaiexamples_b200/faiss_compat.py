"""The handful of ``faiss`` module names the reference's direct-faiss callers use, over the B200 engine.

``community/chat-llama-nemotron/backend-rag/src/rag_service.py`` never goes through LangChain: it calls
``faiss.IndexFlatL2(dim)`` (:88, :447), ``faiss.normalize_L2(x)`` (:194, :292, :439), ``index.add`` (:197, :450),
``index.search`` (:301), ``faiss.write_index(index, path)`` (:376) and ``faiss.read_index(path)`` (:399).
For such a caller the switch is one line::

    import generativeaiexamples_b200.faiss_compat as faiss

Files are faiss's own ``IxF2`` / ``IxFI`` layout both ways, so an index written by either side loads on
the other.  ``B200R_DEVICES=0,1,…`` spreads every index created here over those GPUs (``factory.py``).
"""
from __future__ import annotations

import os
from typing import List, Optional

from .flat_index import (  # noqa: F401
    B200IndexFlat, METRIC_INNER_PRODUCT, METRIC_L2, normalize_L2,
)


def _devices() -> Optional[List[int]]:
    env = os.environ.get("B200R_DEVICES", "").strip()
    return [int(t) for t in env.split(",") if t.strip() != ""] if env else None


def _device_kw() -> dict:
    devs = _devices()
    if devs:
        return {"devices": devs}
    return {"device": int(os.environ.get("B200R_DEVICE", "0"))}


def IndexFlatL2(d: int) -> B200IndexFlat:
    """``faiss.IndexFlatL2(d)``: squared-L2 scores, ascending."""
    return B200IndexFlat(int(d), METRIC_L2, **_device_kw())


def IndexFlatIP(d: int) -> B200IndexFlat:
    """``faiss.IndexFlatIP(d)``: inner-product scores, descending."""
    return B200IndexFlat(int(d), METRIC_INNER_PRODUCT, **_device_kw())


def IndexFlat(d: int, metric: int = METRIC_L2) -> B200IndexFlat:
    """``faiss.IndexFlat(d, metric)``."""
    if metric not in (METRIC_L2, METRIC_INNER_PRODUCT):
        raise ValueError(f"unsupported metric {metric}: the flat index knows METRIC_L2 and METRIC_INNER_PRODUCT")
    return B200IndexFlat(int(d), int(metric), **_device_kw())


def write_index(index: B200IndexFlat, path) -> None:
    """``faiss.write_index(index, path)``."""
    if not isinstance(index, B200IndexFlat):
        raise TypeError("write_index needs an index created through this module")
    index.write(os.fspath(path))


def read_index(path) -> B200IndexFlat:
    """``faiss.read_index(path)`` for flat indexes (``IxF2`` / ``IxFI`` files); raises ``RuntimeError`` with the
    engine's message for anything else, as faiss does for a file it cannot read."""
    return B200IndexFlat.read(os.fspath(path), **_device_kw())
