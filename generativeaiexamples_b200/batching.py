"""Micro-batching of concurrent single-query searches.

The only batched caller in the reference,
``FaissVectorDBResourceService.similarity_search(embeddings, k)``
(community/event-driven-rag-cve-analysis/cyber_dev_day/faiss_vdb_service.py:147-174), fans a
list of B query embeddings out as ``asyncio.gather`` of B ``asimilarity_search_by_vector`` calls,
i.e. B threads each running an nq=1 ``index.search``.  On a CPU that is how faiss gets its
parallelism; on the GPU it would serialise B full corpus scans.  ``SearchCoalescer`` sits between
the store and the index: while one search is in flight the requests that arrive are queued, and
the next launch serves all of them as ONE nq=B search (the tensor-core path from B >= 5).  An idle
coalescer adds no waiting: the first request starts immediately.  After a batch that served several
callers the drain loop lingers for a few per cent of that batch's duration (at most 0.5 ms) before it
forms the next one: the callers it has just released are about to come back, and without the pause
they alternate between two half-size batches (64 callers: 2 x 32 per corpus pass instead of 64).
"""
from __future__ import annotations

import threading
import time
from typing import List, Optional, Tuple

import numpy as np


class _Request:
    __slots__ = ("vec", "k", "event", "finished", "lead", "D", "I", "err")

    def __init__(self, vec, k):
        self.vec, self.k = vec, k
        self.event = threading.Event()  # set when the request is served OR when its caller is made the leader
        self.finished = False
        self.lead = False
        self.D = self.I = self.err = None


class SearchCoalescer:
    def __init__(self, index, max_batch: int = 256, linger_frac: float = 0.03, max_linger_s: float = 5e-4):
        self.index = index
        self.max_batch = int(max_batch)
        self.linger_frac = float(linger_frac)
        self.max_linger_s = float(max_linger_s)
        self._lock = threading.Lock()
        self._pending: List[_Request] = []
        self._busy = False
        self.n_requests = 0
        self.n_launches = 0
        self.n_handoffs = 0

    def search(self, vec: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
        """vec: (1, d) float32.  Blocks until this request has been served; thread-safe."""
        req = _Request(np.ascontiguousarray(vec, dtype=np.float32).reshape(1, -1), int(k))
        with self._lock:
            self._pending.append(req)
            self.n_requests += 1
            lead = not self._busy
            if lead:
                self._busy = True
        while True:
            if lead:
                self._drain(req)
            if req.finished:
                break
            req.event.wait()
            with self._lock:
                lead, req.lead = req.lead, False
                if not req.finished:
                    req.event.clear()
        if req.err is not None:
            raise req.err
        return req.D, req.I

    def _drain(self, own: _Request) -> None:
        """Run by the leader -- the caller that found the coalescer idle, or one that an earlier leader handed
        over to: serve batches until the leader's OWN request is done, then pass the role to a caller that is
        still waiting (so no caller keeps serving strangers under sustained load) or, with nothing queued,
        mark the coalescer idle."""
        while True:
            with self._lock:
                if not self._pending:
                    self._busy = False
                    return
                if own.finished:
                    nxt = self._pending[0]
                    nxt.lead = True
                    nxt.event.set()
                    self.n_handoffs += 1
                    return
                k = self._pending[0].k
                batch = [r for r in self._pending if r.k == k][: self.max_batch]
                taken = set(map(id, batch))
                self._pending = [r for r in self._pending if id(r) not in taken]
            t0 = time.perf_counter()
            try:
                Q = np.concatenate([r.vec for r in batch], axis=0)
                D, I = self.index.search(Q, k)
                self.n_launches += 1
                for j, r in enumerate(batch):
                    r.D, r.I = D[j:j + 1], I[j:j + 1]
            except BaseException as e:  # deliver the failure to every waiter of this batch
                for r in batch:
                    r.err = e
            finally:
                with self._lock:
                    for r in batch:
                        r.finished = True
                        r.event.set()
            if len(batch) > 1 and self.linger_frac > 0:
                time.sleep(min(self.max_linger_s, self.linger_frac * (time.perf_counter() - t0)))
