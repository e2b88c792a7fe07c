"""GPU suite at BASELINE.json's full single-GPU size (C2: 1M x 1024 fp32, IP, top-10),
through size-independent properties -- the CPU oracle would need minutes here:

* planted near-duplicates must come back at rank 0 with the planted id;
* every returned score equals the dot product recomputed from the reconstructed row;
* an independent fp64 brute force on the device (torch, checker only) gives the same ids
  modulo ties within 1e-5;
* splitting the corpus in two indexes and merging with K4 gives the same answer as one
  index (top-k of a union = top-k of the per-shard top-k's).
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_c2_properties(built_lib):
    import torch

    from generativeaiexamples_b200 import B200IndexFlatIP, _lib, synth
    from oracle import assert_topk_parity

    n, d, k, nq = 1_000_000, 1024, 10, 4
    dev = torch.device("cuda", 0)
    ix = B200IndexFlatIP(d, reserve_rows=n)
    half_a = B200IndexFlatIP(d, reserve_rows=n // 2)
    half_b = B200IndexFlatIP(d, reserve_rows=n // 2, id_base=n // 2)
    chunks = []
    for c in range(4):
        x = synth.corpus_chunk_device(c, n // 4, d, dev)
        ix.add(x)
        (half_a if c < 2 else half_b).add(x)
        chunks.append(x)
    torch.cuda.synchronize()
    assert ix.ntotal == n
    X = torch.cat(chunks)
    del chunks
    q, planted = synth.queries_device(nq, d, lambda ids: X[ids.to(dev)], n, dev)
    D, I = ix.search(q, k)
    torch.cuda.synchronize()
    Dh, Ih = D.cpu().numpy(), I.cpu().numpy()
    # planted neighbours
    for i in range(nq):
        if planted[i] >= 0:
            assert Ih[i, 0] == int(planted[i]) and Dh[i, 0] > 0.98
    # scores are real dot products of the rows they name
    for i in range(nq):
        rows = np.stack([ix.reconstruct(int(j)) for j in Ih[i]])
        np.testing.assert_allclose(Dh[i], rows.astype(np.float64) @ q[i].cpu().numpy().astype(np.float64), atol=1e-5)
    # fp64 brute force on the device as an independent checker
    S = torch.empty((nq, n), dtype=torch.float64, device=dev)
    for lo in range(0, n, 100_000):
        S[:, lo:lo + 100_000] = q.double() @ X[lo:lo + 100_000].double().T
    Dt, It = torch.topk(S, k + 8, dim=1)
    assert_topk_parity(Dh, Ih, Dt.cpu().numpy(), It.cpu().numpy(), what="C2 vs device fp64 brute force")
    # union property through the merge kernel
    Da, Ia = half_a.search(q, k)
    Db, Ib = half_b.search(q, k)
    cd = torch.stack([Da, Db]).contiguous()
    ci = torch.stack([Ia, Ib]).contiguous()
    Dm = torch.empty_like(Da)
    Im = torch.empty_like(Ia)
    _lib.check(built_lib.b200r_merge_topk(0, ctypes.c_void_p(cd.data_ptr()), ctypes.c_void_p(ci.data_ptr()), nq, 2, k,
                                          nq * k, nq * k, k, ctypes.c_void_p(Dm.data_ptr()),
                                          ctypes.c_void_p(Im.data_ptr()), 0,
                                          ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    np.testing.assert_array_equal(Im.cpu().numpy(), Ih)
    np.testing.assert_array_equal(Dm.cpu().numpy(), Dh)
    for i in (ix, half_a, half_b):
        i.close()


def test_c3_full_size_both_paths(built_lib):
    """BASELINE config C3 at its full single-GPU size (10M x 1024 fp32 = 41 GB): the scan path
    (nq=4) and the tensor-core path (nq=32) against a float64 brute force computed chunk by chunk
    on the device while the corpus is being ingested (checker only; nothing is kept but top-k)."""
    import torch

    from generativeaiexamples_b200 import B200IndexFlatIP, synth
    from oracle import assert_topk_parity

    free, _ = torch.cuda.mem_get_info()
    if free < 60 * (1 << 30):
        pytest.skip("needs ~50 GB of free HBM")
    n, d, k, nq = 10_000_000, 1024, 10, 32
    dev = torch.device("cuda", 0)
    ix = B200IndexFlatIP(d, reserve_rows=n)
    g = torch.Generator(device=dev)
    g.manual_seed(99)
    q = torch.randn((nq, d), generator=g, device=dev)
    best_s = torch.empty((nq, 0), dtype=torch.float64, device=dev)
    best_i = torch.empty((nq, 0), dtype=torch.int64, device=dev)
    for c in range(n // synth.CHUNK_ROWS):
        x = synth.corpus_chunk_device(c, synth.CHUNK_ROWS, d, dev)
        if c == 3:
            q[::2] = x[:nq // 2] + 0.1 * q[::2] / d ** 0.5   # plant neighbours at rows 3_000_000 ...
            q /= q.norm(dim=1, keepdim=True)
            # scores of the chunks already seen were computed with the old q: redo them
            best_s = torch.empty((nq, 0), dtype=torch.float64, device=dev)
            best_i = torch.empty((nq, 0), dtype=torch.int64, device=dev)
            for c2 in range(c):
                x2 = synth.corpus_chunk_device(c2, synth.CHUNK_ROWS, d, dev)
                s2 = q.double() @ x2.double().T
                ts, ti = torch.topk(torch.cat([best_s, s2], 1), k + 8, dim=1)
                ids = torch.cat([best_i, torch.arange(c2 * synth.CHUNK_ROWS, (c2 + 1) * synth.CHUNK_ROWS,
                                                      device=dev).expand(nq, -1)], 1)
                best_s, best_i = ts, torch.gather(ids, 1, ti)
                del x2, s2
        ix.add(x)
        s = q.double() @ x.double().T
        ts, ti = torch.topk(torch.cat([best_s, s], 1), k + 8, dim=1)
        ids = torch.cat([best_i, torch.arange(c * synth.CHUNK_ROWS, (c + 1) * synth.CHUNK_ROWS,
                                              device=dev).expand(nq, -1)], 1)
        best_s, best_i = ts, torch.gather(ids, 1, ti)
        del x, s
    torch.cuda.synchronize()
    assert ix.ntotal == n
    Dt, It = best_s.cpu().numpy(), best_i.cpu().numpy()
    # the scan path one query at a time (a corpus this large sends 2..4 queries to the GEMM path)
    for i in range(4):
        D1, I1 = ix.search(q[i: i + 1].contiguous(), k)
        assert ix.stats()["last_path"] == 1
        assert_topk_parity(D1.cpu().numpy(), I1.cpu().numpy(), Dt[i: i + 1], It[i: i + 1], what="C3 full size, scan path")
    D4, I4 = ix.search(q[:4].contiguous(), k)
    assert ix.stats()["last_path"] == 2
    assert_topk_parity(D4.cpu().numpy(), I4.cpu().numpy(), Dt[:4], It[:4], what="C3 full size, 4 queries")
    D32, I32 = ix.search(q, k)
    assert ix.stats()["last_path"] == 2
    assert_topk_parity(D32.cpu().numpy(), I32.cpu().numpy(), Dt, It, what="C3 full size, GEMM path")
    assert (I32[::2, 0].cpu() == 3_000_000 + torch.arange(nq // 2)).all()
    ix.close()
