"""GPU test of b200r_search_noalloc, the search entry point for callers that cannot allocate.

Some users of this library run where device memory is not theirs to ask for: a plug-in
inside an inference server that gave the whole HBM to its own pool at start-up (cudaMalloc
there fails, or takes another tenant's share), or a serving loop that cannot tolerate the
device-wide stall of cudaMalloc / cudaFree.  For them the library offers one search call
whose only scratch memory is a fixed array compiled into the library.  This test calls it
through the C ABI with the caller's own device buffers and checks it, twice in a row with
different shapes (the array keeps the previous call's contents, so stale data must not leak),
against the regular b200r_search, with the same tolerance as the other parity tests.

The file name sorts last on purpose: a run that stops at the first failure has then already
run every other test.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _noalloc(lib, X, storage, metric, q, k, device=0, id_base=0):
    import torch

    from generativeaiexamples_b200 import _lib

    nq = q.shape[0]
    D = torch.full((nq, k), float("nan"), dtype=torch.float32, device=q.device)
    I = torch.full((nq, k), -7, dtype=torch.int64, device=q.device)
    rc = lib.b200r_search_noalloc(ctypes.c_void_p(X.data_ptr()), X.shape[0], X.shape[1], storage, metric,
                                  ctypes.c_void_p(q.data_ptr()), nq, k, ctypes.c_void_p(D.data_ptr()),
                                  ctypes.c_void_p(I.data_ptr()), id_base, device,
                                  ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream))
    _lib.check(rc)
    torch.cuda.synchronize()
    return D.cpu().numpy(), I.cpu().numpy()


@pytest.mark.parametrize("metric", [0, 1])
def test_noalloc_matches_regular_search_twice(built_lib, metric):
    import torch

    from generativeaiexamples_b200 import B200IndexFlat
    from oracle import assert_topk_parity

    torch.manual_seed(1234 + metric)
    n, d = 60_000, 1024
    X = torch.randn((n, d), device="cuda")
    q = (X[[11, 30_000, 59_999]] + 0.05 * torch.randn((3, d), device="cuda")).contiguous()
    index = B200IndexFlat(d, metric)
    index.add(X)
    # regular entry point first, so that its workspace exists before free memory is sampled
    want1 = [t.cpu().numpy() for t in index.search(q, 10)]
    want2 = [t.cpu().numpy() for t in index.search(q[1:2].contiguous(), 33)]

    got1 = _noalloc(built_lib, X, 0, metric, q, 10)               # nq=3, k=10 (padded to 4 query slots)
    free_before = torch.cuda.mem_get_info()[0]
    got2 = _noalloc(built_lib, X, 0, metric, q[1:2].contiguous(), 33)   # nq=1, k=33: other kernel instance, same scratch
    free_after = torch.cuda.mem_get_info()[0]

    assert_topk_parity(got1[0], got1[1], want1[0], want1[1], what="no-alloc call 1")
    assert_topk_parity(got2[0], got2[1], want2[0], want2[1], what="no-alloc call 2")
    # ... and against the float64 brute force, so that a fault shared with b200r_search (both run the scan
    # kernel) cannot hide behind the comparison above
    from oracle import truth_topk

    name = "l2" if metric == 1 else "ip"
    Xh, qh = X.cpu().numpy(), q.cpu().numpy()
    assert_topk_parity(got1[0], got1[1], *truth_topk(name, qh, Xh, 18), what="no-alloc call 1 vs f64 truth")
    assert_topk_parity(got2[0], got2[1], *truth_topk(name, qh[1:2], Xh, 41), what="no-alloc call 2 vs f64 truth")
    assert got1[1][:, 0].tolist() == [11, 30_000, 59_999]
    assert free_after == free_before, "the no-alloc entry point changed the amount of free device memory"
    index.close()


def test_noalloc_bf16_idbase_and_limits(built_lib):
    import torch

    from generativeaiexamples_b200 import B200IndexFlat, _lib
    from oracle import assert_topk_parity

    torch.manual_seed(99)
    n, d = 20_000, 256
    X = torch.randn((n, d), device="cuda")
    Xb = X.to(torch.bfloat16).contiguous()
    q = X[:2].contiguous()
    index = B200IndexFlat(d, 0, storage="bf16", id_base=1_000_000)
    index.add(X)
    want = [t.cpu().numpy() for t in index.search(q, 5)]
    got = _noalloc(built_lib, Xb, 1, 0, q, 5, id_base=1_000_000)
    assert_topk_parity(got[0], got[1], want[0], want[1], what="no-alloc bf16 + id_base")
    index.close()
    # empty corpus: padding only
    Dp, Ip = _noalloc(built_lib, X[:0], 0, 1, q, 3)
    assert (Ip == -1).all() and (Dp == np.finfo(np.float32).max).all()

    # sizes beyond the built-in scratch are refused with a clear message, nothing is launched
    def rc_msg(*a):
        rc = built_lib.b200r_search_noalloc(*a)
        return rc, built_lib.b200r_last_error().decode()

    D = torch.empty((8, 200), dtype=torch.float32, device="cuda")
    I = torch.empty((8, 200), dtype=torch.int64, device="cuda")
    p = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    big = torch.zeros((4, 8192), device="cuda")
    rc, msg = rc_msg(p(big), 4, 8192, 0, 0, p(big), 1, 5, p(D), p(I), 0, 0, None)
    assert rc == _lib.E_UNSUPPORTED and "dim 8192 exceeds" in msg
    rc, msg = rc_msg(p(X), n, d, 0, 0, p(X), 5, 5, p(D), p(I), 0, 0, None)
    assert rc == _lib.E_UNSUPPORTED and "nq 5 outside 1..4" in msg
    rc, msg = rc_msg(p(X), n, d, 0, 0, p(X), 1, 129, p(D), p(I), 0, 0, None)
    assert rc == _lib.E_UNSUPPORTED and "k 129 outside 1..128" in msg
    odd = torch.zeros((10, 102), device="cuda")
    rc, msg = rc_msg(p(odd), 10, 102, 0, 0, p(odd), 1, 5, p(D), p(I), 0, 0, None)
    assert rc == _lib.E_UNSUPPORTED and "16-byte" in msg
    rc, msg = rc_msg(p(X), n, d, 0, 0, None, 1, 5, p(D), p(I), 0, 0, None)
    assert rc == _lib.E_INVALID
