"""yoda_normalize_scores_noalloc: the allocation-free NormalizeScore entry point.

It serves callers that cannot call cudaMalloc -- a scheduler thread working while a FULL-mode batch
owns most of HBM, a host with a fixed device-memory budget, code without a yoda_ctx -- by keeping its
scratch in a fixed array inside the library.  This file is collected last on purpose: if it fails, an
`-x` run has already executed every other test.

The check: through the C ABI, its output must equal the existing entry point's
(yoda_normalize_scores) exactly -- integers, so the tolerance is zero, as everywhere in this suite --
on two calls in a row.  The second call is shorter and holds different values, so anything the first
call left in the shared scratch would show.  The input/output device buffers are the caller's (torch
tensors here); the entry point itself allocates nothing.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_SCORES = 1 << 20


def _noalloc(L, d_in, d_out, n, stream=None):
    return L.yoda_normalize_scores_noalloc(C.c_void_p(d_in), C.c_void_p(d_out), n, C.c_void_p(stream) if stream else None)


def test_noalloc_matches_normalize_scores_twice_in_a_row():
    import torch
    from oracle import oracle as ora
    from yoda_scheduler_b200 import _lib, api

    L = _lib.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    rng = np.random.default_rng(42)
    first = rng.integers(-(2**40), 2**40, 300_000, dtype=np.int64)     # 293 scratch slots
    second = rng.integers(5_000, 9_000, 1_500, dtype=np.int64)          # 2 slots, narrower range
    with api.Context(8) as ctx:
        for scores in (first, second):
            want = ctx.normalize_scores(scores)                           # existing entry point
            d_in = torch.from_numpy(scores).to(dev)
            d_out = torch.empty_like(d_in)
            torch.cuda.synchronize()
            api.check(_noalloc(L, d_in.data_ptr(), d_out.data_ptr(), len(scores)))
            torch.cuda.synchronize()
            assert np.array_equal(d_out.cpu().numpy(), want)
            assert d_out.cpu().numpy().tolist() == ora.normalize_score(scores.tolist())   # and the CPU oracle's
            # in place, on a caller-chosen stream
            st = torch.cuda.Stream(device=dev)
            st.wait_stream(torch.cuda.current_stream(dev))
            api.check(_noalloc(L, d_in.data_ptr(), d_in.data_ptr(), len(scores), st.cuda_stream))
            st.synchronize()
            assert np.array_equal(d_in.cpu().numpy(), want)
        # edge values of the Go arithmetic: all equal -> 100, single element, negative-only lists
        for scores in ([7, 7, 7], [5], [-5, -5], [0, 10, 3, 7], [-(2**63), 2**63 - 1, 0]):
            a = np.array(scores, dtype=np.int64)
            d = torch.from_numpy(a).to(dev)
            api.check(_noalloc(L, d.data_ptr(), d.data_ptr(), len(a)))
            torch.cuda.synchronize()
            assert np.array_equal(d.cpu().numpy(), ctx.normalize_scores(a))
            assert d.cpu().numpy().tolist() == ora.normalize_score(list(scores))           # wrap cases included


def test_noalloc_rejects_sizes_beyond_the_scratch():
    import torch
    from yoda_scheduler_b200 import _lib, api

    L = _lib.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    d = torch.zeros(16, dtype=torch.int64, device=dev)
    assert _noalloc(L, d.data_ptr(), d.data_ptr(), MAX_SCORES + 1) == -3        # YODA_E_RANGE
    assert b"exceed the fixed scratch" in L.yoda_last_error()
    assert _noalloc(L, d.data_ptr(), d.data_ptr(), 0) == -1                     # YODA_E_ARG
    assert _noalloc(L, 0, d.data_ptr(), 4) == -1
    # the largest accepted size works
    big = torch.arange(MAX_SCORES, dtype=torch.int64, device=dev)
    api.check(_noalloc(L, big.data_ptr(), big.data_ptr(), MAX_SCORES))
    torch.cuda.synchronize()
    got = big.cpu().numpy()
    assert got[0] == 0 and got[-1] == 100 and got[MAX_SCORES // 2] == (MAX_SCORES // 2) * 100 // (MAX_SCORES - 1)
