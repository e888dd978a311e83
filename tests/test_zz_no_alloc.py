"""ll_topk_static: the allocation-free top-k entry point (include/ll_gpu.h) against the existing one, ll_local_topk.

No counterpart in the reference; added for callers that may not call cudaMalloc.  The file
name sorts last on purpose: a failure here must not hide another test.
All comparisons are bit-exact (integer / index work, the bar the other top-k tests use)."""
import ctypes as C

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ll(built):
    import lunar_lander_b200 as m
    return m


def _static_topk(torch, abi, fit_dev, k, first_global=0):
    """-> (fitness[k], index[k]) from ll_topk_static on torch's current stream; inputs and outputs are the caller's device buffers."""
    out = torch.zeros(2 * k, dtype=torch.int64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = abi.lib.ll_topk_static(-1, stream, C.c_void_p(fit_dev.data_ptr()), fit_dev.numel(), first_global, k, C.c_void_p(out.data_ptr()))
    abi.check(rc)
    torch.cuda.synchronize()
    raw = out.cpu().numpy().reshape(k, 2)
    return raw[:, 0].copy().view(np.float64), raw[:, 1].astype(np.uint64)


def test_static_topk_matches_local_topk_on_two_calls_in_a_row(ll):
    import torch
    from lunar_lander_b200 import _abi
    sizes = util.TOPOLOGIES["8_16_8"]
    pop, top = 5000, 37
    eng = ll.Engine(util.cfg_for(sizes), population=pop, top=top, precision=64, mode=0)
    eng.create_population(True)  # random nets: many crash identically -> exact fitness ties, resolved by index
    spad, lpad, d = util.layouts(8)
    free_before = None
    for call, (lo, hi, max_steps) in enumerate(((0, 2, 300), (2, 5, 200))):
        fit, _, _, _ = eng.evaluate(spad[lo:hi], lpad[lo:hi], d[lo:hi], max_steps)
        assert len(np.unique(fit)) < pop
        want_f, want_i = eng.local_topk()  # the existing entry point (context-owned, cudaMalloc'ed scratch)
        fit_dev = torch.tensor(fit, dtype=torch.float64, device="cuda")
        if call == 1:
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
        got_f, got_i = _static_topk(torch, _abi, fit_dev, top)
        if call == 1:  # the second call: nothing left to load lazily, so device memory in use must not have moved
            assert free_before - torch.cuda.mem_get_info()[0] < (1 << 20)
        assert np.array_equal(got_i, want_i), call
        assert np.array_equal(got_f.view(np.uint64), want_f.view(np.uint64)), call
    eng.close()


def test_static_topk_forgets_the_previous_call_and_honours_first_global(ll):
    import torch
    from lunar_lander_b200 import _abi
    rng = np.random.default_rng(7)
    # a large call first (leaves 2^20 keys, 4096 candidates and a finished select state behind), then small ones over the same array
    for n, k, first in ((1 << 20, 4096, 0), (1000, 1000, 123456789), (3, 1, 0), (70001, 512, 1 << 40)):
        fit = rng.integers(-50, 50, size=n).astype(np.float64) * 0.25  # heavy ties
        fit[rng.integers(0, n, size=max(1, n // 100))] = -0.0          # -0.0 ties with +0.0 (the reference's `<`)
        order = np.lexsort((np.arange(n), fit + 0.0))[:k]
        got_f, got_i = _static_topk(torch, _abi, torch.tensor(fit, device="cuda"), k, first)
        assert np.array_equal(got_i, order.astype(np.uint64) + np.uint64(first)), (n, k)
        assert np.array_equal(got_f, fit[order]), (n, k)


def test_static_topk_rejects_sizes_beyond_the_static_array(ll):
    import torch
    from lunar_lander_b200 import _abi
    buf = torch.zeros(16, dtype=torch.float64, device="cuda")
    out = torch.zeros(32, dtype=torch.int64, device="cuda")
    fn = _abi.lib.ll_topk_static
    args = lambda n, k: (-1, None, C.c_void_p(buf.data_ptr()), n, 0, k, C.c_void_p(out.data_ptr()))
    assert fn(*args((1 << 20) + 1, 4)) == 1  # LL_EINVAL, before anything is launched (the pointer is never read)
    assert "LL_TOPK_STATIC_MAX" in _abi.lib.ll_last_error(None).decode()
    assert fn(*args(16, 17)) == 1 and fn(*args(16, 0)) == 1 and fn(*args(0, 1)) == 1
    assert fn(-1, None, None, 16, 0, 4, C.c_void_p(out.data_ptr())) == 1
    assert fn(*args(16, 4)) == 0  # and the entry point still works afterwards
    torch.cuda.synchronize()
