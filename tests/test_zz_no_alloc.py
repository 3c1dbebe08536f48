"""cpprl_b200_advantages_noalloc -- the allocation-free variant of cpprl_b200_advantages (scratch in a file-scope device
array of the library; no counterpart in the reference) -- against the
existing entry point, through the C ABI, on two calls in a row with different inputs (the shared scratch keeps the
previous call's contents).  Same kernels and launch shape, so the outputs must be bit-identical; the assertion uses the
1e-5 relative tolerance of the other advantage tests as its bound and reports exact equality on top.

Collected last (file name) so that a failure here cannot hide other tests under `pytest -x`; it needs nothing from
the other test files and passes alone in a fresh process as well as after them."""
import ctypes as C

import numpy as np
import pytest
import torch

from pytorch_cpp_rl_b200 import cabi

pytestmark = pytest.mark.gpu

MAX_ROWS = 1 << 22  # CPPRL_B200_ADVANTAGES_NOALLOC_MAX_ROWS (include/cpprl_b200.h)


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _run(fn, returns, values, T, N, normalize):
    adv = torch.full((T, N), float("nan"), device="cuda")
    moments = torch.full((4,), float("nan"), dtype=torch.float64, device="cuda")
    rc = fn(_ptr(returns), _ptr(values), T, N, normalize, _ptr(adv), _ptr(moments), None)
    assert rc == 0, cabi.load().cpprl_b200_last_error().decode()
    torch.cuda.synchronize()
    return adv.cpu().numpy(), moments.cpu().numpy()


@pytest.mark.parametrize("T,N", [(128, 8), (256, 1024), (5, 4099), (1, 2)])
def test_noalloc_matches_allocating_entry_point_twice_in_a_row(T, N):
    lib = cabi.load()
    g = torch.Generator(device="cuda").manual_seed(1234 + T + N)
    for call in range(2):  # second call: new inputs over the scratch the first call left behind
        returns = torch.randn(T + 1, N, generator=g, device="cuda") * (1.0 + call)
        values = torch.randn(T + 1, N, generator=g, device="cuda") + 0.25 * call
        for normalize in (1, 0):
            want_adv, want_m = _run(lib.cpprl_b200_advantages, returns, values, T, N, normalize)
            got_adv, got_m = _run(lib.cpprl_b200_advantages_noalloc, returns, values, T, N, normalize)
            np.testing.assert_allclose(got_adv, want_adv, rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(got_m, want_m, rtol=1e-12)
            assert np.array_equal(got_adv, want_adv) and np.array_equal(got_m, want_m), "same kernels, same launch shape: expected identical bits"
            # and against a float64 restatement of ppo.cpp:51-57
            a = (returns[:T].double() - values[:T].double()).cpu().numpy()
            ref = (a - a.mean()) / (a.std(ddof=1) + 1e-5) if normalize else a
            if T * N > 2:
                np.testing.assert_allclose(got_adv, ref, rtol=1e-5, atol=2e-6)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    lib = cabi.load()
    x = torch.zeros(8, device="cuda")
    m = torch.zeros(4, dtype=torch.float64, device="cuda")
    rc = lib.cpprl_b200_advantages_noalloc(_ptr(x), _ptr(x), MAX_ROWS // 1024 + 1, 1024, 1, _ptr(x), _ptr(m), None)
    assert rc == cabi.EUNSUPPORTED
    msg = lib.cpprl_b200_last_error().decode()
    assert "exceeds" in msg and str(MAX_ROWS) in msg
    assert lib.cpprl_b200_advantages_noalloc(_ptr(x), _ptr(x), 0, 4, 1, _ptr(x), _ptr(m), None) == cabi.EINVAL
    assert lib.cpprl_b200_advantages_noalloc(None, _ptr(x), 2, 2, 1, _ptr(x), _ptr(m), None) == cabi.EINVAL


def test_noalloc_at_the_limit():
    lib = cabi.load()
    T, N = 1024, MAX_ROWS // 1024
    g = torch.Generator(device="cuda").manual_seed(7)
    returns = torch.randn(T + 1, N, generator=g, device="cuda")
    values = torch.randn(T + 1, N, generator=g, device="cuda")
    want_adv, want_m = _run(lib.cpprl_b200_advantages, returns, values, T, N, 1)
    got_adv, got_m = _run(lib.cpprl_b200_advantages_noalloc, returns, values, T, N, 1)
    assert np.array_equal(got_adv, want_adv) and np.array_equal(got_m, want_m)
