"""The library's host permutation generator against torch.randperm on the CPU generator (the reference's
FeedForwardGenerator / RecurrentGenerator draw torch::randperm(n, kLong) from it: feed_forward_generator.cpp:31-34,
recurrent_generator.cpp:38): bit-identical permutations AND a bit-identical generator state afterwards."""
import ctypes as C
import time

import numpy as np
import pytest
import torch

from pytorch_cpp_rl_b200 import cabi


def lib_randperm(n, count, threads=0):
    lib = cabi.load()
    state = torch.get_rng_state()
    assert state.numel() == lib.cpprl_b200_rng_state_bytes()
    out = torch.empty(count, n, dtype=torch.int64)
    rc = lib.cpprl_b200_host_randperm(C.c_void_p(state.data_ptr()), n, count, C.c_void_p(out.data_ptr()), threads)
    assert rc == 0, cabi.last_error()
    torch.set_rng_state(state)
    return out


@pytest.mark.parametrize("seed,n,count", [(0, 1, 3), (1, 2, 4), (2, 8, 5), (3, 623, 3), (4, 624, 2), (5, 625, 4), (6, 1024, 4),
                                          (7, 1249, 7), (1000, 262144, 3), (12345, 99991, 10)])
def test_matches_torch_randperm(seed, n, count):
    torch.manual_seed(seed)
    want = torch.stack([torch.randperm(n, dtype=torch.int64) for _ in range(count)])
    after_want = torch.rand(5)
    torch.manual_seed(seed)
    got = lib_randperm(n, count)
    after_got = torch.rand(5)
    assert torch.equal(got, want)
    assert torch.equal(after_got, after_want), "generator state after the draws differs from torch's"


def test_mid_stream_state_and_single_thread():
    # a generator that is not freshly seeded (mid-block position) and the one-worker path
    torch.manual_seed(77)
    torch.rand(1000)
    torch.randint(0, 10, (333,))
    s = torch.get_rng_state()
    want = torch.stack([torch.randperm(5000, dtype=torch.int64) for _ in range(6)])
    tail_want = torch.randperm(100)
    torch.set_rng_state(s)
    got = lib_randperm(5000, 6, threads=2)
    tail_got = torch.randperm(100)
    assert torch.equal(got, want) and torch.equal(tail_got, tail_want)


def test_empty_and_errors():
    lib = cabi.load()
    torch.manual_seed(3)
    before = torch.rand(3)
    torch.manual_seed(3)
    out = lib_randperm(0, 4)       # randperm(0) draws nothing
    assert out.numel() == 0 and torch.equal(torch.rand(3), before)
    state = torch.get_rng_state()
    big = (2 ** 32 - 1) // 20      # torch switches to 64-bit draws here: not this algorithm
    dummy = torch.empty(8, dtype=torch.int64)  # refused before anything is written
    assert lib.cpprl_b200_host_randperm(C.c_void_p(state.data_ptr()), big, 1, C.c_void_p(dummy.data_ptr()), 0) == cabi.EUNSUPPORTED
    bad = torch.zeros_like(state)  # not a seeded generator state
    assert lib.cpprl_b200_host_randperm(C.c_void_p(bad.data_ptr()), 8, 1, C.c_void_p(dummy.data_ptr()), 0) == cabi.EINVAL


def _stream(seed, n, count, used, reseed_before=None, reseed_seed=None):
    """Consecutive updates on one permutation stream (the way cpprl_b200_ppo_update_rng drives it)."""
    lib = cabi.load()
    torch.manual_seed(seed)
    state = torch.get_rng_state()
    reseed_state = None
    if reseed_seed is not None:
        torch.manual_seed(reseed_seed)
        reseed_state = torch.get_rng_state()
    used_a = np.asarray(used, np.int32)
    flags = np.asarray(reseed_before if reseed_before is not None else [0] * len(used), np.int32)
    out = torch.empty(int(used_a.sum()), n, dtype=torch.int64)
    restarts = C.c_longlong(0)
    rc = lib.cpprl_b200_debug_perm_stream(C.c_void_p(state.data_ptr()), C.c_void_p(reseed_state.data_ptr()) if reseed_state is not None else None,
                                          n, count, len(used), C.c_void_p(used_a.ctypes.data), C.c_void_p(flags.ctypes.data),
                                          C.c_void_p(out.data_ptr()), C.byref(restarts))
    assert rc == 0, cabi.last_error()
    return out, state, restarts.value


def test_stream_continues_across_updates():
    # five full updates of four permutations each: the stream starts once, every later update finds its permutations
    # already drawn ahead, and what comes out is torch's stream
    seed, n, count = 21, 3001, 4
    got, state, restarts = _stream(seed, n, count, [count] * 5)
    torch.manual_seed(seed)
    want = torch.stack([torch.randperm(n, dtype=torch.int64) for _ in range(5 * count)])
    after_want = torch.rand(4)
    assert torch.equal(got, want)
    assert restarts == 1
    torch.set_rng_state(state)
    assert torch.equal(torch.rand(4), after_want), "state handed back after the last update differs from torch's"


def test_stream_after_early_stops_and_foreign_draws():
    # updates that stop early use fewer permutations: the next update continues from where the LAST USED one left the
    # generator (the unused ones drawn ahead are the right ones for it); a state nobody recorded restarts the stream
    seed, n, count = 5, 777, 5
    used = [5, 2, 5, 0, 1, 5, 3]
    got, _, restarts = _stream(seed, n, count, used)
    torch.manual_seed(seed)
    want = torch.stack([torch.randperm(n, dtype=torch.int64) for _ in range(sum(used))])
    assert torch.equal(got, want)
    assert restarts == 2, "only the first update and the one after the update that drew nothing restart the stream"
    # someone reseeds the generator before the third update
    used = [3, 3, 3, 3]
    got, _, restarts = _stream(seed, n, 3, used, reseed_before=[0, 0, 1, 0], reseed_seed=99)
    torch.manual_seed(seed)
    a = [torch.randperm(n, dtype=torch.int64) for _ in range(6)]
    torch.manual_seed(99)
    b = [torch.randperm(n, dtype=torch.int64) for _ in range(6)]
    assert torch.equal(got, torch.stack(a + b))
    assert restarts == 2


def test_faster_than_torch():
    n, count = 262144, 10
    torch.manual_seed(0)
    t0 = time.perf_counter()
    for _ in range(count):
        torch.randperm(n, dtype=torch.int64)
    t_torch = time.perf_counter() - t0
    torch.manual_seed(0)
    lib_randperm(n, count)
    t0 = time.perf_counter()
    lib_randperm(n, count)
    t_lib = time.perf_counter() - t0
    print(f"torch.randperm x{count}: {t_torch * 1e3:.1f} ms; library: {t_lib * 1e3:.1f} ms")
    assert t_lib < t_torch
