"""GPU parity tests of the write side of the rollout buffer: RolloutStorage::insert,
after_update, set_first_observation and the concatenating constructor as single fused-copy launches through the
C ABI (cpprl_b200_storage_*), bit-exact against the oracle's restatement of src/storage.cpp:51-111,176-193,220-223
(pinned by the reference's own scenarios in tests/test_oracle_cpu.py)."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import cpprl_oracle as O
from pytorch_cpp_rl_b200 import api, cabi

pytestmark = pytest.mark.gpu

FIELDS = ("observations", "hidden_states", "rewards", "value_predictions", "returns", "action_log_probs", "actions",
          "masks")


def assert_same(st: api.RolloutStorage, ro: O.Rollout):
    for name in FIELDS:
        got, want = getattr(st, name).cpu().numpy(), getattr(ro, name)
        assert got.dtype == want.dtype and got.shape == want.shape, name
        assert np.array_equal(got, want), name  # bit-exact: byte movement only


def random_step(rng, N, obs_shape, discrete, A, Hs):
    action = rng.integers(0, A, (N, 1)) if discrete else rng.standard_normal((N, A)).astype(np.float32)
    f = lambda *s: rng.standard_normal(s).astype(np.float32)
    return (f(N, *obs_shape), f(N, Hs), action, f(N, 1), f(N, 1), f(N, 1),
            (rng.random((N, 1)) < 0.9).astype(np.float32))


# shapes chosen to hit the 16-byte path (aligned rows), the 4-byte path (odd N / odd widths) and image observations
CASES = [
    dict(T=3, N=4, obs_shape=(5, 2), discrete=True, A=3, Hs=10),      # the reference's own test shape, storage.cpp:298
    dict(T=5, N=7, obs_shape=(3,), discrete=False, A=3, Hs=1),        # nothing 16-byte aligned
    dict(T=4, N=64, obs_shape=(17,), discrete=False, A=6, Hs=1),      # cfg3-shaped rows
    dict(T=6, N=8, obs_shape=(8,), discrete=True, A=4, Hs=64),        # cfg1 / cfg4-shaped rows
    dict(T=2, N=16, obs_shape=(4, 84, 84), discrete=True, A=6, Hs=1),  # cfg5 frames
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"T{c['T']}N{c['N']}D{'x'.join(map(str, c['obs_shape']))}")
def test_insert_after_update_bit_exact(case):
    T, N, obs_shape, discrete, A, Hs = (case[k] for k in ("T", "N", "obs_shape", "discrete", "A", "Hs"))
    rng = np.random.default_rng(T * 1000 + N)
    space = api.ActionSpace("Discrete" if discrete else "Box", [A])
    st = api.RolloutStorage(T, N, obs_shape, space, Hs)
    ro = O.new_rollout(T, N, obs_shape, discrete, A, Hs)
    first = rng.standard_normal((N, *obs_shape)).astype(np.float32)
    st.set_first_observation(torch.from_numpy(first).cuda())
    ro.observations[0] = first
    launches = cabi.load().cpprl_b200_launch_count()
    step = 0
    for _ in range(2 * T + 1):  # wraps around: step = (step + 1) % T
        row = random_step(rng, N, obs_shape, discrete, A, Hs)
        st.insert(*[torch.from_numpy(x).cuda() for x in row])
        step = O.rollout_insert(ro, step, *row)
        assert st.step == step
    assert cabi.load().cpprl_b200_launch_count() - launches == 2 * T + 1  # ONE launch per insert
    assert_same(st, ro)
    st.after_update()
    O.rollout_after_update(ro)
    assert_same(st, ro)


def test_insert_converts_like_copy_():
    """Tensor::copy_ semantics the reference relies on (storage.cpp:183-189 and its tests): float actions into an
    int64 buffer, broadcast of a [N] / scalar-shaped input."""
    st = api.RolloutStorage(3, 4, (5,), api.ActionSpace("Discrete", [3]), 2)
    ro = O.new_rollout(3, 4, (5,), True, 1, 2)
    rng = np.random.default_rng(3)
    obs, hid = rng.random((4, 5), np.float32), rng.random((4, 2), np.float32)
    act = rng.integers(1, 3, (4, 1)).astype(np.float32)
    one = np.ones((1, 1), np.float32)
    st.insert(*[torch.from_numpy(x).cuda() for x in (obs, hid, act, one, one * 2, one * 3, one * 0)])
    O.rollout_insert(ro, 0, obs, hid, act, one, one * 2, one * 3, one * 0)
    assert_same(st, ro)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        st.insert(*[torch.from_numpy(x) for x in (obs, hid, act, one, one, one, one)])


@pytest.mark.parametrize("widths", [(1, 1, 1, 1, 1), (3, 5, 8), (4, 4)])
def test_concatenate_bit_exact(widths):
    """src/storage.cpp:51-104 / its test :504-581: storages of differing environment counts joined along axis 1."""
    rng = np.random.default_rng(sum(widths))
    T, obs_shape, Hs = 3, (4,), 5
    space = api.ActionSpace("Discrete", [3])
    gpu_parts, ref_parts = [], []
    for n in widths:
        st, ro = api.RolloutStorage(T, n, obs_shape, space, Hs), O.new_rollout(T, n, obs_shape, True, 1, Hs)
        step = 0
        for _ in range(T):
            row = random_step(rng, n, obs_shape, True, 3, Hs)
            st.insert(*[torch.from_numpy(x).cuda() for x in row])
            step = O.rollout_insert(ro, step, *row)
        st.returns.copy_(torch.from_numpy(rng.standard_normal((T + 1, n, 1)).astype(np.float32)))
        ro.returns[:] = st.returns.cpu().numpy()
        gpu_parts.append(st)
        ref_parts.append(ro)
    launches = cabi.load().cpprl_b200_launch_count()
    joined = api.RolloutStorage.concatenate(gpu_parts)
    assert cabi.load().cpprl_b200_launch_count() - launches == len(widths)  # one launch per part, all eight fields
    assert_same(joined, O.rollout_concatenate(ref_parts))
    assert joined.step == 0


def test_concatenate_image_shards_round_trip():
    """The scatter of per-GPU shards into a global rollout: cfg5-shaped frames, two shards, then an update-path call
    (compute_returns) on the joined storage to show the buffers are usable as they are."""
    rng = np.random.default_rng(5)
    T, obs_shape = 2, (4, 84, 84)
    space = api.ActionSpace("Discrete", [6])
    parts, refs = [], []
    for n in (8, 8):
        st = api.RolloutStorage(T, n, obs_shape, space, 1)
        st.observations.copy_(torch.from_numpy(np.floor(rng.random((T + 1, n, *obs_shape), np.float32) * 256)))
        st.rewards.copy_(torch.from_numpy(rng.standard_normal((T, n, 1)).astype(np.float32)))
        st.value_predictions.copy_(torch.from_numpy(rng.standard_normal((T + 1, n, 1)).astype(np.float32)))
        parts.append(st)
        ro = O.new_rollout(T, n, obs_shape, True, 1, 1)
        for name in FIELDS:
            getattr(ro, name)[:] = getattr(st, name).cpu().numpy()
        refs.append(ro)
    joined = api.RolloutStorage.concatenate(parts)
    want = O.rollout_concatenate(refs)
    assert_same(joined, want)
    nv = rng.standard_normal((16, 1)).astype(np.float32)
    joined.compute_returns(torch.from_numpy(nv).cuda(), True, 0.99, 0.95)
    O.compute_returns(want.rewards, want.value_predictions, want.masks, want.returns, nv, True, 0.99, 0.95)
    assert np.array_equal(joined.returns.cpu().numpy(), want.returns)


def test_storage_argument_errors():
    lib = cabi.load()
    st = api.RolloutStorage(3, 2, (4,), api.ActionSpace("Discrete", [3]), 5)
    d = st.storage_desc()
    p = C.c_void_p(st.masks.data_ptr())
    assert lib.cpprl_b200_storage_insert(C.byref(d), 3, p, p, p, p, p, p, p, None) == cabi.EINVAL  # step outside [0, T)
    assert b"step" in lib.cpprl_b200_last_error()
    assert lib.cpprl_b200_storage_insert(C.byref(d), 0, None, p, p, p, p, p, p, None) == cabi.EINVAL
    other = api.RolloutStorage(4, 2, (4,), api.ActionSpace("Discrete", [3]), 5).storage_desc()
    assert lib.cpprl_b200_storage_concat(C.byref(d), C.byref(other), 0, None) == cabi.EINVAL  # T differs
    same = api.RolloutStorage(3, 2, (4,), api.ActionSpace("Discrete", [3]), 5).storage_desc()
    assert lib.cpprl_b200_storage_concat(C.byref(d), C.byref(same), 1, None) == cabi.EINVAL  # columns fall outside


def test_rollout_step_is_graph_capturable():
    """act -> insert as one CUDA graph: the whole collection step replays without host work."""
    torch.manual_seed(0)
    N, D, A, T = 64, 17, 6, 4
    space = api.ActionSpace("Box", [A])
    policy = api.Policy(space, D, normalize_observations=False)
    policy.params.copy_(torch.randn_like(policy.params) * 0.1)
    st = api.RolloutStorage(T, N, (D,), space, 1)
    obs = torch.randn(N, D, device="cuda")
    hxs, masks, reward = torch.zeros(N, 1, device="cuda"), torch.ones(N, 1, device="cuda"), torch.randn(N, 1, device="cuda")
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        value, action, logp, _ = policy.act(obs, hxs, masks, seed=7, offset=0)  # warm-up outside capture
        st.insert(obs, hxs, action, logp, value, reward, masks)
        st.step = 0
        side.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            value, action, logp, _ = policy.act(obs, hxs, masks, seed=7, offset=0)
            st.insert(obs, hxs, action, logp, value, reward, masks)
    torch.cuda.current_stream().wait_stream(side)
    st.observations.zero_(), st.actions.zero_(), st.value_predictions.zero_()
    obs.copy_(torch.randn(N, D, device="cuda"))  # new inputs, same addresses
    graph.replay()
    torch.cuda.synchronize()
    want_value, want_action, want_logp, _ = policy.act(obs, hxs, masks, seed=7, offset=0)
    assert torch.equal(st.observations[1], obs)
    assert torch.equal(st.actions[0], want_action.view(N, A)) and torch.equal(st.value_predictions[0], want_value.view(N, 1))
    assert torch.equal(st.action_log_probs[0], want_logp.view(N, 1)) and torch.equal(st.rewards[0], reward)


# ---- bootstrap value + returns as one call ---------------------------------------------------
@pytest.mark.parametrize("cfg", ["cfg1", "cfg3", "cfg4", "cnn_small"])
@pytest.mark.parametrize("use_gae", [True, False])
def test_bootstrap_returns_equals_two_calls(cfg, use_gae):
    """compute_returns(get_values(obs[T], hxs[T], masks[T])) (gym_client.cpp:266-275) fused into one C-ABI call gives
    the bits of the two separate calls, and get_values gives the bits of evaluate_actions' value."""
    from pytorch_cpp_rl_b200.workloads import CONFIGS, synth_rollout
    from tests import helpers as H
    import dataclasses
    if cfg == "cnn_small":
        w = dataclasses.replace(CONFIGS["cfg5"], T=4, N=8)
    else:
        w = CONFIGS[cfg]
        w = dataclasses.replace(w, T=min(w.T, 16), N=min(w.N, 64))
    r = synth_rollout(w, 3)
    rng = np.random.default_rng(4)
    n_params = api.Policy(api.ActionSpace("Discrete" if w.head == "categorical" else "Box", [w.num_actions]), w.obs_dim,
                          normalize_observations=w.normalize_obs, hidden=w.hidden, recurrent=(w.base == "gru"),
                          obs_shape=w.obs_shape if w.base == "cnn" else None).params.numel()
    params = (rng.standard_normal(n_params) * 0.05).astype(np.float32)
    policy, st = H.make_gpu_side(w, r, params)
    policy2, st2 = H.make_gpu_side(w, r, params)
    T, N = w.T, w.N
    obs_T, hxs_T, masks_T = st.observations[T], st.hidden_states[T], st.masks[T]
    next_value = policy.get_values(obs_T, hxs_T, masks_T)
    if w.head == "categorical":
        placeholder = torch.zeros(N, 1, dtype=torch.int64, device="cuda")
    else:
        placeholder = torch.zeros(N, w.num_actions, device="cuda")
    want_value = policy.evaluate_actions(obs_T.contiguous(), hxs_T.contiguous() if w.base == "gru" else None,
                                         masks_T.contiguous() if w.base == "gru" else None, placeholder)[0]
    assert torch.equal(next_value, want_value)
    st.compute_returns(next_value, use_gae, w.gamma, w.tau)
    st2.bootstrap_returns(policy2, use_gae, w.gamma, w.tau)
    assert torch.equal(st.returns, st2.returns)
    assert torch.equal(st.value_predictions, st2.value_predictions)
    assert float(st2.returns[:T].abs().sum()) > 0


# ---- the combining constructor of the observation normaliser ------------------------------------
def test_observation_normalizer_combine():
    """observation_normalizer.cpp:19-67 and its test :145-190 on device normalisers (1, 2 and 3 kernel updates)."""
    rng = np.random.default_rng(8)
    gpu, ref = [], []
    for i in range(3):
        n, rms = api.ObservationNormalizer(3), O.RunningMeanStd(3)
        for _ in range(i + 1):
            x = rng.random((1, 3), np.float32)
            n.update(torch.from_numpy(x).cuda())
            rms.update(x)
        gpu.append(n)
        ref.append((10.0, rms))
    combined = api.ObservationNormalizer.combine(gpu)
    clip, want = O.combine_normalizers(ref)
    np.testing.assert_allclose(combined.rms.mean.cpu().numpy(), want.mean, rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(combined.rms.variance.cpu().numpy(), want.variance, rtol=1e-5, atol=1e-7)
    assert combined.get_step_count() == 6 and float(combined.clip) == float(clip)
    y = combined.process_observation(torch.from_numpy(rng.random((4, 3), np.float32)).cuda())  # usable on the device
    assert y.shape == (4, 3) and bool(torch.isfinite(y).all())
