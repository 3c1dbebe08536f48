"""Shared helpers for the parity tests: build matching oracle / CUDA sessions for a workload."""
from __future__ import annotations

import numpy as np
import torch

from pytorch_cpp_rl_b200 import api
from pytorch_cpp_rl_b200.workloads import Workload, synth_rollout


def make_ref_session(ref, w: Workload, init_seed=0, **overrides):
    kw = dict(algo=w.algo, base=w.base, head=w.head, T=w.T, N=w.N, obs_shape=w.obs_shape, num_actions=w.num_actions,
              hidden=w.hidden, storage_hidden=w.storage_hidden, normalize_obs=w.normalize_obs, num_epoch=w.num_epoch,
              num_mini_batch=w.num_mini_batch, clip_param=w.clip_param, actor_loss_coef=w.actor_loss_coef,
              value_loss_coef=w.value_loss_coef, entropy_coef=w.entropy_coef, learning_rate=w.learning_rate,
              epsilon=w.epsilon, alpha=w.alpha, kl_target=w.kl_target, init_seed=init_seed)
    kw.update(overrides)
    return ref.RefSession(**kw)


def fill_policy_outputs(sess, w: Workload, r):
    """value_predictions[:T] and action_log_probs from the oracle policy at its current weights (§8d)."""
    T, N = w.T, w.N
    hs = w.hidden if w.base == "gru" else 1
    if w.base == "gru":
        hxs = r["hidden_states"][0].reshape(N, hs)
    else:
        hxs = np.zeros((T * N, hs), np.float32)
    v, lp, _ = sess.evaluate_actions(r["observations"][:T].reshape((T * N,) + tuple(w.obs_shape)), hxs,
                                     r["masks"][:T].reshape(-1, 1), r["actions"].reshape(T * N, -1))
    r["value_predictions"][:T] = v.reshape(T, N, 1)
    r["action_log_probs"][:] = lp.reshape(T, N, 1)


def ref_load_rollout(sess, r):
    sess.set_rollout(r["observations"], r["hidden_states"], r["rewards"], r["value_predictions"], r["returns"],
                     r["action_log_probs"], r["actions"], r["masks"])


def make_gpu_side(w: Workload, r, params: np.ndarray, normalizer=None):
    space = api.ActionSpace("Discrete" if w.head == "categorical" else "Box", [w.num_actions])
    policy = api.Policy(space, w.obs_dim, normalize_observations=w.normalize_obs, hidden=w.hidden,
                        recurrent=(w.base == "gru"), obs_shape=w.obs_shape if w.base == "cnn" else None)
    policy.params.copy_(torch.from_numpy(params))
    if normalizer is not None and policy.observation_normalizer is not None:
        _, count, mean, var = normalizer
        n = policy.observation_normalizer
        n.rms.count.copy_(torch.from_numpy(count))
        n.rms.mean.copy_(torch.from_numpy(mean))
        n.rms.variance.copy_(torch.from_numpy(var))
    st = api.RolloutStorage(w.T, w.N, w.obs_shape, space, w.storage_hidden)
    st.observations.copy_(torch.from_numpy(r["observations"]))
    st.hidden_states.copy_(torch.from_numpy(r["hidden_states"]))
    st.rewards.copy_(torch.from_numpy(r["rewards"]))
    st.value_predictions.copy_(torch.from_numpy(r["value_predictions"]))
    st.returns.copy_(torch.from_numpy(r["returns"]))
    st.action_log_probs.copy_(torch.from_numpy(r["action_log_probs"]))
    st.actions.copy_(torch.from_numpy(r["actions"]))
    st.masks.copy_(torch.from_numpy(r["masks"]))
    return policy, st


def rel_err(a, b, floor=0.0):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor if floor > 0 else 1e-30)))


def draw_perms(seed: int, n: int, epochs: int) -> torch.Tensor:
    """The permutations the reference's generator draws after torch::manual_seed(seed): one randperm per epoch."""
    torch.manual_seed(seed)
    return torch.stack([torch.randperm(n, dtype=torch.int64) for _ in range(epochs)])
