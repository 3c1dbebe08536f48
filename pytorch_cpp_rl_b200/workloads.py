"""BASELINE.json workloads and the synthetic rollouts they are measured on (§8d).

Synthetic inputs are generated on the host from a seed so the oracle and the CUDA path see identical
tensors.  `value_predictions[:T]` and `action_log_probs` are NOT produced here: §8d wants them to be the
policy's own outputs at its initial weights (ratio = 1 on the first minibatch), so the caller fills them
from an evaluate_actions pass (the oracle's in tests, the CUDA path's in bench.py).
"""
from __future__ import annotations

from dataclasses import dataclass, replace
from typing import Dict, Tuple

import numpy as np


@dataclass(frozen=True)
class Workload:
    name: str
    description: str
    algo: str                 # "ppo" | "a2c"
    base: str                 # "mlp" | "gru" | "cnn"
    head: str                 # "categorical" | "normal"
    T: int
    N: int
    obs_shape: Tuple[int, ...]
    num_actions: int
    hidden: int = 64
    storage_hidden: int = 1
    normalize_obs: bool = True
    num_epoch: int = 4
    num_mini_batch: int = 4
    use_gae: bool = True
    gamma: float = 0.99
    tau: float = 0.95
    clip_param: float = 0.2
    actor_loss_coef: float = 1.0
    value_loss_coef: float = 0.5
    entropy_coef: float = 0.01
    learning_rate: float = 3e-4
    epsilon: float = 1e-8
    alpha: float = 0.99
    kl_target: float = 1e9    # throughput runs disable the KL stop (§8d); parity has a case with 0.01

    @property
    def rows(self) -> int:
        return self.T * self.N

    @property
    def obs_dim(self) -> int:
        return int(np.prod(self.obs_shape))

    @property
    def action_cols(self) -> int:
        return 1 if self.head == "categorical" else self.num_actions

    @property
    def optimizer_steps(self) -> int:
        if self.algo == "a2c":
            return 1
        return self.num_epoch * ((self.rows) // (self.rows // self.num_mini_batch))

    def scaled(self, **kw) -> "Workload":
        return replace(self, **kw)


CONFIGS: Dict[str, Workload] = {
    # BASELINE.json configs[0]
    "cfg1": Workload("cfg1", "PPO discrete MLPBase 64x2, LunarLander-shaped rollout (obs 8, 4 actions, 8 envs x 128 steps), "
                     "4 epochs x 4 minibatches", "ppo", "mlp", "categorical", T=128, N=8, obs_shape=(8,), num_actions=4,
                     storage_hidden=64, num_epoch=4, num_mini_batch=4),
    # configs[1]
    "cfg2": Workload("cfg2", "A2C discrete MLPBase at scale: 4096 envs x 5 steps, obs 8, 4 actions", "a2c", "mlp",
                     "categorical", T=5, N=4096, obs_shape=(8,), num_actions=4, use_gae=False, learning_rate=7e-4,
                     num_epoch=1, num_mini_batch=1),
    # configs[2]
    "cfg3": Workload("cfg3", "PPO continuous DiagGaussian MLP, MuJoCo-shaped obs 17 / act 6, 1024 envs x 256 steps, "
                     "10 epochs x 32 minibatches", "ppo", "mlp", "normal", T=256, N=1024, obs_shape=(17,), num_actions=6,
                     num_epoch=10, num_mini_batch=32),
    # configs[3]
    "cfg4": Workload("cfg4", "PPO recurrent GRU policy (hidden 64, RecurrentGenerator), 512 envs x 128 steps, obs 8",
                     "ppo", "gru", "categorical", T=128, N=512, obs_shape=(8,), num_actions=4, storage_hidden=64,
                     normalize_obs=False, num_epoch=4, num_mini_batch=512),
    # configs[4]
    "cfg5": Workload("cfg5", "PPO Atari CNNBase 84x84x4 frames, 6 actions, 256 envs x 128 steps, 4 epochs x 4 minibatches",
                     "ppo", "cnn", "categorical", T=128, N=256, obs_shape=(4, 84, 84), num_actions=6, hidden=512,
                     normalize_obs=False, num_epoch=4, num_mini_batch=4),
}


def synth_rollout(w: Workload, seed: int) -> Dict[str, np.ndarray]:
    """Synthetic rollout of §8d: obs ~ N(0,1) (pixels floor(U[0,256)) for the CNN), rewards ~ N(0,1),
    masks = Bernoulli(0.98) with row 0 all ones, random actions, zero hidden states, next_value ~ N(0,1)."""
    rng = np.random.default_rng(seed)
    T, N = w.T, w.N
    if w.base == "cnn":
        obs = np.floor(rng.random((T + 1, N) + tuple(w.obs_shape), dtype=np.float32) * 256.0).astype(np.float32)
    else:
        obs = rng.standard_normal((T + 1, N) + tuple(w.obs_shape), dtype=np.float32)
    rewards = rng.standard_normal((T, N, 1), dtype=np.float32)
    masks = (rng.random((T + 1, N, 1)) < 0.98).astype(np.float32)
    masks[0] = 1.0
    if w.head == "categorical":
        actions = rng.integers(0, w.num_actions, size=(T, N, 1), dtype=np.int64)
    else:
        actions = rng.standard_normal((T, N, w.num_actions), dtype=np.float32)
    hidden = np.zeros((T + 1, N, w.storage_hidden), np.float32)
    if w.base == "gru":
        hidden[0] = 0.1 * rng.standard_normal((N, w.storage_hidden), dtype=np.float32)
    next_value = rng.standard_normal((N, 1), dtype=np.float32)
    return dict(observations=obs, rewards=rewards, masks=masks, actions=actions, hidden_states=hidden,
                next_value=next_value,
                value_predictions=np.zeros((T + 1, N, 1), np.float32), returns=np.zeros((T + 1, N, 1), np.float32),
                action_log_probs=np.zeros((T, N, 1), np.float32))


def flops_per_row(w: Workload) -> int:
    """Algorithmic fwd+bwd FLOPs per transition per epoch = 6 x MACs/row."""
    H, D, A = w.hidden, w.obs_dim, w.num_actions
    if w.base == "mlp":
        macs = 2 * (D * H + H * H) + H + A * H
    elif w.base == "gru":
        macs = 3 * H * (D + H) + 2 * (H * H + H * H) + H + A * H
    else:
        macs = 7640576 + 512 + A * 512
    return 6 * macs


def _orthogonal(rng, rows: int, cols: int, gain: float) -> np.ndarray:
    a = rng.standard_normal((max(rows, cols), min(rows, cols)))
    q, r = np.linalg.qr(a)
    q = q * np.sign(np.diag(r))
    if rows < cols:
        q = q.T
    return (gain * q[:rows, :cols]).astype(np.float32)


def init_params(w: Workload, seed: int) -> np.ndarray:
    """Random-init flat parameters of the workload's architecture, in the reference's registration order, with
    the reference's init scheme (orthogonal weights with the gains of mlp_base.cpp:40-42 /
    output_layers.cpp:37,51, zero biases, scale_log = 0).  Used by bench.py (no checkpoints offline)."""
    from .api import cnn_param_shapes, mlp_param_shapes
    rng = np.random.default_rng(seed)
    head = "Discrete" if w.head == "categorical" else "Box"
    out = []
    if w.base == "cnn":
        shapes = cnn_param_shapes(w.obs_shape[0], w.num_actions, head, w.hidden)
    else:
        shapes = mlp_param_shapes(w.obs_dim, w.num_actions, head, w.hidden, w.base == "gru")
    for name, shape in shapes:
        if len(shape) >= 2:
            gain = np.sqrt(2.0)
            if name.startswith("output.linear_loc") or "gru" in name or ("critic_linear" in name and w.base == "cnn"):
                gain = 1.0
            elif name.startswith("output.linear"):
                gain = 0.01
            out.append(_orthogonal(rng, shape[0], int(np.prod(shape[1:])), gain).ravel())
        else:
            out.append(np.zeros(shape, np.float32).ravel())
    return np.concatenate(out)
