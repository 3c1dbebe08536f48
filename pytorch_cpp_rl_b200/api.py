"""Python host-side mirror of the reference's operator interface for the update hot path.

Same names and argument meaning as the reference's C++ classes (`RolloutStorage::compute_returns`,
`PPO::update`, `A2C::update`, `Policy::evaluate_actions`, `FeedForwardGenerator`, `ObservationNormalizer`,
`RunningMeanStd`) so the parity tests read like the reference's own.  Every method forwards to the C ABI
(include/cpprl_b200.h); torch is used only to own device memory and streams.  The C++ drop-in
(include/cpprl/*.h, libcpprl.so) sits on the same C ABI.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch

from . import cabi
from .cabi import check


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _need_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("cpprl_b200: tensors must live on a CUDA device (there is no CPU fallback)")
        if t is not None and not t.is_contiguous():
            raise RuntimeError("cpprl_b200: tensors must be contiguous")


# ---------------------------------------------------------------------------------------------------
# spaces / policy
# ---------------------------------------------------------------------------------------------------
@dataclass
class ActionSpace:  # include/cpprl/spaces.h:9-13 in the reference
    type: str
    shape: Sequence[int]


def mlp_param_shapes(num_inputs: int, num_actions: int, head: str, hidden: int = 64, recurrent: bool = False):
    """(name, shape) list in the reference's module registration order."""
    shapes = []
    tower_in = num_inputs
    if recurrent:
        shapes += [("base.gru.weight_ih_l0", (3 * hidden, num_inputs)), ("base.gru.weight_hh_l0", (3 * hidden, hidden)),
                   ("base.gru.bias_ih_l0", (3 * hidden,)), ("base.gru.bias_hh_l0", (3 * hidden,))]
        tower_in = hidden
    for tower in ("actor", "critic"):
        shapes += [(f"base.{tower}.0.weight", (hidden, tower_in)), (f"base.{tower}.0.bias", (hidden,)),
                   (f"base.{tower}.2.weight", (hidden, hidden)), (f"base.{tower}.2.bias", (hidden,))]
    shapes += [("base.critic_linear.weight", (1, hidden)), ("base.critic_linear.bias", (1,))]
    if head == "Discrete":
        shapes += [("output.linear.weight", (num_actions, hidden)), ("output.linear.bias", (num_actions,))]
    elif head == "Box":
        shapes += [("output.scale_log", (num_actions,)), ("output.linear_loc.weight", (num_actions, hidden)),
                   ("output.linear_loc.bias", (num_actions,))]
    else:
        raise RuntimeError(f"Action space {head} not supported")
    return shapes


def cnn_param_shapes(channels: int, num_actions: int, head: str, hidden: int = 512):
    """CnnBase (cnn_base.cpp:9-31): three convolutions, the 1568 -> hidden linear layer, critic_linear, the head."""
    shapes = [("base.main.0.weight", (32, channels, 8, 8)), ("base.main.0.bias", (32,)),
              ("base.main.2.weight", (64, 32, 4, 4)), ("base.main.2.bias", (64,)),
              ("base.main.4.weight", (32, 64, 3, 3)), ("base.main.4.bias", (32,)),
              ("base.main.7.weight", (hidden, 32 * 7 * 7)), ("base.main.7.bias", (hidden,)),
              ("base.critic_linear.weight", (1, hidden)), ("base.critic_linear.bias", (1,))]
    if head == "Discrete":
        shapes += [("output.linear.weight", (num_actions, hidden)), ("output.linear.bias", (num_actions,))]
    elif head == "Box":
        shapes += [("output.scale_log", (num_actions,)), ("output.linear_loc.weight", (num_actions, hidden)),
                   ("output.linear_loc.bias", (num_actions,))]
    else:
        raise RuntimeError(f"Action space {head} not supported")
    return shapes


class RunningMeanStd:
    """Device buffers count[1], mean[D], variance[D] (running_mean_std.cpp:8-11) updated by the rms kernel."""

    def __init__(self, size: int, device="cuda"):
        self.count = torch.full((1,), 1e-4, dtype=torch.float32, device=device)
        self.mean = torch.zeros(size, dtype=torch.float32, device=device)
        self.variance = torch.ones(size, dtype=torch.float32, device=device)
        self._ws = None

    def update(self, observation: torch.Tensor) -> None:
        D = self.mean.numel()
        x = observation.reshape(-1, D)
        _need_cuda(x)
        lib = cabi.load()
        need = lib.cpprl_b200_rms_workspace_bytes(x.shape[0], D)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=x.device)
        check(lib.cpprl_b200_rms_update(_ptr(x), x.shape[0], D, _ptr(self.count), _ptr(self.mean),
                                        _ptr(self.variance), _ptr(self._ws), _stream()))


class ObservationNormalizer:
    def __init__(self, size: int, clip: float = 10.0, device="cuda"):
        self.clip = torch.full((1,), clip, dtype=torch.float32, device=device)
        self.rms = RunningMeanStd(size, device)

    def process_observation(self, observation: torch.Tensor) -> torch.Tensor:
        D = self.rms.mean.numel()
        x = observation.reshape(-1, D)
        _need_cuda(x)
        out = torch.empty_like(x)
        check(cabi.load().cpprl_b200_normalize_observations(_ptr(x), x.shape[0], D, _ptr(self.rms.mean),
                                                            _ptr(self.rms.variance), _ptr(self.clip), _ptr(out),
                                                            _stream()))
        return out.view(observation.shape)

    def update(self, observations: torch.Tensor) -> None:
        self.rms.update(observations)

    def get_step_count(self) -> int:  # observation_normalizer.h / running_mean_std.h:25: truncated to int
        return int(self.rms.count.item())

    @classmethod
    def combine(cls, others: Sequence["ObservationNormalizer"], device="cuda") -> "ObservationNormalizer":
        """The combining constructor (observation_normalizer.cpp:19-67): UNWEIGHTED means of the others' clips, means
        and variances, step counts truncated and summed — host bookkeeping over small buffers, as in the reference."""
        n = len(others)
        out = cls(others[0].rms.mean.numel(), 0.0, device)
        clip = torch.zeros(1, dtype=torch.float32)
        mean = torch.zeros(out.rms.mean.numel(), dtype=torch.float32)
        var = torch.zeros_like(mean)
        total = 0
        for o in others:
            clip += o.clip.cpu()
            mean += o.rms.mean.cpu()
            var += o.rms.variance.cpu()
            total += o.get_step_count()
        out.clip.copy_(clip / n)
        out.rms.mean.copy_(mean / n)
        out.rms.variance.copy_(var / n)
        out.rms.count.fill_(total + 1e-8)  # set_count, running_mean_std.h:28
        return out


class Policy:
    """Flat-parameter policy (MlpBase, GRU-fronted MlpBase or CnnBase + Categorical / Normal head).  `params` is
    ONE fp32 device buffer in the reference's registration order; `named_parameters()` returns views into it.
    `obs_shape` = (C, 84, 84) selects CnnBase (policy.cpp / gym_client.cpp:117-131 pick the base by observation rank)."""

    def __init__(self, action_space: ActionSpace, num_inputs: int, normalize_observations: bool = False,
                 hidden: int = 64, recurrent: bool = False, device="cuda", obs_shape: Optional[Sequence[int]] = None):
        if action_space.type not in ("Discrete", "Box"):
            raise RuntimeError("Action space " + action_space.type + " not supported")
        self.action_space = action_space
        self.num_inputs = num_inputs
        self.hidden = hidden
        self.recurrent = recurrent
        self.num_actions = int(action_space.shape[0])
        self.obs_chw = tuple(int(x) for x in obs_shape) if obs_shape is not None and len(obs_shape) == 3 else None
        if self.obs_chw is not None:
            if recurrent:
                raise RuntimeError("cpprl_b200: recurrent CnnBase is not implemented")
            self.shapes = cnn_param_shapes(self.obs_chw[0], self.num_actions, action_space.type, hidden)
        else:
            self.shapes = mlp_param_shapes(num_inputs, self.num_actions, action_space.type, hidden, recurrent)
        n = sum(int(torch.Size(s).numel()) for _, s in self.shapes)
        self.params = torch.zeros(n, dtype=torch.float32, device=device)
        self.observation_normalizer = ObservationNormalizer(num_inputs, device=device) if normalize_observations else None

    def named_parameters(self):
        out, o = [], 0
        for name, shape in self.shapes:
            n = int(torch.Size(shape).numel())
            out.append((name, self.params[o:o + n].view(shape)))
            o += n
        return out

    def is_recurrent(self) -> bool:
        return self.recurrent

    def using_observation_normalizer(self) -> bool:
        return self.observation_normalizer is not None

    def update_observation_normalizer(self, observations: torch.Tensor) -> None:
        assert self.observation_normalizer is not None
        self.observation_normalizer.update(observations)

    def desc(self) -> cabi.Policy:
        p = cabi.Policy()
        p.base_kind = cabi.BASE_GRU_MLP if self.recurrent else cabi.BASE_MLP
        if self.obs_chw is not None:
            p.base_kind = cabi.BASE_CNN
            for i in range(3):
                p.obs_chw[i] = self.obs_chw[i]
        p.head_kind = cabi.HEAD_CATEGORICAL if self.action_space.type == "Discrete" else cabi.HEAD_NORMAL
        p.obs_dim = self.num_inputs
        p.hidden = self.hidden
        p.num_actions = self.num_actions
        p.d_params = self.params.data_ptr()
        p.n_params = self.params.numel()
        if self.observation_normalizer is not None:
            n = self.observation_normalizer
            p.d_norm_clip = n.clip.data_ptr()
            p.d_norm_count = n.rms.count.data_ptr()
            p.d_norm_mean = n.rms.mean.data_ptr()
            p.d_norm_var = n.rms.variance.data_ptr()
        return p

    def act(self, inputs, rnn_hxs, masks, seed: Optional[int] = None, offset: int = 0) -> List[torch.Tensor]:
        """PolicyImpl::act (policy.cpp:52-81): {value [B,1], action [B,1] int64 or [B,A], log-prob [B,1], hxs}.
        The sample comes from a Philox stream keyed by (seed, row); without a seed one is drawn from torch's CPU
        generator (so torch.manual_seed makes a run reproducible, as it does for the reference)."""
        _need_cuda(inputs)
        B = inputs.shape[0]
        obs = inputs.reshape(B, -1).float().contiguous()
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        discrete = self.action_space.type == "Discrete"
        A = self.action_space.shape[0]
        value = torch.empty(B, 1, device=obs.device)
        logp = torch.empty(B, 1, device=obs.device)
        action = torch.empty((B, 1), dtype=torch.int64, device=obs.device) if discrete else torch.empty(B, A, device=obs.device)
        hxs = msk = hxs_out = None
        if self.is_recurrent():
            hxs = rnn_hxs.float().contiguous()
            msk = masks.float().reshape(-1).contiguous()
            hxs_out = torch.empty_like(hxs)
        pd = self.desc()
        check(cabi.load().cpprl_b200_act(C.byref(pd), _ptr(obs), _ptr(hxs), 0 if hxs is None else hxs.shape[0], _ptr(msk), B,
                                         C.c_uint64(seed), C.c_uint64(offset), _ptr(value), _ptr(action), _ptr(logp),
                                         _ptr(hxs_out), _stream()))
        return [value, action, logp, hxs_out if hxs_out is not None else rnn_hxs]

    def get_values(self, inputs, rnn_hxs, masks) -> torch.Tensor:
        """PolicyImpl::get_values (policy.cpp:132-144): value [B,1] of the fused forward, no actions needed."""
        _need_cuda(inputs)
        B = inputs.shape[0]
        obs = inputs.reshape(B, -1).float().contiguous()
        value = torch.empty(B, 1, device=obs.device)
        hxs = msk = None
        if self.is_recurrent():
            _need_cuda(rnn_hxs, masks)
            hxs, msk = rnn_hxs.float().contiguous(), masks.float().reshape(-1).contiguous()
        pd = self.desc()
        check(cabi.load().cpprl_b200_get_values(C.byref(pd), _ptr(obs), _ptr(hxs), 0 if hxs is None else hxs.shape[0],
                                                _ptr(msk), B, _ptr(value), _stream()))
        return value

    def evaluate_actions(self, inputs, rnn_hxs, masks, actions) -> List[torch.Tensor]:
        """{value [B,1], action_log_probs [B,1], entropy [], rnn_hxs} — policy.cpp:83-115 (forward only)."""
        _need_cuda(inputs, actions)
        B = inputs.shape[0]
        value = torch.empty(B, 1, dtype=torch.float32, device=inputs.device)
        logp = torch.empty(B, 1, dtype=torch.float32, device=inputs.device)
        entropy = torch.empty(1, dtype=torch.float32, device=inputs.device)
        d = self.desc()
        n_hxs = 0 if rnn_hxs is None else rnn_hxs.shape[0]
        hxs_out = None
        if self.recurrent:
            _need_cuda(rnn_hxs, masks)
            rnn_hxs, masks = rnn_hxs.contiguous(), masks.contiguous()
            hxs_out = torch.empty_like(rnn_hxs)  # new state in step mode, a copy of rnn_hxs in sequence mode
        check(cabi.load().cpprl_b200_evaluate_actions(C.byref(d), _ptr(inputs), _ptr(rnn_hxs), n_hxs, _ptr(masks),
                                                      _ptr(actions), B, _ptr(value), _ptr(logp), _ptr(entropy),
                                                      _ptr(hxs_out), _stream()))
        return [value, logp, entropy.view(()), hxs_out if self.recurrent else rnn_hxs]


# ---------------------------------------------------------------------------------------------------
# rollout storage
# ---------------------------------------------------------------------------------------------------
class RolloutStorage:
    """Time-major fp32 buffers exactly as storage.cpp:15-49 lays them out, resident in HBM."""

    def __init__(self, num_steps: int, num_processes: int, obs_shape: Sequence[int], action_space: ActionSpace,
                 hidden_state_size: int, device="cuda"):
        f = dict(dtype=torch.float32, device=device)
        T, N = num_steps, num_processes
        self.num_steps, self.num_processes = T, N
        self.action_space = action_space
        self.observations = torch.zeros(T + 1, N, *obs_shape, **f)
        self.hidden_states = torch.zeros(T + 1, N, hidden_state_size, **f)
        self.rewards = torch.zeros(T, N, 1, **f)
        self.value_predictions = torch.zeros(T + 1, N, 1, **f)
        self.returns = torch.zeros(T + 1, N, 1, **f)
        self.action_log_probs = torch.zeros(T, N, 1, **f)
        if action_space.type == "Discrete":
            self.actions = torch.zeros(T, N, 1, dtype=torch.int64, device=device)
        else:
            self.actions = torch.zeros(T, N, int(action_space.shape[0]), **f)
        self.masks = torch.ones(T + 1, N, 1, **f)
        self.step = 0

    def compute_returns(self, next_value: torch.Tensor, use_gae: bool, gamma: float, tau: float) -> None:
        _need_cuda(self.rewards, self.value_predictions, self.masks, next_value, self.returns)
        check(cabi.load().cpprl_b200_compute_returns(_ptr(self.rewards), _ptr(self.value_predictions), _ptr(self.masks),
                                                     _ptr(next_value), _ptr(self.returns), self.num_steps,
                                                     self.num_processes, int(use_gae), gamma, tau, _stream()))

    def storage_desc(self) -> cabi.Storage:
        """Writable description of the eight buffers for the write-side entry points (cpprl_b200_storage_*)."""
        bufs = (self.observations, self.hidden_states, self.rewards, self.value_predictions, self.returns,
                self.action_log_probs, self.actions, self.masks)
        _need_cuda(*bufs)
        s = cabi.Storage()
        s.T, s.N = self.num_steps, self.num_processes
        s.obs_dim = self.observations[0, 0].numel()
        s.hidden_dim = self.hidden_states.shape[-1]
        s.action_cols = self.actions.shape[-1]
        s.action_elem_bytes = self.actions.element_size()
        (s.d_observations, s.d_hidden_states, s.d_rewards, s.d_value_preds, s.d_returns, s.d_action_log_probs,
         s.d_actions, s.d_masks) = (t.data_ptr() for t in bufs)
        return s

    def _row(self, value: torch.Tensor, like: torch.Tensor, what: str) -> torch.Tensor:
        # Tensor::copy_ converts dtype / device and broadcasts (storage.cpp:183-189); the kernel takes dense rows
        # of the storage's own dtype, so anything else is converted here first
        if not value.is_cuda:
            raise RuntimeError(f"cpprl_b200: RolloutStorage.insert: {what} must live on a CUDA device "
                               "(there is no CPU fallback)")
        if value.dtype != like.dtype:
            value = value.to(like.dtype)
        if value.shape != like.shape:
            value = value.expand(like.shape)
        return value.contiguous()

    def insert(self, observation, hidden_state, action, action_log_prob, value_prediction, reward, mask) -> None:
        """RolloutStorage::insert (storage.cpp:176-193): all seven fields written by ONE launch."""
        k = self.step
        rows = (self._row(observation, self.observations[k + 1], "observation"),
                self._row(hidden_state, self.hidden_states[k + 1], "hidden_state"),
                self._row(action, self.actions[k], "action"),
                self._row(action_log_prob, self.action_log_probs[k], "action_log_prob"),
                self._row(value_prediction, self.value_predictions[k], "value_prediction"),
                self._row(reward, self.rewards[k], "reward"),
                self._row(mask, self.masks[k + 1], "mask"))
        st = self.storage_desc()
        check(cabi.load().cpprl_b200_storage_insert(C.byref(st), k, *[_ptr(r) for r in rows], _stream()))
        self.step = (k + 1) % self.num_steps

    def bootstrap_returns(self, policy: "Policy", use_gae: bool, gamma: float, tau: float) -> None:
        """`compute_returns(policy.get_values(obs[T], hxs[T], masks[T]), ...)` (gym_client.cpp:266-275) as one call:
        the value forward of time row T lands where compute_returns stores next_value, then the scan; no temporary."""
        st, pd = self.storage_desc(), policy.desc()
        check(cabi.load().cpprl_b200_bootstrap_returns(C.byref(pd), C.byref(st), int(use_gae), gamma, tau, _stream()))

    def after_update(self) -> None:
        """storage.cpp:106-111: time row 0 <- time row T of observations, hidden states and masks, one launch."""
        st = self.storage_desc()
        check(cabi.load().cpprl_b200_storage_after_update(C.byref(st), _stream()))

    def set_first_observation(self, observation: torch.Tensor) -> None:  # storage.cpp:220-223
        row = self._row(observation, self.observations[0], "observation")
        st = self.storage_desc()
        check(cabi.load().cpprl_b200_storage_set_first_observation(C.byref(st), _ptr(row), _stream()))

    @classmethod
    def concatenate(cls, parts: Sequence["RolloutStorage"], device="cuda") -> "RolloutStorage":
        """The concatenating constructor (storage.cpp:51-104): the workers' storages joined along the environment
        axis, one launch per part covering all eight fields."""
        first = parts[0]
        total = sum(p.num_processes for p in parts)
        out = cls(first.num_steps, total, tuple(first.observations.shape[2:]), first.action_space,
                  first.hidden_states.shape[-1], device=device)
        dst = out.storage_desc()
        offset = 0
        for p in parts:
            src = p.storage_desc()
            check(cabi.load().cpprl_b200_storage_concat(C.byref(dst), C.byref(src), offset, _stream()))
            offset += p.num_processes
        return out

    def desc(self) -> cabi.Rollout:
        r = cabi.Rollout()
        r.T, r.N = self.num_steps, self.num_processes
        r.hidden_dim = self.hidden_states.shape[-1]
        r.action_cols = self.actions.shape[-1]
        r.d_observations = self.observations.data_ptr()
        r.d_hidden_states = self.hidden_states.data_ptr()
        r.d_actions = self.actions.data_ptr()
        r.d_value_preds = self.value_predictions.data_ptr()
        r.d_returns = self.returns.data_ptr()
        r.d_masks = self.masks.data_ptr()
        r.d_action_log_probs = self.action_log_probs.data_ptr()
        return r

    def feed_forward_generator(self, advantages: torch.Tensor, num_mini_batch: int, indices: torch.Tensor):
        """Minibatches as FeedForwardGenerator::next yields them (feed_forward_generator.cpp:42-80) for a
        caller-supplied permutation (`indices`, int64, host or device).  Returns a list of dicts."""
        T, N = self.num_steps, self.num_processes
        batch = T * N
        if batch < num_mini_batch:
            raise RuntimeError(
                f"PPO needs the number of processes ({N}) * the number of steps ({T}) = {batch} to be greater than "
                f"or equal to the number of minibatches ({num_mini_batch})")
        mbs = batch // num_mini_batch
        idx = indices.to(self.observations.device).view(-1, mbs)
        fields = {
            "observations": self.observations[:T].reshape(batch, -1),
            "hidden_states": self.hidden_states[:T].reshape(batch, -1),
            "actions": self.actions.reshape(batch, -1),
            "value_predictions": self.value_predictions[:T].reshape(batch, 1),
            "returns": self.returns[:T].reshape(batch, 1),
            "masks": self.masks[:T].reshape(batch, 1),
            "action_log_probs": self.action_log_probs.reshape(batch, 1),
            "advantages": advantages.reshape(batch, 1),
        }
        out = []
        for k in range(idx.shape[0]):
            mb = {name: gather_rows(src.contiguous(), idx[k].contiguous()) for name, src in fields.items()}
            mb["observations"] = mb["observations"].view(mbs, *self.observations.shape[2:])
            out.append(mb)
        return out


# ---------------------------------------------------------------------------------------------------
# stand-alone stages
# ---------------------------------------------------------------------------------------------------
def gather_rows(src: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """dst[i] = src[idx[i]] for 2-D src (bit-exact copy; replaces Tensor::index on flattened rows)."""
    _need_cuda(src, idx)
    assert idx.dtype == torch.int64 and src.dim() == 2
    dst = torch.empty((idx.numel(), src.shape[1]), dtype=src.dtype, device=src.device)
    row_bytes = src.shape[1] * src.element_size()
    check(cabi.load().cpprl_b200_gather_rows(_ptr(src), src.shape[0], row_bytes, _ptr(idx), idx.numel(), _ptr(dst),
                                             _stream()))
    return dst


def advantages(returns: torch.Tensor, value_preds: torch.Tensor, normalize: bool = True):
    """returns[:-1] - value_preds[:-1], normalised as ppo.cpp:51-57.  Returns (adv [T,N,1], moments[4] f64)."""
    _need_cuda(returns, value_preds)
    T, N = returns.shape[0] - 1, returns.shape[1]
    adv = torch.empty(T, N, 1, dtype=torch.float32, device=returns.device)
    moments = torch.empty(4, dtype=torch.float64, device=returns.device)
    check(cabi.load().cpprl_b200_advantages(_ptr(returns), _ptr(value_preds), T, N, int(normalize), _ptr(adv),
                                            _ptr(moments), _stream()))
    return adv, moments


def adam_step(params, grads, exp_avg, exp_avg_sq, step: int, lr: float, eps: float = 1e-8, beta1=0.9, beta2=0.999):
    _need_cuda(params, grads, exp_avg, exp_avg_sq)
    check(cabi.load().cpprl_b200_adam_step(_ptr(params), _ptr(grads), _ptr(exp_avg), _ptr(exp_avg_sq), params.numel(),
                                           step, lr, beta1, beta2, eps, _stream()))


def rmsprop_step(params, grads, square_avg, lr: float, alpha: float = 0.99, eps: float = 1e-8):
    _need_cuda(params, grads, square_avg)
    check(cabi.load().cpprl_b200_rmsprop_step(_ptr(params), _ptr(grads), _ptr(square_avg), params.numel(), lr, alpha,
                                              eps, _stream()))


# ---------------------------------------------------------------------------------------------------
# multi-GPU: environments sharded over the GPUs of one box (include/cpprl_b200.h, section (e))
# ---------------------------------------------------------------------------------------------------
def shard_range(num_envs: int, rank: int, world: int):
    """Environments [lo, hi) owned by `rank`: contiguous, sizes differ by at most one."""
    base, extra = divmod(num_envs, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def exchange_handles(handle: bytes, group=None):
    """All-gather the 64-byte IPC handles of every rank's exchange area (torch.distributed; any backend)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    mine = torch.tensor(list(handle), dtype=torch.uint8, device=dev)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return b"".join(bytes(t.cpu().tolist()) for t in out)


class Communicator:
    """One per rank (one process per GPU).  Owns the peer-mapped exchange area the fused update kernel reduces
    gradients through; torch.distributed is only used to hand the IPC handles around."""

    def __init__(self, max_params: int, group=None):
        import torch.distributed as dist
        distributed = dist.is_available() and dist.is_initialized()
        self.rank, self.world = (dist.get_rank(group), dist.get_world_size(group)) if distributed else (0, 1)
        lib = cabi.load()
        self._h = C.c_void_p()
        check(lib.cpprl_b200_comm_create(C.byref(self._h), self.rank, self.world, max_params))
        if self.world > 1:
            buf = C.create_string_buffer(64)
            check(lib.cpprl_b200_comm_export(self._h, buf))
            self._handles = exchange_handles(buf.raw, group)
            check(lib.cpprl_b200_comm_connect(self._h, C.c_char_p(self._handles)))
            dist.barrier(group)

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value and cabi is not None:
            cabi.load().cpprl_b200_comm_destroy(self._h)
            self._h = C.c_void_p()


# ---------------------------------------------------------------------------------------------------
# algorithms
# ---------------------------------------------------------------------------------------------------
@dataclass
class UpdateDatum:  # algorithm.h:10-14
    name: str
    value: float


class PPO:
    """cpprl::PPO (ppo.h:24-36) over the fused CUDA update.  `update` needs the per-epoch permutations the
    reference's generator would draw; by default they are drawn here with torch.randperm on the CPU generator,
    one per epoch, exactly as feed_forward_generator.cpp:31-34 does."""

    def __init__(self, policy: Policy, clip_param: float, num_epoch: int, num_mini_batch: int, actor_loss_coef: float,
                 value_loss_coef: float, entropy_coef: float, learning_rate: float, epsilon: float = 1e-8,
                 max_grad_norm: float = 0.5, kl_target: float = 0.01):
        self.policy = policy
        self.original_clip_param = clip_param
        self.num_epoch, self.num_mini_batch = num_epoch, num_mini_batch
        self.actor_loss_coef, self.value_loss_coef, self.entropy_coef = actor_loss_coef, value_loss_coef, entropy_coef
        self.original_learning_rate = learning_rate
        self.epsilon, self.max_grad_norm, self.kl_target = epsilon, max_grad_norm, kl_target
        self.device_permutations = False  # opt-in: draw the per-epoch permutations with the CUDA generator (see update)
        self.host_rng_threads = True      # draw torch's CPU-stream permutations in the library (False: torch.randperm here)
        self._h = C.c_void_p()
        check(cabi.load().cpprl_b200_ppo_create(C.byref(self._h), policy.params.numel()))

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value and cabi is not None:
            cabi.load().cpprl_b200_ppo_destroy(self._h)
            self._h = C.c_void_p()

    def _options(self, decay_level: float) -> cabi.PPOOptions:
        import numpy as np
        o = cabi.PPOOptions()
        # float32 products, as the reference forms them (ppo.cpp:47-48)
        o.clip_param = float(np.float32(self.original_clip_param) * np.float32(decay_level))
        o.learning_rate = float(np.float32(self.original_learning_rate) * np.float32(decay_level))
        o.num_epoch, o.num_mini_batch = self.num_epoch, self.num_mini_batch
        o.actor_loss_coef, o.value_loss_coef, o.entropy_coef = self.actor_loss_coef, self.value_loss_coef, self.entropy_coef
        o.epsilon, o.kl_target = self.epsilon, self.kl_target
        return o

    def set_communicator(self, comm: Optional["Communicator"], env_offset: int = 0, global_envs: int = 0):
        """Env-sharded update: this rank's storage holds envs [env_offset, env_offset + N) of `global_envs`;
        permutations passed to update() are then over the GLOBAL rollout (T * global_envs ids)."""
        self._comm, self._global_envs = comm, global_envs
        check(cabi.load().cpprl_b200_ppo_set_comm(self._h, comm._h if comm else None, env_offset, global_envs))

    def optimizer_state(self):
        m, v, step = C.c_void_p(), C.c_void_p(), C.c_int64()
        check(cabi.load().cpprl_b200_ppo_state(self._h, C.byref(m), C.byref(v), C.byref(step)))
        return m.value, v.value, step.value

    def update(self, rollouts: RolloutStorage, decay_level: float = 1.0, perms: Optional[torch.Tensor] = None,
               return_stats: bool = False):
        """PPO::update (ppo.cpp:44-178).  With `perms` (host int64 [num_epoch, n]) the whole update is one C-ABI call.
        Without, each epoch's permutation is drawn here with torch.randperm on the CPU generator exactly when the
        reference's generator constructor would draw it (one per epoch, none after a KL early stop), while the
        previous epoch's kernel is still running on the device."""
        lib = cabi.load()
        T, N = rollouts.num_steps, rollouts.num_processes
        pd, rd, od = self.policy.desc(), rollouts.desc(), self._options(decay_level)
        stats = cabi.PPOStats()
        if perms is not None:
            perms = perms.contiguous()
            assert perms.dtype == torch.int64 and not perms.is_cuda
            check(lib.cpprl_b200_ppo_update(self._h, C.byref(pd), C.byref(rd), C.byref(od), C.c_void_p(perms.data_ptr()),
                                            perms.shape[1], C.byref(stats), _stream()))
        elif (getattr(self, "_comm", None) is None and not self.device_permutations and self.host_rng_threads
              and (N if self.policy.is_recurrent() else T * N) < (2 ** 32 - 1) // 20):
            # the reference's own random stream, drawn by the library: torch's CPU generator state goes in, the state
            # after the epochs' torch::randperm calls comes out (bit-identical permutations, tests/test_host_rng_cpu.py);
            # the epochs' Fisher-Yates walks run on background threads while the device runs the previous epoch
            state = torch.get_rng_state()
            check(lib.cpprl_b200_ppo_update_rng(self._h, C.byref(pd), C.byref(rd), C.byref(od), C.c_void_p(state.data_ptr()),
                                                C.byref(stats), _stream()))
            torch.set_rng_state(state)
        else:
            n = N if self.policy.is_recurrent() else T * (getattr(self, "_global_envs", 0) or N)
            can_stop = self.kl_target * 1.5 < 1e3  # a mean log-prob difference cannot reach this: no stop possible
            keep = []
            check(lib.cpprl_b200_ppo_begin(self._h, C.byref(pd), C.byref(rd), C.byref(od), _stream()))
            for epoch in range(self.num_epoch):
                if can_stop and epoch > 0:
                    stopped = C.c_int()
                    check(lib.cpprl_b200_ppo_stopped(self._h, C.byref(stopped)))
                    if stopped.value:
                        break
                if getattr(self, "_comm", None) is not None or self.device_permutations:
                    # env-sharded: the permutation is over the GLOBAL rollout (T * global_envs ids, identical on every
                    # rank).  Drawing it on the host costs every rank tens of ms per epoch and does not shrink with
                    # the number of GPUs, so it comes from the CUDA generator instead: identically seeded ranks
                    # (torch.manual_seed) on identical GPUs draw identical permutations.  (The reference has no
                    # distributed path whose CPU random stream this could be compared with.)  On one GPU the same
                    # path is the opt-in `device_permutations`: statistically equivalent minibatches, but no longer
                    # the reference's CPU random stream (the default keeps torch.randperm on the CPU generator).
                    perm = torch.randperm(n, dtype=torch.int64, device=self.policy.params.device)
                    keep.append(perm)
                    check(lib.cpprl_b200_ppo_epoch_device(self._h, _ptr(perm), n))
                else:
                    perm = torch.randperm(n, dtype=torch.int64)
                    check(lib.cpprl_b200_ppo_epoch(self._h, C.c_void_p(perm.data_ptr()), n))
            check(lib.cpprl_b200_ppo_end(self._h, C.byref(stats)))
            keep.clear()
        if return_stats:
            return stats
        data = [UpdateDatum("Value loss", stats.value_loss), UpdateDatum("Action loss", stats.action_loss),
                UpdateDatum("Clip fraction", stats.clip_fraction), UpdateDatum("Entropy", stats.entropy),
                UpdateDatum("KL divergence", stats.kl_divergence)]
        if stats.kl_early_stop_update > -1:
            data.append(UpdateDatum("KL divergence early stop update", stats.kl_early_stop_update))
        return data

    # staged interface (device-resident permutations, no synchronisation until fetch) used by bench.py
    def begin(self, rollouts: RolloutStorage, decay_level: float = 1.0):
        self._keep = (self.policy.desc(), rollouts.desc(), self._options(decay_level))
        pd, rd, od = self._keep
        check(cabi.load().cpprl_b200_ppo_begin(self._h, C.byref(pd), C.byref(rd), C.byref(od), _stream()))

    def epoch_device(self, d_perm: torch.Tensor):
        check(cabi.load().cpprl_b200_ppo_epoch_device(self._h, _ptr(d_perm), d_perm.numel()))

    def end_async(self):
        check(cabi.load().cpprl_b200_ppo_end_async(self._h))

    def fetch_stats(self) -> cabi.PPOStats:
        stats = cabi.PPOStats()
        check(cabi.load().cpprl_b200_ppo_fetch_stats(self._h, C.byref(stats)))
        return stats


class A2C:
    """cpprl::A2C (a2c.h:23-32) over the fused CUDA update."""

    def __init__(self, policy: Policy, actor_loss_coef: float, value_loss_coef: float, entropy_coef: float,
                 learning_rate: float, epsilon: float = 1e-8, alpha: float = 0.99, max_grad_norm: float = 0.5):
        self.policy = policy
        self.actor_loss_coef, self.value_loss_coef, self.entropy_coef = actor_loss_coef, value_loss_coef, entropy_coef
        self.original_learning_rate = learning_rate
        self.epsilon, self.alpha, self.max_grad_norm = epsilon, alpha, max_grad_norm
        self._h = C.c_void_p()
        check(cabi.load().cpprl_b200_a2c_create(C.byref(self._h), policy.params.numel()))

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value and cabi is not None:
            cabi.load().cpprl_b200_a2c_destroy(self._h)
            self._h = C.c_void_p()

    def set_communicator(self, comm: Optional["Communicator"], global_envs: int = 0):
        self._comm = comm
        check(cabi.load().cpprl_b200_a2c_set_comm(self._h, comm._h if comm else None, global_envs))

    def _options(self, decay_level: float) -> cabi.A2COptions:
        import numpy as np
        o = cabi.A2COptions()
        o.value_loss_coef, o.entropy_coef = self.value_loss_coef, self.entropy_coef
        o.learning_rate = float(np.float32(self.original_learning_rate) * np.float32(decay_level))
        o.epsilon, o.alpha = self.epsilon, self.alpha
        return o

    def update(self, rollouts: RolloutStorage, decay_level: float = 1.0):
        pd, rd, od = self.policy.desc(), rollouts.desc(), self._options(decay_level)
        stats = cabi.A2CStats()
        check(cabi.load().cpprl_b200_a2c_update(self._h, C.byref(pd), C.byref(rd), C.byref(od), C.byref(stats), _stream()))
        return [UpdateDatum("Value loss", stats.value_loss), UpdateDatum("Action loss", stats.action_loss),
                UpdateDatum("Entropy", stats.entropy)]

    def update_async(self, rollouts: RolloutStorage, decay_level: float = 1.0):
        self._keep = (self.policy.desc(), rollouts.desc(), self._options(decay_level))
        pd, rd, od = self._keep
        check(cabi.load().cpprl_b200_a2c_update_async(self._h, C.byref(pd), C.byref(rd), C.byref(od), _stream()))

    def fetch_stats(self) -> cabi.A2CStats:
        stats = cabi.A2CStats()
        check(cabi.load().cpprl_b200_a2c_fetch_stats(self._h, C.byref(stats)))
        return stats
