"""ctypes binding of oracle/_ref/libcpprl_ref.so — the reference itself, compiled from /root/reference by
oracle/build_ref.sh, driven through oracle/ref_driver.cpp.

TEST INFRASTRUCTURE ONLY.  Importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; never from the product package.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "_ref" / "libcpprl_ref.so"

ALGO_PPO, ALGO_A2C = 0, 1
BASE_MLP, BASE_GRU_MLP, BASE_CNN = 0, 1, 2
HEAD_CATEGORICAL, HEAD_NORMAL = 0, 1


class RefConfig(C.Structure):  # mirrors oracle/ref_driver.h
    _fields_ = [
        ("algo", C.c_int32), ("base", C.c_int32), ("head", C.c_int32),
        ("T", C.c_int32), ("N", C.c_int32), ("obs_ndim", C.c_int32),
        ("obs_shape", C.c_int64 * 3), ("num_actions", C.c_int64),
        ("hidden", C.c_int32), ("storage_hidden", C.c_int32), ("normalize_obs", C.c_int32),
        ("num_epoch", C.c_int32), ("num_mini_batch", C.c_int32),
        ("clip_param", C.c_float), ("actor_loss_coef", C.c_float), ("value_loss_coef", C.c_float),
        ("entropy_coef", C.c_float), ("learning_rate", C.c_float), ("epsilon", C.c_float),
        ("alpha", C.c_float), ("max_grad_norm", C.c_float), ("kl_target", C.c_float),
    ]


_lib = None


def available() -> bool:
    return LIB_PATH.exists()


def load():
    global _lib
    if _lib is not None:
        return _lib
    import torch  # noqa: F401  (loads libtorch into the process before the oracle library)
    lib = C.CDLL(str(LIB_PATH))
    lib.ref_last_error.restype = C.c_char_p
    lib.ref_torch_version.restype = C.c_char_p
    lib.ref_create.restype = C.c_void_p
    lib.ref_create.argtypes = [C.POINTER(RefConfig), C.c_uint64]
    lib.ref_destroy.argtypes = [C.c_void_p]
    lib.ref_num_params.restype = C.c_int64
    lib.ref_num_params.argtypes = [C.c_void_p]
    lib.ref_manual_seed.argtypes = [C.c_uint64]
    _lib = lib
    return lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class RefError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise RefError(load().ref_last_error().decode())


class RefSession:
    """One reference Policy + RolloutStorage + PPO/A2C, all CPU libtorch, driven with numpy buffers."""

    def __init__(self, *, algo="ppo", base="mlp", head="categorical", T, N, obs_shape, num_actions, hidden=64,
                 storage_hidden=1, normalize_obs=False, num_epoch=4, num_mini_batch=4, clip_param=0.2,
                 actor_loss_coef=1.0, value_loss_coef=0.5, entropy_coef=0.01, learning_rate=3e-4, epsilon=1e-8,
                 alpha=0.99, max_grad_norm=0.5, kl_target=1e9, init_seed=0):
        lib = load()
        c = RefConfig()
        c.algo = {"ppo": ALGO_PPO, "a2c": ALGO_A2C}[algo]
        c.base = {"mlp": BASE_MLP, "gru": BASE_GRU_MLP, "cnn": BASE_CNN}[base]
        c.head = {"categorical": HEAD_CATEGORICAL, "normal": HEAD_NORMAL}[head]
        c.T, c.N = T, N
        c.obs_ndim = len(obs_shape)
        for i, s in enumerate(obs_shape):
            c.obs_shape[i] = s
        c.num_actions = num_actions
        c.hidden, c.storage_hidden, c.normalize_obs = hidden, storage_hidden, int(normalize_obs)
        c.num_epoch, c.num_mini_batch = num_epoch, num_mini_batch
        c.clip_param, c.actor_loss_coef, c.value_loss_coef = clip_param, actor_loss_coef, value_loss_coef
        c.entropy_coef, c.learning_rate, c.epsilon = entropy_coef, learning_rate, epsilon
        c.alpha, c.max_grad_norm, c.kl_target = alpha, max_grad_norm, kl_target
        self.cfg = c
        self.discrete = head == "categorical"
        self.obs_shape = tuple(obs_shape)
        self.lib = lib
        self.h = lib.ref_create(C.byref(c), init_seed)
        if not self.h:
            raise RefError(lib.ref_last_error().decode())
        self.h = C.c_void_p(self.h)
        self.n_params = lib.ref_num_params(self.h)

    def close(self):
        if self.h:
            self.lib.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters / buffers ----
    def param_layout(self):
        buf = C.create_string_buffer(1 << 16)
        n = self.lib.ref_param_layout(self.h, buf, len(buf))
        assert n >= 0
        out = []
        for line in buf.value.decode().strip().split("\n"):
            name, numel = line.rsplit(" ", 1)
            out.append((name, int(numel)))
        return out

    def get_params(self) -> np.ndarray:
        out = np.empty(self.n_params, dtype=np.float32)
        _check(self.lib.ref_get_params(self.h, _p(out)))
        return out

    def set_params(self, p) -> None:
        p = _f32(p)
        assert p.size == self.n_params
        _check(self.lib.ref_set_params(self.h, _p(p)))

    def get_normalizer(self):
        D = int(np.prod(self.obs_shape))
        clip, count = np.zeros(1, np.float32), np.zeros(1, np.float32)
        mean, var = np.zeros(D, np.float32), np.zeros(D, np.float32)
        _check(self.lib.ref_get_normalizer(self.h, _p(clip), _p(count), _p(mean), _p(var)))
        return clip, count, mean, var

    def set_normalizer(self, count, mean, var):
        count, mean, var = _f32(count), _f32(mean), _f32(var)
        _check(self.lib.ref_set_normalizer(self.h, _p(count), _p(mean), _p(var)))

    # ---- rollout ----
    def set_rollout(self, obs, hidden, rewards, value_preds, returns, logp, actions, masks):
        actions = np.ascontiguousarray(actions, dtype=np.int64 if self.discrete else np.float32)
        arrs = [_f32(obs), _f32(hidden), _f32(rewards), _f32(value_preds), _f32(returns), _f32(logp), actions, _f32(masks)]
        _check(self.lib.ref_set_rollout(self.h, *[_p(a) for a in arrs]))

    def get_returns(self):
        out = np.empty((self.cfg.T + 1, self.cfg.N, 1), np.float32)
        _check(self.lib.ref_get_returns(self.h, _p(out)))
        return out

    def get_value_predictions(self):
        out = np.empty((self.cfg.T + 1, self.cfg.N, 1), np.float32)
        _check(self.lib.ref_get_value_predictions(self.h, _p(out)))
        return out

    def compute_returns(self, next_value, use_gae, gamma, tau) -> float:
        nv = _f32(next_value)
        sec = C.c_double()
        _check(self.lib.ref_compute_returns(self.h, _p(nv), int(use_gae), C.c_float(gamma), C.c_float(tau), C.byref(sec)))
        return sec.value

    def update(self, decay_level=1.0, perm_seed=-1):
        stats = np.zeros(8, np.float32)
        n, sec = C.c_int(), C.c_double()
        _check(self.lib.ref_update(self.h, C.c_float(decay_level), C.c_int64(perm_seed), _p(stats), C.byref(n), C.byref(sec)))
        return stats[: n.value].copy(), sec.value

    def after_update(self):
        _check(self.lib.ref_after_update(self.h))

    # ---- policy ----
    def evaluate_actions(self, obs, hxs, masks, actions):
        obs = _f32(obs)
        B = obs.shape[0]
        hxs, masks = _f32(hxs), _f32(masks)
        actions = np.ascontiguousarray(actions, dtype=np.int64 if self.discrete else np.float32)
        value, logp, ent = np.empty((B, 1), np.float32), np.empty((B, 1), np.float32), np.empty(1, np.float32)
        _check(self.lib.ref_evaluate_actions(self.h, _p(obs), _p(hxs), C.c_int64(hxs.shape[0]), _p(masks), _p(actions),
                                             C.c_int64(B), _p(value), _p(logp), _p(ent)))
        return value, logp, float(ent[0])

    def get_values(self, obs, hxs, masks):
        obs, hxs, masks = _f32(obs), _f32(hxs), _f32(masks)
        B = obs.shape[0]
        value = np.empty((B, 1), np.float32)
        _check(self.lib.ref_get_values(self.h, _p(obs), _p(hxs), _p(masks), C.c_int64(B), _p(value)))
        return value

    def generate_minibatches(self, advantages, num_mini_batch, perm_seed):
        c = self.cfg
        rows = c.T * c.N
        D = int(np.prod(self.obs_shape))
        adv = _f32(advantages)
        A = 1 if self.discrete else int(c.num_actions)
        out = {
            "observations": np.zeros((rows, D), np.float32),
            "hidden_states": np.zeros((rows, c.storage_hidden), np.float32),
            "actions": np.zeros((rows, A), np.int64 if self.discrete else np.float32),
            "value_predictions": np.zeros((rows, 1), np.float32), "returns": np.zeros((rows, 1), np.float32),
            "masks": np.zeros((rows, 1), np.float32), "action_log_probs": np.zeros((rows, 1), np.float32),
            "advantages": np.zeros((rows, 1), np.float32),
        }
        n_mb, mbs = C.c_int(), C.c_int()
        _check(self.lib.ref_generate_minibatches(
            self.h, _p(adv), num_mini_batch, C.c_int64(perm_seed), _p(out["observations"]), _p(out["hidden_states"]),
            _p(out["actions"]), _p(out["value_predictions"]), _p(out["returns"]), _p(out["masks"]),
            _p(out["action_log_probs"]), _p(out["advantages"]), C.byref(n_mb), C.byref(mbs)))
        return out, n_mb.value, mbs.value


# ---- stand-alone known-answer helpers ----
def categorical(logits, actions):
    lib = load()
    logits = _f32(logits)
    B, A = logits.shape
    actions = np.ascontiguousarray(actions, dtype=np.int64)
    logp, ent, probs = np.empty(B, np.float32), np.empty(B, np.float32), np.empty((B, A), np.float32)
    _check(lib.ref_categorical(_p(logits), _p(actions), C.c_int64(B), C.c_int64(A), _p(logp), _p(ent), _p(probs)))
    return logp, ent, probs


def normal(loc, scale, value):
    lib = load()
    loc, scale, value = _f32(loc), _f32(scale), _f32(value)
    B, A = loc.shape
    logp, ent = np.empty((B, A), np.float32), np.empty(B, np.float32)
    _check(lib.ref_normal(_p(loc), _p(scale), _p(value), C.c_int64(B), C.c_int64(A), _p(logp), _p(ent)))
    return logp, ent


def obs_normalizer(D, clip, count, mean, var, x=None, y=None):
    """Returns (count, mean, var, y_out) after an optional update with x [M,D] and an optional normalise of y."""
    lib = load()
    count, mean, var = _f32(count).copy(), _f32(mean).copy(), _f32(var).copy()
    x = None if x is None else _f32(x).reshape(-1, D)
    y = None if y is None else _f32(y).reshape(-1, D)
    y_out = None if y is None else np.empty_like(y)
    _check(lib.ref_obs_normalizer(C.c_int64(D), C.c_float(clip), _p(count), _p(mean), _p(var), _p(x),
                                  C.c_int64(0 if x is None else x.shape[0]), _p(y),
                                  C.c_int64(0 if y is None else y.shape[0]), _p(y_out)))
    return count, mean, var, y_out
