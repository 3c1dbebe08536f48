"""ctypes binding of include/cpprl_b200.h (the C-ABI shared library with the sm_100a kernels).

Loading fails loudly: there is no Python or CPU fallback for any entry point.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "lib" / "libcpprl_b200.so"

OK, EINVAL, ECUDA, EUNSUPPORTED, ENCCL, ESTATE = range(6)
BASE_MLP, BASE_GRU_MLP, BASE_CNN = 0, 1, 2
HEAD_CATEGORICAL, HEAD_NORMAL = 0, 1

c_float_p = C.POINTER(C.c_float)
c_i64_p = C.POINTER(C.c_int64)


class Policy(C.Structure):
    _fields_ = [
        ("base_kind", C.c_int32), ("head_kind", C.c_int32),
        ("obs_dim", C.c_int64), ("obs_chw", C.c_int64 * 3),
        ("hidden", C.c_int64), ("num_actions", C.c_int64),
        ("d_params", C.c_void_p), ("n_params", C.c_int64),
        ("d_norm_clip", C.c_void_p), ("d_norm_count", C.c_void_p),
        ("d_norm_mean", C.c_void_p), ("d_norm_var", C.c_void_p),
    ]


class Rollout(C.Structure):
    _fields_ = [
        ("T", C.c_int64), ("N", C.c_int64), ("hidden_dim", C.c_int64), ("action_cols", C.c_int64),
        ("d_observations", C.c_void_p), ("d_hidden_states", C.c_void_p), ("d_actions", C.c_void_p),
        ("d_value_preds", C.c_void_p), ("d_returns", C.c_void_p), ("d_masks", C.c_void_p),
        ("d_action_log_probs", C.c_void_p),
    ]


class Storage(C.Structure):
    _fields_ = [
        ("T", C.c_int64), ("N", C.c_int64), ("obs_dim", C.c_int64), ("hidden_dim", C.c_int64),
        ("action_cols", C.c_int64), ("action_elem_bytes", C.c_int32),
        ("d_observations", C.c_void_p), ("d_hidden_states", C.c_void_p), ("d_rewards", C.c_void_p),
        ("d_value_preds", C.c_void_p), ("d_returns", C.c_void_p), ("d_action_log_probs", C.c_void_p),
        ("d_actions", C.c_void_p), ("d_masks", C.c_void_p),
    ]


class PPOOptions(C.Structure):
    _fields_ = [
        ("clip_param", C.c_float), ("num_epoch", C.c_int32), ("num_mini_batch", C.c_int32),
        ("actor_loss_coef", C.c_float), ("value_loss_coef", C.c_float), ("entropy_coef", C.c_float),
        ("learning_rate", C.c_float), ("epsilon", C.c_float), ("kl_target", C.c_float),
    ]


class PPOStats(C.Structure):
    _fields_ = [
        ("value_loss", C.c_float), ("action_loss", C.c_float), ("clip_fraction", C.c_float),
        ("entropy", C.c_float), ("kl_divergence", C.c_float), ("kl_early_stop_update", C.c_float),
        ("num_updates", C.c_int32),
    ]


class A2COptions(C.Structure):
    _fields_ = [
        ("value_loss_coef", C.c_float), ("entropy_coef", C.c_float), ("learning_rate", C.c_float),
        ("epsilon", C.c_float), ("alpha", C.c_float),
    ]


class A2CStats(C.Structure):
    _fields_ = [("value_loss", C.c_float), ("action_loss", C.c_float), ("entropy", C.c_float)]


# name -> (restype, argtypes); every symbol include/cpprl_b200.h declares
SIGNATURES = {
    "cpprl_b200_abi_version": (C.c_int, []),
    "cpprl_b200_last_error": (C.c_char_p, []),
    "cpprl_b200_launch_count": (C.c_uint64, []),
    "cpprl_b200_device_info": (C.c_int, [C.POINTER(C.c_int)] * 3 + [C.c_char_p, C.c_int]),
    "cpprl_b200_compute_returns": (C.c_int, [C.c_void_p] * 5 + [C.c_int64, C.c_int64, C.c_int, C.c_float, C.c_float, C.c_void_p]),
    "cpprl_b200_advantages": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cpprl_b200_advantages_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cpprl_b200_gather_rows": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "cpprl_b200_rms_workspace_bytes": (C.c_int64, [C.c_int64, C.c_int64]),
    "cpprl_b200_rms_update": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64] + [C.c_void_p] * 5),
    "cpprl_b200_normalize_observations": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64] + [C.c_void_p] * 5),
    "cpprl_b200_adam_step": (C.c_int, [C.c_void_p] * 4 + [C.c_int64, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p]),
    "cpprl_b200_rmsprop_step": (C.c_int, [C.c_void_p] * 3 + [C.c_int64, C.c_float, C.c_float, C.c_float, C.c_void_p]),
    "cpprl_b200_policy_num_params": (C.c_int64, [C.POINTER(Policy)]),
    "cpprl_b200_evaluate_actions": (C.c_int, [C.POINTER(Policy), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64] + [C.c_void_p] * 5),
    "cpprl_b200_act": (C.c_int, [C.POINTER(Policy), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_uint64, C.c_uint64]
                       + [C.c_void_p] * 5),
    "cpprl_b200_ppo_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int64]),
    "cpprl_b200_ppo_destroy": (None, [C.c_void_p]),
    "cpprl_b200_ppo_state": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]),
    "cpprl_b200_ppo_set_step": (C.c_int, [C.c_void_p, C.c_int64]),
    "cpprl_b200_ppo_update": (C.c_int, [C.c_void_p, C.POINTER(Policy), C.POINTER(Rollout), C.POINTER(PPOOptions), C.c_void_p, C.c_int64, C.POINTER(PPOStats), C.c_void_p]),
    "cpprl_b200_ppo_update_rng": (C.c_int, [C.c_void_p, C.POINTER(Policy), C.POINTER(Rollout), C.POINTER(PPOOptions), C.c_void_p, C.POINTER(PPOStats), C.c_void_p]),
    "cpprl_b200_host_randperm": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int]),
    "cpprl_b200_rng_state_bytes": (C.c_int64, []),
    "cpprl_b200_ppo_begin": (C.c_int, [C.c_void_p, C.POINTER(Policy), C.POINTER(Rollout), C.POINTER(PPOOptions), C.c_void_p]),
    "cpprl_b200_ppo_epoch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "cpprl_b200_ppo_epoch_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "cpprl_b200_ppo_stopped": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "cpprl_b200_ppo_end": (C.c_int, [C.c_void_p, C.POINTER(PPOStats)]),
    "cpprl_b200_ppo_end_async": (C.c_int, [C.c_void_p]),
    "cpprl_b200_ppo_fetch_stats": (C.c_int, [C.c_void_p, C.POINTER(PPOStats)]),
    "cpprl_b200_a2c_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int64]),
    "cpprl_b200_a2c_destroy": (None, [C.c_void_p]),
    "cpprl_b200_a2c_state": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "cpprl_b200_a2c_update": (C.c_int, [C.c_void_p, C.POINTER(Policy), C.POINTER(Rollout), C.POINTER(A2COptions), C.POINTER(A2CStats), C.c_void_p]),
    "cpprl_b200_a2c_update_async": (C.c_int, [C.c_void_p, C.POINTER(Policy), C.POINTER(Rollout), C.POINTER(A2COptions), C.c_void_p]),
    "cpprl_b200_a2c_fetch_stats": (C.c_int, [C.c_void_p, C.POINTER(A2CStats)]),
    "cpprl_b200_comm_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int64]),
    "cpprl_b200_comm_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "cpprl_b200_comm_connect": (C.c_int, [C.c_void_p, C.c_void_p]),
    "cpprl_b200_comm_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "cpprl_b200_comm_destroy": (None, [C.c_void_p]),
    "cpprl_b200_ppo_set_comm": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]),
    "cpprl_b200_a2c_set_comm": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "cpprl_b200_rms_update_sharded": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64] + [C.c_void_p] * 5),
    "cpprl_b200_comm_check": (C.c_int, [C.c_void_p, C.c_void_p]),
    "cpprl_b200_storage_insert": (C.c_int, [C.POINTER(Storage), C.c_int64] + [C.c_void_p] * 8),
    "cpprl_b200_storage_after_update": (C.c_int, [C.POINTER(Storage), C.c_void_p]),
    "cpprl_b200_storage_set_first_observation": (C.c_int, [C.POINTER(Storage), C.c_void_p, C.c_void_p]),
    "cpprl_b200_storage_concat": (C.c_int, [C.POINTER(Storage), C.POINTER(Storage), C.c_int64, C.c_void_p]),
    "cpprl_b200_get_values": (C.c_int, [C.POINTER(Policy), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "cpprl_b200_bootstrap_returns": (C.c_int, [C.POINTER(Policy), C.POINTER(Storage), C.c_int, C.c_float, C.c_float, C.c_void_p]),
    "cpprl_b200_kernel_time_ms": (C.c_int, [C.c_char_p, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "cpprl_b200_kernel_time_reset": (None, []),
    "cpprl_b200_kernel_time_enable": (None, [C.c_int]),
    "cpprl_b200_debug_phase_clocks": (None, [C.c_void_p, C.c_int]),
    "cpprl_b200_mlp_update_variant": (C.c_int, [C.c_int64, C.c_int64, C.c_int64]),
    "cpprl_b200_debug_perm_stream": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p]),
    "cpprl_b200_debug_tc_gemm": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
}

_lib = None


class CpprlB200Error(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(message)
        self.code = code


def load() -> C.CDLL:
    """Load lib/libcpprl_b200.so (built by pytorch_cpp_rl_b200.build); raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA extension has not been built "
            "(run `python -m pytorch_cpp_rl_b200.build`); there is no CPU fallback")
    lib = C.CDLL(str(LIB_PATH))
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().cpprl_b200_last_error().decode()
        raise CpprlB200Error(rc, msg or f"cpprl_b200 error {rc}")
