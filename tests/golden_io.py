"""Load a tests/golden/*.npz fixture (written by tests/golden/make_golden.py from the compiled reference)."""
from __future__ import annotations

import ast
from pathlib import Path

import numpy as np

from pytorch_cpp_rl_b200.workloads import Workload

GOLDEN = Path(__file__).resolve().parent / "golden"
HYPER = ["T", "N", "num_actions", "hidden", "storage_hidden", "num_epoch", "num_mini_batch", "gamma", "tau", "clip_param",
         "actor_loss_coef", "value_loss_coef", "entropy_coef", "learning_rate", "epsilon", "alpha", "kl_target"]
INT_HYPER = {"T", "N", "num_actions", "hidden", "storage_hidden", "num_epoch", "num_mini_batch"}
CASES = sorted(p.stem for p in GOLDEN.glob("*.npz") if p.stem != "kats")


def load(name: str):
    d = np.load(GOLDEN / f"{name}.npz")
    algo, base, head, norm, gae, shape = (str(x) for x in d["meta"])
    kw = {k: (int(v) if k in INT_HYPER else float(v)) for k, v in zip(HYPER, d["hyper"])}
    w = Workload(name, "golden fixture", algo, base, head, obs_shape=ast.literal_eval(shape), normalize_obs=bool(int(norm)),
                 use_gae=bool(int(gae)), **kw)
    return w, d


def initial_params(w, d):
    """Initial flat parameters of a fixture: stored, or (CnnBase: 866k values) regenerated from the recorded seed."""
    if "params_init" in d:
        return d["params_init"]
    from pytorch_cpp_rl_b200.workloads import init_params
    return init_params(w, int(d["params_seed"]))


def expected_params(d, u: int, flat):
    """(got, want) for update u: the CnnBase fixture stores the post-update parameters at `param_idx` only."""
    want = d[f"params{u}"]
    return (flat[d["param_idx"]], want) if "param_idx" in d else (flat, want)


def rollout_arrays(d):
    r = {k: d[k].copy() for k in ("observations", "hidden_states", "rewards", "value_predictions", "action_log_probs",
                                  "actions", "masks", "next_value")}
    r["observations"] = r["observations"].astype(np.float32)
    r["returns"] = np.zeros_like(r["value_predictions"])
    return r
