#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by running the COMPILED REFERENCE (oracle/_ref, built from
/root/reference by oracle/build_ref.sh) on small seeded rollouts.  Run in the build container (the reference
sources do not exist on the GPU box):

    bash oracle/build_ref.sh && python tests/golden/make_golden.py

Each .npz holds every input (rollout buffers, initial parameters, normaliser state, permutations, hyper-parameters)
and every output (returns, per-update statistics, post-update parameters and normaliser state) of
`compute_returns` + `update`, repeated for `n_updates` consecutive updates, so both the numpy restatement
(oracle/cpprl_oracle.py, CPU tests) and the CUDA path (GPU tests) can be checked without the reference present.
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import ref_binding as ref  # noqa: E402
from pytorch_cpp_rl_b200.workloads import CONFIGS, init_params, synth_rollout  # noqa: E402
from tests import helpers as H  # noqa: E402

HERE = Path(__file__).resolve().parent

CASES = {
    # name: (workload, seed, n_updates)
    "ppo_mlp_categorical_cfg1": (CONFIGS["cfg1"], 0, 2),
    "ppo_mlp_normal_small": (CONFIGS["cfg3"].scaled(T=8, N=16, num_epoch=2, num_mini_batch=4), 1, 2),
    # default kl_target with a large learning rate: the KL stop (ppo.cpp:100-104) trips after 1, 2 and 0 steps
    "ppo_mlp_categorical_klstop": (CONFIGS["cfg1"].scaled(T=16, N=8, kl_target=0.01, learning_rate=3e-2), 5, 3),
    "a2c_mlp_categorical_small": (CONFIGS["cfg2"].scaled(T=5, N=64), 2, 2),
    "ppo_gru_categorical_small": (CONFIGS["cfg4"].scaled(T=16, N=8, num_mini_batch=8, num_epoch=2), 3, 2),
    "a2c_gru_categorical_small": (CONFIGS["cfg4"].scaled(T=12, N=6, algo="a2c", num_epoch=1, num_mini_batch=1,
                                                         use_gae=False, learning_rate=7e-4), 4, 2),
    # CnnBase: 866 439 parameters.  To keep the fixture small the initial parameters are regenerated from the seed
    # (workloads.init_params) and the post-update parameters are stored at every 16th index + the whole tail
    "ppo_cnn_categorical_small": (CONFIGS["cfg5"].scaled(T=4, N=4, num_epoch=2, num_mini_batch=2), 6, 1),
}

HYPER = ["T", "N", "num_actions", "hidden", "storage_hidden", "num_epoch", "num_mini_batch", "gamma", "tau", "clip_param",
         "actor_loss_coef", "value_loss_coef", "entropy_coef", "learning_rate", "epsilon", "alpha", "kl_target"]


def make(name, w, seed, n_updates):
    r = synth_rollout(w, seed)
    sess = H.make_ref_session(ref, w, init_seed=seed)
    cnn = w.base == "cnn"
    if cnn:
        sess.set_params(init_params(w, seed))
    H.fill_policy_outputs(sess, w, r)
    H.ref_load_rollout(sess, r)
    out = {k: r[k] for k in ("observations", "hidden_states", "rewards", "value_predictions", "action_log_probs",
                             "actions", "masks", "next_value")}
    if cnn:
        out["params_seed"] = np.array(seed)
        n = sess.n_params
        out["param_idx"] = np.unique(np.concatenate([np.arange(0, n, 16), np.arange(n - 4096, n)]))
        out["observations"] = r["observations"].astype(np.uint8)  # pixels are integers 0..255
    else:
        out["params_init"] = sess.get_params()
    out["meta"] = np.array([w.algo, w.base, w.head, str(int(w.normalize_obs)), str(int(w.use_gae)), str(tuple(w.obs_shape))])
    out["hyper"] = np.array([float(getattr(w, k)) for k in HYPER], np.float64)
    n_perm = w.N if w.base == "gru" else w.rows
    for u in range(n_updates):
        sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
        out[f"returns{u}"] = sess.get_returns()
        perm_seed = 1000 + seed + u
        stats, _ = sess.update(1.0, perm_seed=perm_seed)
        out[f"stats{u}"] = stats
        out[f"params{u}"] = sess.get_params()[out["param_idx"]] if cnn else sess.get_params()
        if w.algo == "ppo":
            out[f"perms{u}"] = H.draw_perms(perm_seed, n_perm, w.num_epoch).numpy()
        if w.normalize_obs:
            _, c, m, v = sess.get_normalizer()
            out[f"norm{u}"] = np.concatenate([c, m, v])
    out["n_updates"] = np.array(n_updates)
    np.savez_compressed(HERE / f"{name}.npz", **out)
    print(f"{name}: {sess.n_params} params, {n_updates} updates, last stats {out[f'stats{n_updates - 1}']}")


def make_kats():
    """Stand-alone known answers of the reference's distributions / normaliser on random inputs."""
    rng = np.random.default_rng(7)
    logits = (rng.standard_normal((33, 6)) * 3).astype(np.float32)
    actions = rng.integers(0, 6, 33)
    c_logp, c_ent, c_probs = ref.categorical(logits, actions)
    loc = rng.standard_normal((21, 5)).astype(np.float32)
    scale = np.exp(rng.standard_normal(5) * 0.3).astype(np.float32)
    value = rng.standard_normal((21, 5)).astype(np.float32)
    n_logp, n_ent = ref.normal(loc, np.broadcast_to(scale, loc.shape).copy(), value)
    D = 7
    x = (rng.standard_normal((3, 50, D)) * 2 + 1).astype(np.float32)
    y = (rng.standard_normal((9, D)) * 4).astype(np.float32)
    c, m, v = np.array([1e-4], np.float32), np.zeros(D, np.float32), np.ones(D, np.float32)
    states, ys = [], []
    for i in range(3):
        c, m, v, yo = ref.obs_normalizer(D, 2.5, c, m, v, x=x[i], y=y)
        states.append(np.concatenate([c, m, v]))
        ys.append(yo)
    np.savez_compressed(HERE / "kats.npz", logits=logits, actions=actions, c_logp=c_logp, c_ent=c_ent, c_probs=c_probs,
                        loc=loc, scale=scale, value=value, n_logp=n_logp, n_ent=n_ent, rms_x=x, rms_y=y,
                        rms_states=np.stack(states), rms_y_out=np.stack(ys), rms_clip=np.array(2.5))
    print("kats: categorical / normal / observation normaliser")


if __name__ == "__main__":
    assert ref.available(), "build oracle/_ref first (bash oracle/build_ref.sh)"
    only = sys.argv[1:]
    for name, (w, seed, n) in CASES.items():
        if not only or name in only:
            make(name, w, seed, n)
    if not only:
        make_kats()
