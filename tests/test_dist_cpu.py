"""world_size-2 `gloo` tests (CPU) of the host-side logic of the env-sharded update:
 * shard_range / handle exchange plumbing of api.Communicator (torch.distributed all_gather of the 64-byte handles);
 * the sharding SEMANTICS the CUDA path implements (include/cpprl_b200.h section (e)): every rank filters the GLOBAL
   permutation to the environments it owns, divides by the GLOBAL minibatch size, all-reduces gradient + statistics
   and takes the same Adam step — checked here with the numpy oracle over gloo against the single-process oracle."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import cpprl_oracle as O
from pytorch_cpp_rl_b200 import api
from pytorch_cpp_rl_b200.workloads import CONFIGS, init_params, synth_rollout


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_range_partitions():
    for n, world in [(8, 2), (10, 4), (1024, 8), (3, 8)]:
        spans = [api.shard_range(n, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [hi - lo for lo, hi in spans]
        assert max(sizes) - min(sizes) <= 1


def _rollout(w, seed):
    r = synth_rollout(w, seed)
    rng = np.random.default_rng(seed + 1)
    r["value_predictions"][: w.T] = rng.standard_normal((w.T, w.N, 1)).astype(np.float32)
    r["action_log_probs"][:] = -np.abs(rng.standard_normal((w.T, w.N, 1))).astype(np.float32) - 1.0
    return r


def _single(w, r, params, perms):
    pol = O.Policy(w.obs_dim, w.num_actions, w.head, w.hidden, False, w.normalize_obs)
    pol.set_params(params)
    ro = O.Rollout(**{k: r[k].copy() for k in ("observations", "hidden_states", "rewards", "value_predictions", "returns",
                                               "action_log_probs", "actions", "masks")})
    O.compute_returns(ro.rewards, ro.value_predictions, ro.masks, ro.returns, r["next_value"], True, w.gamma, w.tau)
    algo = O.PPO(pol, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef, w.entropy_coef,
                 w.learning_rate, w.epsilon, 0.5, w.kl_target)
    stats = algo.update(ro, perms)
    return pol, stats


def _worker(rank, world, port, w, seed, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # plumbing: every rank contributes a distinct 64-byte "handle"; all ranks see them in rank order
        handles = api.exchange_handles(bytes([rank + 1] * 64))
        assert handles == b"".join(bytes([r + 1] * 64) for r in range(world))

        r = _rollout(w, seed)
        params = init_params(w, seed)
        perms = np.stack([np.random.default_rng(100 + e).permutation(w.rows) for e in range(w.num_epoch)])
        lo, hi = api.shard_range(w.N, rank, world)
        n_loc = hi - lo
        pol = O.Policy(w.obs_dim, w.num_actions, w.head, w.hidden, False, w.normalize_obs)
        pol.set_params(params)
        loc = {k: np.ascontiguousarray(v[:, lo:hi]) for k, v in r.items() if k != "next_value"}
        O.compute_returns(loc["rewards"], loc["value_predictions"], loc["masks"], loc["returns"], r["next_value"][lo:hi], True,
                          w.gamma, w.tau)                                       # per-env independent: no communication
        # global advantage moments: all-reduce {sum, sum of squares} in float64 (ppo.cpp:57 over ALL rows)
        adv_raw = (loc["returns"][:-1] - loc["value_predictions"][:-1]).astype(np.float32)
        mom = torch.tensor([adv_raw.astype(np.float64).sum(), (adv_raw.astype(np.float64) ** 2).sum()], dtype=torch.float64)
        dist.all_reduce(mom)
        cnt = float(w.rows)
        mean = mom[0].item() / cnt
        std = np.sqrt(max((mom[1].item() - mom[0].item() * mean) / (cnt - 1.0), 0.0))
        adv = ((adv_raw - np.float32(mean)) / (np.float32(std) + np.float32(1e-5))).reshape(w.T * n_loc, 1)
        opt = O.Adam(pol.n_params, w.learning_rate, w.epsilon)
        mbs = w.rows // w.num_mini_batch
        flat = {k: loc[k][: w.T].reshape((w.T * n_loc,) + loc[k].shape[2:]) for k in ("observations", "actions", "returns",
                                                                                     "action_log_probs")}
        tot = np.zeros(4)
        n_upd = 0
        for e in range(w.num_epoch):
            for k in range(w.num_mini_batch):
                g = perms[e][k * mbs:(k + 1) * mbs]
                t, env = g // w.N, g % w.N
                mine = (env >= lo) & (env < hi)                                  # stable filter of the GLOBAL minibatch
                rows = t[mine] * n_loc + (env[mine] - lo)
                B = float(mbs)                                                    # every mean divides by the GLOBAL count
                value, logp, _, _, tape = pol.evaluate_actions(flat["observations"][rows], None, None, flat["actions"][rows],
                                                               keep=True)
                old, a, ret = flat["action_log_probs"][rows], adv[rows], flat["returns"][rows]
                ratio = np.exp(logp - old)
                lo_c, hi_c = np.float32(1 - w.clip_param), np.float32(1 + w.clip_param)
                s1, s2 = ratio * a, np.clip(ratio, lo_c, hi_c) * a
                passes = ((ratio >= lo_c) & (ratio <= hi_c)) | (s1 < s2)
                d_logp = np.where(passes, -np.float32(w.actor_loss_coef) / np.float32(B) * a * ratio, np.float32(0))
                d_value = np.float32(w.value_loss_coef) * (value - ret) / np.float32(B)
                # entropy mean over the global minibatch: this rank contributes rows/B of it
                grad = pol.backward(tape, d_value.astype(np.float32), d_logp.astype(np.float32),
                                    -w.entropy_coef * len(rows) / B)
                ent_rows = tape["ent_rows"] if w.head == "categorical" else O.normal_entropy(tape["loc"], tape["scale"])
                part = np.array([((ret - value) ** 2).sum(), np.minimum(s1, s2).sum(), ent_rows.sum(), (old - logp).sum()])
                buf = torch.from_numpy(np.concatenate([grad.astype(np.float64), part]))
                dist.all_reduce(buf)                                             # gradient + statistics in ONE message
                pol.flat = opt.step(pol.flat, buf[:-4].numpy().astype(np.float32))
                tot += np.array([0.5 * buf[-4].item() / B, -buf[-3].item() / B, buf[-2].item() / B, 0.0])
                n_upd += 1
        out[rank] = (pol.flat.copy(), tot / n_upd)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("cfg", ["cfg1", "cfg3"])
def test_sharded_semantics_two_ranks_gloo(cfg):
    w = CONFIGS[cfg].scaled(T=8, N=6, num_epoch=2, num_mini_batch=4, normalize_obs=False)
    seed = 4
    r = _rollout(w, seed)
    perms = np.stack([np.random.default_rng(100 + e).permutation(w.rows) for e in range(w.num_epoch)])
    pol, stats = _single(w, r, init_params(w, seed), perms)
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(rank, 2, port, w, seed, out)) for rank in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert np.array_equal(out[0][0], out[1][0]), "ranks must take identical steps"
    err = np.abs(out[0][0] - pol.flat) / np.maximum(np.abs(pol.flat), w.learning_rate)
    assert err.max() < 5e-4, err.max()
    np.testing.assert_allclose(out[0][1][:3], [v for _, v in stats][0:1] + [[v for _, v in stats][1]] + [[v for _, v in stats][3]],
                               rtol=1e-4, atol=1e-6)
