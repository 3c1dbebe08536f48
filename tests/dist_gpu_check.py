#!/usr/bin/env python
"""Multi-GPU parity check, run under torchrun on N GPUs of one box (not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/dist_gpu_check.py

Every rank holds an environment shard of ONE global rollout; the env-sharded PPO / A2C update (gradient all-reduce
inside the fused kernel over NVLink peer memory) must reproduce the compiled reference's single-process update of the
global rollout with the same global permutations, and all ranks must end with bit-identical parameters."""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from pytorch_cpp_rl_b200 import api  # noqa: E402
from pytorch_cpp_rl_b200.workloads import CONFIGS, synth_rollout  # noqa: E402
from tests import helpers as H  # noqa: E402


def shard(r, lo, hi):
    out = {}
    for k, v in r.items():
        out[k] = np.ascontiguousarray(v[lo:hi] if k == "next_value" else v[:, lo:hi])
    return out


def run_case(w, seed, rank, world, n_updates=2):
    from oracle import ref_binding as ref
    if w.N < world:  # every rank needs at least one environment
        if rank == 0:
            print(f"[dist] {w.name} {w.algo}: skipped, {w.N} environments < world={world}", flush=True)
        return True
    r = synth_rollout(w, seed)
    sess = H.make_ref_session(ref, w, init_seed=seed)          # every rank computes the same reference (CPU)
    H.fill_policy_outputs(sess, w, r)
    H.ref_load_rollout(sess, r)
    lo, hi = api.shard_range(w.N, rank, world)
    wl = w.scaled(N=hi - lo)
    policy, st = H.make_gpu_side(wl, shard(r, lo, hi), sess.get_params(), sess.get_normalizer() if w.normalize_obs else None)
    comm = api.Communicator(policy.params.numel())
    if w.algo == "ppo":
        algo = api.PPO(policy, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef,
                       w.entropy_coef, w.learning_rate, w.epsilon, 0.5, w.kl_target)
        algo.set_communicator(comm, lo, w.N)
    else:
        algo = api.A2C(policy, 1.0, w.value_loss_coef, w.entropy_coef, w.learning_rate, w.epsilon, w.alpha)
        algo.set_communicator(comm, w.N)
    nv = torch.from_numpy(np.ascontiguousarray(r["next_value"][lo:hi])).cuda()
    ok = True
    for upd in range(n_updates):
        sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
        want, _ = sess.update(1.0, perm_seed=1000 + seed + upd)
        st.compute_returns(nv, w.use_gae, w.gamma, w.tau)
        assert np.array_equal(st.returns.cpu().numpy(), sess.get_returns()[:, lo:hi])
        if w.algo == "ppo":
            got = algo.update(st, 1.0, perms=H.draw_perms(1000 + seed + upd, w.rows, w.num_epoch))
        else:
            got = algo.update(st, 1.0)
        got = np.array([d.value for d in got], np.float32)
        np.testing.assert_allclose(got[:4] if w.algo == "ppo" else got, want[:4] if w.algo == "ppo" else want,
                                   rtol=2e-4 if upd == 0 or w.base != "cnn" else 1e-3, atol=5e-6)
        want_p, got_p = sess.get_params(), policy.params.cpu().numpy()
        err = np.abs(got_p - want_p) / np.maximum(np.abs(want_p), w.learning_rate)
        budget = 1e-4 * max(1, w.optimizer_steps) ** 0.5 + 2e-4
        if w.base == "cnn":
            # CnnBase: the bulk within 1e-4, the near-dead-ReLU tail as in tests/test_cnn_gpu.py
            bulk_ok = np.quantile(err, 0.99) <= 1e-4 and np.quantile(err, 0.999) <= 2e-3
            budget = 0.5 if bulk_ok else 0.0
        # all ranks bit-identical
        gathered = [torch.empty_like(policy.params) for _ in range(world)]
        dist.all_gather(gathered, policy.params)
        same = all(torch.equal(gathered[0], g) for g in gathered)
        if w.normalize_obs:
            _, c, m, v = sess.get_normalizer()
            n = policy.observation_normalizer
            np.testing.assert_allclose(n.rms.count.cpu().numpy(), c, rtol=1e-6)
            np.testing.assert_allclose(n.rms.mean.cpu().numpy(), m, rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(n.rms.variance.cpu().numpy(), v, rtol=1e-5, atol=1e-6)
        if rank == 0:
            print(f"[dist] {w.name} {w.algo} world={world} update {upd}: max param err {err.max():.3e} (budget {budget:.1e}) "
                  f"ranks identical={same} stats {got}", flush=True)
        ok = ok and same and err.max() <= budget
    return ok


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    ok = True
    ok &= run_case(CONFIGS["cfg1"].scaled(N=16), 0, rank, world)
    ok &= run_case(CONFIGS["cfg3"].scaled(T=32, N=64, num_epoch=3, num_mini_batch=4), 1, rank, world)
    ok &= run_case(CONFIGS["cfg3"].scaled(T=16, N=1536, num_epoch=2, num_mini_batch=2), 2, rank, world, n_updates=1)
    ok &= run_case(CONFIGS["cfg1"].scaled(N=10, kl_target=0.01, learning_rate=3e-2), 5, rank, world, n_updates=3)
    ok &= run_case(CONFIGS["cfg2"].scaled(N=512), 3, rank, world)
    # shards that differ by one environment (api.shard_range): at world=2 the ranks' own row counts would pick
    # different grids (41 vs 40 tiles); the launch plan must come from the rank-independent mean shard
    ok &= run_case(CONFIGS["cfg2"].scaled(N=513), 6, rank, world)
    ok &= run_case(CONFIGS["cfg2"].scaled(N=4099), 8, rank, world, n_updates=1)
    cnn = CONFIGS["cfg5"].scaled(T=8, N=8, num_epoch=2, num_mini_batch=2)
    ok &= run_case(cnn, 0, rank, world, n_updates=1)
    ok &= run_case(cnn.scaled(T=6, N=6, num_mini_batch=3, kl_target=0.003, learning_rate=3e-3, num_epoch=3), 4, rank, world)
    ok &= run_case(cnn.scaled(T=5, N=16, algo="a2c", num_epoch=1, num_mini_batch=1, use_gae=False, learning_rate=7e-4), 7,
                   rank, world, n_updates=1)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("DIST CHECK", "PASSED" if int(flag) else "FAILED", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) else 1)


if __name__ == "__main__":
    main()
