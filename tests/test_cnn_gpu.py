"""GPU parity of the CnnBase path (cnn_base.cpp:9-47) against the compiled reference (oracle/_ref): forward
evaluation, PPO and A2C updates, multi-pass gradient accumulation, Normal head."""
import os

import numpy as np
import pytest
import torch

from pytorch_cpp_rl_b200 import api
from pytorch_cpp_rl_b200.workloads import CONFIGS, synth_rollout
from tests import helpers as H

pytestmark = pytest.mark.gpu


def _cnn(**kw):
    base = dict(T=8, N=8, num_epoch=2, num_mini_batch=2)
    base.update(kw)
    return CONFIGS["cfg5"].scaled(**base)


@pytest.mark.parametrize("head,B,C", [("categorical", 37, 4), ("categorical", 256, 4), ("normal", 19, 3)])
def test_evaluate_actions_cnn_vs_reference(ref, head, B, C):
    w = _cnn(T=1, N=B, head=head, num_actions=6 if head == "categorical" else 3, obs_shape=(C, 84, 84))
    sess = H.make_ref_session(ref, w, init_seed=3)
    rng = np.random.default_rng(5)
    params = sess.get_params() + (rng.standard_normal(sess.n_params) * 0.01).astype(np.float32)
    sess.set_params(params)
    obs = np.floor(rng.random((B,) + tuple(w.obs_shape)) * 256).astype(np.float32)
    if head == "categorical":
        actions = rng.integers(0, w.num_actions, (B, 1))
    else:
        actions = rng.standard_normal((B, w.num_actions)).astype(np.float32)
    v, lp, ent = sess.evaluate_actions(obs, np.zeros((B, 1), np.float32), np.ones((B, 1), np.float32), actions)
    r = synth_rollout(w.scaled(T=1, N=1), 0)
    policy, _ = H.make_gpu_side(w.scaled(T=1, N=1), r, params)
    dact = torch.from_numpy(np.ascontiguousarray(actions)).cuda()
    out = policy.evaluate_actions(torch.from_numpy(obs).cuda().reshape(B, -1), None, None, dact)
    np.testing.assert_allclose(out[0].cpu().numpy(), v, rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(out[1].cpu().numpy(), lp, rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(float(out[2]), ent, rtol=5e-5)


def _run_cnn_ppo(ref, w, seed, n_updates=2, mutate=None):
    r = synth_rollout(w, seed)
    if mutate is not None:
        mutate(r)
    sess = H.make_ref_session(ref, w, init_seed=seed)
    H.fill_policy_outputs(sess, w, r)
    H.ref_load_rollout(sess, r)
    policy, st = H.make_gpu_side(w, r, sess.get_params())
    algo = api.PPO(policy, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef,
                   w.entropy_coef, w.learning_rate, w.epsilon, 0.5, w.kl_target)
    nv = torch.from_numpy(r["next_value"]).cuda()
    for upd in range(n_updates):
        sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
        want, _ = sess.update(1.0, perm_seed=300 + seed + upd)
        st.compute_returns(nv, w.use_gae, w.gamma, w.tau)
        got = algo.update(st, 1.0, perms=H.draw_perms(300 + seed + upd, w.rows, w.num_epoch))
        got = np.array([d.value for d in got], np.float32)
        assert len(got) == len(want), (got, want)
        # the first update starts from identical parameters: 1e-4.  Later updates inherit the parameter tail of the
        # previous one (a ReLU pre-activation within rounding of zero flips one unit's gradient; near-dead units take
        # sign-like Adam steps), so their loss means get 1e-3
        tol = 1e-4 if upd == 0 else 1e-3
        np.testing.assert_allclose(got[:4], want[:4], rtol=tol, atol=5e-6)
        np.testing.assert_allclose(got[4], want[4], rtol=10 * tol, atol=5e-6)
        if len(want) > 5:
            assert got[5] == want[5]
        want_p, got_p = sess.get_params(), policy.params.cpu().numpy()
        err = np.abs(got_p - want_p) / np.maximum(np.abs(want_p), w.learning_rate)
        print(f"[parity] cnn ppo {w.head} update {upd}: max {err.max():.2e} q99 {np.quantile(err, 0.99):.2e} "
              f"q999 {np.quantile(err, 0.999):.2e}")
        # 866k parameters, 170k of them behind dead or nearly dead ReLU units with |gradient| ~ eps = 1e-5, where
        # Adam's step lr * g / (|g| + eps) amplifies a 1e-7 rounding difference by lr / eps per step: after one step
        # the worst parameter is 3e-3 * lr away (test_ppo_update_cnn_single_step), after several the tail reaches
        # tenths of lr on a handful -- the bulk stays at 1e-4
        tail = (0.5 if w.optimizer_steps > 1 else 2e-2)
        assert np.quantile(err, 0.99) <= 1e-4 and np.quantile(err, 0.999) <= 2e-3 and err.max() <= tail, upd
    return policy, sess


def test_ppo_update_cnn(ref):
    _run_cnn_ppo(ref, _cnn(), 0, n_updates=1)  # two updates of this case: test_ppo_update_cnn_float64_tiebreak


def test_ppo_update_cnn_single_step(ref):
    """One optimiser step over all 64 rows: every tensor within 1e-4 (the gradient path itself), tight tail."""
    _run_cnn_ppo(ref, _cnn(num_epoch=1, num_mini_batch=1), 0, n_updates=1)


def test_ppo_update_cnn_multi_pass(ref):
    """Minibatches of 32 rows processed in passes of 12, 12 and 8 rows: gradients and loss sums accumulate."""
    os.environ["CPPRL_B200_CNN_CHUNK"] = "12"
    try:
        _run_cnn_ppo(ref, _cnn(), 1, n_updates=1)
    finally:
        del os.environ["CPPRL_B200_CNN_CHUNK"]


def test_ppo_update_cnn_fractional_pixels(ref):
    """Observations OFF the TF32 grid (frames with fractional values): the per-update scan must flag them and conv1
    must keep all three TF32 products — same tolerances against the reference as the integer-pixel case."""
    def jitter(r):
        rng = np.random.default_rng(9)
        r["observations"] += rng.random(r["observations"].shape, np.float32) * 0.999
        assert (r["observations"].view(np.uint32) & 0x1fff).any()
    _run_cnn_ppo(ref, _cnn(), 0, n_updates=1, mutate=jitter)


def test_cnn_exact_pixel_path_same_bits():
    """Raw 0..255 pixels lie on the TF32 grid, so conv1 drops the A_lo x B_hi product (cnn_gemm_tc.cuh): the skipped
    product is exactly zero, hence the results are those of the full three-product path (up to the run-to-run order of
    the weight gradients' fp32 atomics)."""
    w = _cnn()
    r = synth_rollout(w, 5)
    assert not (r["observations"].view(np.uint32) & 0x1fff).any()
    results = []
    for switch in ("0", "1"):
        os.environ["CPPRL_B200_CNN_NO_EXACT"] = switch
        try:
            n_params = sum(int(np.prod(shape)) for _, shape in api.cnn_param_shapes(w.obs_shape[0], w.num_actions, "Discrete"))
            params = (np.random.default_rng(6).standard_normal(n_params) * 0.03).astype(np.float32)
            policy, st = H.make_gpu_side(w, r, params)
            algo = api.PPO(policy, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef,
                           w.entropy_coef, w.learning_rate, w.epsilon, 0.5, w.kl_target)
            st.compute_returns(torch.from_numpy(r["next_value"]).cuda(), w.use_gae, w.gamma, w.tau)
            stats = algo.update(st, 1.0, perms=H.draw_perms(77, w.rows, w.num_epoch))
            a2c = api.A2C(policy, 1.0, w.value_loss_coef, w.entropy_coef, 7e-4, w.epsilon, w.alpha)
            stats += a2c.update(st, 1.0)
            results.append((policy.params.clone(), [d.value for d in stats]))
        finally:
            del os.environ["CPPRL_B200_CNN_NO_EXACT"]
    # not bitwise: the weight gradients are reduced with fp32 atomics, whose order changes from run to run
    err = (results[0][0] - results[1][0]).abs() / results[1][0].abs().clamp_min(w.learning_rate)
    e = err.cpu().numpy()  # near-dead units take sign-like Adam steps (see _run_cnn_ppo): bulk tight, bounded tail
    assert np.quantile(e, 0.99) <= 1e-5 and e.max() <= 0.5, (np.quantile(e, 0.99), e.max())
    np.testing.assert_allclose(results[0][1], results[1][1], rtol=1e-5, atol=1e-7)
    assert float((results[0][0] - torch.from_numpy(params).cuda()).abs().max()) > 0


def test_ppo_update_cnn_normal_head(ref):
    _run_cnn_ppo(ref, _cnn(head="normal", num_actions=3, obs_shape=(3, 84, 84), T=6, N=6, num_mini_batch=3), 2, n_updates=1)


def test_ppo_update_cnn_kl_early_stop(ref):
    _run_cnn_ppo(ref, _cnn(kl_target=0.003, learning_rate=3e-3, num_epoch=3), 4)


def test_a2c_update_cnn(ref):
    w = _cnn(T=5, N=16, algo="a2c", num_epoch=1, num_mini_batch=1, use_gae=False, learning_rate=7e-4)
    r = synth_rollout(w, 7)
    sess = H.make_ref_session(ref, w, init_seed=7)
    H.fill_policy_outputs(sess, w, r)
    H.ref_load_rollout(sess, r)
    policy, st = H.make_gpu_side(w, r, sess.get_params())
    algo = api.A2C(policy, 1.0, w.value_loss_coef, w.entropy_coef, w.learning_rate, w.epsilon, w.alpha)
    nv = torch.from_numpy(r["next_value"]).cuda()
    for upd in range(2):
        sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
        want, _ = sess.update(1.0)
        st.compute_returns(nv, w.use_gae, w.gamma, w.tau)
        got = np.array([d.value for d in algo.update(st, 1.0)], np.float32)
        np.testing.assert_allclose(got, want, rtol=1e-4, atol=5e-6)
        want_p, got_p = sess.get_params(), policy.params.cpu().numpy()
        err = np.abs(got_p - want_p) / np.maximum(np.abs(want_p), w.learning_rate)
        print(f"[parity] cnn a2c update {upd}: max {err.max():.2e} q99 {np.quantile(err, 0.99):.2e}")
        assert np.quantile(err, 0.99) <= 1e-4 and np.quantile(err, 0.999) <= 2e-3 and err.max() <= 0.5, upd


@pytest.mark.parametrize("seed", [1, 0])
def test_ppo_update_cnn_float64_tiebreak(ref, seed):
    """Judged by the float64 numpy restatement (oracle/cpprl_oracle.py): the CUDA path must be about as close to
    float64 as the compiled reference is, tensor by tensor.

    seed 1 is a clean trajectory: two updates, 8 optimiser steps, every tensor as close to float64 as the reference.
    On seed 0 (one epoch of two minibatches) one conv2 pre-activation lies within fp32 rounding of zero: this path and
    the reference put it on different sides of the ReLU, which changes dL/dz of that one unit and therefore the
    gradients UPSTREAM of it (conv2's filter / bias of that channel, all of conv1) by up to 1e-2 of a step, while
    everything downstream (conv3, the linear layer, the heads) still agrees to 1e-7.  The reference shows the same
    effect against itself between 1 and 16 host threads (tools/scratch/cnn_ref_self.py), and from the third step on
    the perturbed filters move every tensor.  For that seed the upstream tensors get the loose bound and the
    downstream ones the strict one, which also pins where the difference enters."""
    from oracle import cpprl_oracle as O
    w = _cnn() if seed != 0 else _cnn(num_epoch=1)
    r = synth_rollout(w, seed)
    sess = H.make_ref_session(ref, w, init_seed=seed)
    H.fill_policy_outputs(sess, w, r)
    H.ref_load_rollout(sess, r)
    p0 = sess.get_params()
    policy, st = H.make_gpu_side(w, r, p0)
    algo = api.PPO(policy, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef,
                   w.entropy_coef, w.learning_rate, w.epsilon, 0.5, w.kl_target)
    pol64 = O.Policy(w.obs_dim, w.num_actions, w.head, w.hidden, False, False, np.float64, cnn_shape=w.obs_shape)
    pol64.set_params(p0)
    cast = lambda a: a.astype(np.float64) if a.dtype == np.float32 else a
    ro64 = O.Rollout(**{k: cast(r[k].copy()) for k in ("observations", "hidden_states", "rewards", "value_predictions",
                                                       "returns", "action_log_probs", "actions", "masks")})
    algo64 = O.PPO(pol64, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef,
                   w.entropy_coef, w.learning_rate, w.epsilon, 0.5, w.kl_target)
    nv = torch.from_numpy(r["next_value"]).cuda()
    upstream_of_flip = ("base.main.0.", "base.main.2.") if seed == 0 else ()
    for upd in range(2 if seed != 0 else 1):
        perms = H.draw_perms(300 + seed + upd, w.rows, w.num_epoch)
        sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
        want, _ = sess.update(1.0, perm_seed=300 + seed + upd)
        st.compute_returns(nv, w.use_gae, w.gamma, w.tau)
        got = np.array([d.value for d in algo.update(st, 1.0, perms=perms)], np.float32)
        O.compute_returns(ro64.rewards, ro64.value_predictions, ro64.masks, ro64.returns, cast(r["next_value"]), w.use_gae,
                          w.gamma, w.tau)
        s64 = np.array([v for _, v in algo64.update(ro64, perms.numpy())])
        p64, p_ref, p_gpu = pol64.flat, sess.get_params(), policy.params.cpu().numpy()
        o = 0
        for name, n in sess.param_layout():
            floor = np.maximum(np.abs(p64[o:o + n]), w.learning_rate)
            e_ref = np.abs(p_ref[o:o + n] - p64[o:o + n]) / floor
            e_gpu = np.abs(p_gpu[o:o + n] - p64[o:o + n]) / floor
            o += n
            q = 0.99 if n >= 100 else 0.5
            a, b = np.quantile(e_gpu, q), np.quantile(e_ref, q)
            print(f"[tiebreak] seed {seed} update {upd} {name:28s} q{q}: reference {b:.1e}  cuda {a:.1e}")
            if name.startswith(upstream_of_flip) and upstream_of_flip:
                assert a <= 2e-2, (upd, name, a)
            else:
                assert a <= 3.0 * b + 1e-5, (upd, name, a, b)  # 1e-5: a tenth of the 1e-4 parity budget
        for i in range(4):
            d_ref, d_gpu = abs(want[i] - s64[i]), abs(got[i] - s64[i])
            slack = 1e-3 if seed == 0 and upd > 0 else 1e-4
            assert d_gpu <= 3.0 * d_ref + slack * abs(s64[i]) + 5e-6, (upd, i, got[i], want[i], s64[i])
