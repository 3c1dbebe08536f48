"""The tensor-core variant of the fused MlpBase update (csrc/mlp_tc.cuh: tcgen05 kind::tf32 as 3xTF32; 56- or 64-row tiles
for minibatches of more than 32 rows per SM, 32- and 16-row tiles for smaller ones) against the compiled reference and against the fp32 CUDA-core variant of the same kernel
(CPPRL_B200_MLP_TC=0), over the shapes the variant has to get right: observation widths around the 8-column k-step,
both heads, minibatches that end in a ragged tile, several tiles per CTA, PPO with several optimiser steps and A2C."""
import os

import numpy as np
import pytest
import torch

from pytorch_cpp_rl_b200 import api, cabi
from pytorch_cpp_rl_b200.workloads import CONFIGS
from pytorch_cpp_rl_b200.workloads import synth_rollout
from tests import helpers as H

pytestmark = pytest.mark.gpu


def _variant(rows, obs, act):
    return cabi.load().cpprl_b200_mlp_update_variant(rows, obs, act)


def test_variant_selection():
    assert _variant(8192, 17, 6) == 56         # cfg3's minibatch
    assert _variant(20480, 8, 4) == 56         # cfg2
    assert _variant(9000, 17, 6) == 64         # fewer waves with 64-row tiles
    assert _variant(256, 8, 4) == 16           # cfg1: a minibatch that does not fill the chip -> 16 tiles of 16 rows
    assert _variant(16 * 148, 8, 4) == 16      # one 16-row tile on every SM
    assert _variant(16 * 148 + 1, 8, 4) == 32  # 32-row tiles up to 32 rows per SM
    assert _variant(32 * 148, 17, 6) == 32
    assert _variant(32 * 148 + 1, 17, 6) == 56
    assert _variant(256, 33, 4) == 0
    assert _variant(8192, 33, 6) == 0          # more than 32 observation columns: CUDA-core variant
    os.environ["CPPRL_B200_MLP_TC"] = "0"
    try:
        assert _variant(8192, 17, 6) == 0
        assert _variant(256, 8, 4) == 0
    finally:
        del os.environ["CPPRL_B200_MLP_TC"]


def _update_once(w, r, params, normalizer, perms, n_updates=1):
    policy, st = H.make_gpu_side(w, r, params, normalizer)
    if w.algo == "ppo":
        algo = api.PPO(policy, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef, w.entropy_coef,
                       w.learning_rate, w.epsilon, 0.5, w.kl_target)
    else:
        algo = api.A2C(policy, 1.0, w.value_loss_coef, w.entropy_coef, w.learning_rate, w.epsilon, w.alpha)
    nv = torch.from_numpy(r["next_value"]).cuda()
    stats = []
    for u in range(n_updates):
        st.compute_returns(nv, w.use_gae, w.gamma, w.tau)
        got = algo.update(st, 1.0, perms=perms[u]) if w.algo == "ppo" else algo.update(st, 1.0)
        stats.append(np.array([d.value for d in got], np.float32))
    return policy.params.cpu().numpy(), stats


CASES = {
    # name: (base config, overrides): minibatch rows > 32 * 148 = 4736 (56- / 64-row tiles) or <= 16 * 148 = 2368 (16-row tiles)
    "normal_obs17_ragged": ("cfg3", dict(T=13, N=800, num_epoch=2, num_mini_batch=2)),        # 5200 rows: 92 full tiles + 48 rows
    "normal_obs3": ("cfg3", dict(T=8, N=640, obs_shape=(3,), num_epoch=1, num_mini_batch=1)),   # one k-step, 5 padded columns
    "normal_obs32_act1": ("cfg3", dict(T=8, N=640, obs_shape=(32,), num_actions=1, num_epoch=1, num_mini_batch=1)),
    "categorical_obs9_act16": ("cfg1", dict(T=8, N=700, obs_shape=(9,), num_actions=16, num_epoch=2, num_mini_batch=1,
                                            storage_hidden=1)),
    "multi_tile": ("cfg3", dict(T=20, N=1024, num_epoch=1, num_mini_batch=1)),                 # 20480 rows: three tiles per CTA
    # 10240-row minibatches = 183 tiles: some CTAs own two tiles, and stage the next minibatch's first tile under the second
    "multi_tile_two_minibatches": ("cfg3", dict(T=20, N=1024, num_epoch=2, num_mini_batch=2)),
    "a2c_obs8": ("cfg2", dict(T=3, N=2048)),
    # 9000-row minibatches: 161 tiles of 56 rows (two waves over 148 SMs) but 141 of 64 -> the 64-row tile
    "wide_tile_normal": ("cfg3", dict(T=9, N=1000, num_epoch=2, num_mini_batch=1)),
    "wide_tile_categorical_ragged": ("cfg1", dict(T=7, N=1300, obs_shape=(8,), num_actions=4, num_epoch=1, num_mini_batch=1,
                                                  storage_hidden=1)),                            # 9100 rows = 142 tiles + 12 rows
    "small_tile_cfg1": ("cfg1", dict()),                                                         # 256-row minibatches: 16 tiles
    "small_tile_normal_ragged": ("cfg3", dict(T=7, N=30, num_epoch=2, num_mini_batch=2)),        # 105 rows: 6 tiles + 9 rows
    "small_tile_act16_every_sm": ("cfg1", dict(T=16, N=148, obs_shape=(9,), num_actions=16, num_epoch=1, num_mini_batch=1,
                                               storage_hidden=1)),                               # 2368 rows = 148 full tiles
    "small_tile_obs32_act1": ("cfg3", dict(T=5, N=9, obs_shape=(32,), num_actions=1, num_epoch=2, num_mini_batch=1)),  # 45 rows
    "small_tile_a2c": ("cfg2", dict(T=5, N=64)),
    # 80 optimiser steps in one launch: beyond the table of bias corrections the kernel fills at launch (64 steps)
    "small_tile_80_minibatches": ("cfg1", dict(T=80, N=8, num_epoch=1, num_mini_batch=80)),
    "mid_tile_normal": ("cfg3", dict(T=16, N=200, num_epoch=2, num_mini_batch=1)),               # 3200 rows = 100 tiles of 32
    "mid_tile_categorical_ragged": ("cfg1", dict(T=13, N=300, num_epoch=1, num_mini_batch=1, storage_hidden=1)),  # 121 tiles + 28 rows
    "no_normaliser": ("cfg3", dict(T=8, N=640, normalize_obs=False, num_epoch=1, num_mini_batch=1)),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_tc_variant_matches_reference_and_cuda_core_variant(ref, case):
    base, kw = CASES[case]
    w = CONFIGS[base].scaled(**kw)
    assert _variant(w.rows // max(1, w.num_mini_batch), w.obs_dim, w.num_actions) > 0, "case does not reach the tensor-core variant"
    seed = 3
    r = synth_rollout(w, seed)
    sess = H.make_ref_session(ref, w, init_seed=seed)
    H.fill_policy_outputs(sess, w, r)
    H.ref_load_rollout(sess, r)
    params0 = sess.get_params()
    norm0 = sess.get_normalizer() if w.normalize_obs else None
    n_updates = 2
    perms = [H.draw_perms(900 + u, w.rows, w.num_epoch) for u in range(n_updates)] if w.algo == "ppo" else [None] * n_updates
    want_stats = []
    for u in range(n_updates):
        sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
        s, _ = sess.update(1.0, perm_seed=900 + u) if w.algo == "ppo" else sess.update(1.0)
        want_stats.append(np.asarray(s, np.float32))
    want = sess.get_params()

    tc_params, tc_stats = _update_once(w, r, params0, norm0, perms, n_updates)
    os.environ["CPPRL_B200_MLP_TC"] = "0"
    try:
        cc_params, cc_stats = _update_once(w, r, params0, norm0, perms, n_updates)
    finally:
        del os.environ["CPPRL_B200_MLP_TC"]

    floor = w.learning_rate
    err_tc = np.abs(tc_params - want) / np.maximum(np.abs(want), floor)
    err_cc = np.abs(cc_params - want) / np.maximum(np.abs(want), floor)
    ab = np.abs(tc_params - cc_params) / np.maximum(np.abs(cc_params), floor)
    steps = n_updates * max(1, w.optimizer_steps)
    budget = 1e-4 * steps ** 0.5 + 1e-4
    print(f"[tc] {case}: vs reference tc {err_tc.max():.2e} / cuda-core {err_cc.max():.2e} (budget {budget:.1e}); tc vs cuda-core {ab.max():.2e}")
    assert err_tc.max() <= budget, (err_tc.max(), int(np.argmax(err_tc)))
    assert ab.max() <= budget
    for u in range(n_updates):
        k = 4 if w.algo == "ppo" else 3
        np.testing.assert_allclose(tc_stats[u][:k], want_stats[u][:k], rtol=1e-4, atol=2e-6)
        np.testing.assert_allclose(tc_stats[u][:k], cc_stats[u][:k], rtol=2e-5, atol=1e-6)


def test_tc_variant_is_deterministic(ref):
    w = CONFIGS["cfg3"].scaled(T=13, N=800, num_epoch=2, num_mini_batch=2)
    r = synth_rollout(w, 1)
    sess = H.make_ref_session(ref, w, init_seed=1)
    H.fill_policy_outputs(sess, w, r)
    params0, norm0 = sess.get_params(), sess.get_normalizer()
    perms = [H.draw_perms(5, w.rows, w.num_epoch)]
    a, sa = _update_once(w, r, params0, norm0, perms)
    b, sb = _update_once(w, r, params0, norm0, perms)
    assert np.array_equal(a, b) and np.array_equal(sa[0], sb[0])


def test_tc_variant_kl_early_stop(ref):
    """The KL early stop (ppo.cpp:100-104) INSIDE the tensor-core kernel: every CTA must take the same decision, hand its
    tensor memory back and write its optimiser state before leaving; later epochs' launches must return at once; the next
    update must start from a clean state.  Large learning rate + the default kl_target trip the stop mid-update."""
    w = CONFIGS["cfg3"].scaled(T=13, N=800, num_epoch=4, num_mini_batch=2, kl_target=0.01, learning_rate=2e-2)
    assert _variant(w.rows // w.num_mini_batch, w.obs_dim, w.num_actions) > 0
    seed = 11
    r = synth_rollout(w, seed)
    sess = H.make_ref_session(ref, w, init_seed=seed)
    H.fill_policy_outputs(sess, w, r)
    H.ref_load_rollout(sess, r)
    policy, st = H.make_gpu_side(w, r, sess.get_params(), sess.get_normalizer())
    algo = api.PPO(policy, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef, w.entropy_coef,
                   w.learning_rate, w.epsilon, 0.5, w.kl_target)
    nv = torch.from_numpy(r["next_value"]).cuda()
    stops = 0
    for upd in range(3):
        sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
        want, _ = sess.update(1.0, perm_seed=70 + upd)
        st.compute_returns(nv, w.use_gae, w.gamma, w.tau)
        got = np.array([d.value for d in algo.update(st, 1.0, perms=H.draw_perms(70 + upd, w.rows, w.num_epoch))], np.float32)
        assert len(got) == len(want), (got, want)
        if len(want) > 5:
            stops += 1
            assert got[5] == want[5], "early-stop index differs from the reference's"
        np.testing.assert_allclose(got[:5], want[:5], rtol=5e-4, atol=5e-6, equal_nan=True)
        want_p, got_p = sess.get_params(), policy.params.cpu().numpy()
        err = np.abs(got_p - want_p) / np.maximum(np.abs(want_p), w.learning_rate)
        print(f"[tc] kl stop update {upd}: stats {got} max param err {err.max():.2e}")
        assert err.max() <= 1e-3
    assert stops >= 1, "the case was meant to trip the KL stop"
