"""CPU tests pinning the numpy restatement (oracle/cpprl_oracle.py):
 1. the reference's own known-answer vectors (file:line cited per test);
 2. the golden fixtures generated from the compiled reference (tests/golden/make_golden.py);
 3. when oracle/_ref is present, a live cross-check against the compiled reference on another seed.
"""
import numpy as np
import pytest

from oracle import cpprl_oracle as O
from tests import golden_io as G


# ---- 1. the reference's known answers ------------------------------------------------------------------
@pytest.mark.parametrize("use_gae", [True, False])
def test_compute_returns_known_answer(use_gae):
    """src/storage.cpp:341-437: 3 steps x 2 envs, gamma = tau = 0.6, env 1 ends an episode at step 1."""
    T, N = 3, 2
    rewards = np.array([[0, 1], [1, 2], [2, 3]], np.float32).reshape(T, N, 1)
    values = np.zeros((T + 1, N, 1), np.float32)
    values[:T, :, 0] = [[0, 1], [1, 2], [2, 3]]
    masks = np.ones((T + 1, N, 1), np.float32)
    masks[1:, :, 0] = [[1, 1], [1, 0], [1, 1]]
    returns = np.zeros((T + 1, N, 1), np.float32)
    O.compute_returns(rewards, values, masks, returns, np.array([[0.0], [1.0]], np.float32), use_gae, 0.6, 0.6)
    if use_gae:
        want = [[1.032, 2.2], [2.2, 2.0], [2.0, 3.6], [0.0, 0.0]]  # returns[T] stays untouched
        assert np.array_equal(values[T, :, 0], [0.0, 1.0])          # value_predictions[T] <- next_value
    else:
        want = [[1.32, 2.2], [2.2, 2.0], [2.0, 3.6], [0.0, 1.0]]
    np.testing.assert_allclose(returns[..., 0], want, rtol=1e-6)


def test_categorical_known_answer():
    """src/distributions/categorical.cpp:142-195 (probs constructor; entropy and broadcast log_prob)."""
    probs = np.array([[0.5, 0.5, 0.0, 0.0], [0.25, 0.25, 0.25, 0.25]], np.float32)
    logits, p = O.categorical_from_probs(probs)
    np.testing.assert_allclose(O.categorical_entropy(logits, p), [0.6931, 1.3863], rtol=1e-3)
    # log_prob(actions [2,2]) broadcasts the two distributions along the LAST action axis (categorical.cpp:54-61):
    # out[i][j] = logits[j][actions[i][j]]
    actions = np.array([[0, 1], [2, 3]])
    got = np.array([[logits[j, actions[i, j]] for j in range(2)] for i in range(2)])
    np.testing.assert_allclose(got, [[-0.6931, -1.3863], [-15.9424, -1.3863]], rtol=1e-3)


def test_normal_known_answer():
    """src/distributions/normal.cpp:46-104."""
    loc = np.array([[0, 1, 2], [3, 4, 5]], np.float32)
    scale = np.array([[5, 4, 3], [2, 1, 0]], np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        ent = O.normal_entropy(loc, scale)
        lp = O.normal_log_prob(loc, scale, np.array([[0, 1, 2], [0, 1, 2]], np.float32))
    np.testing.assert_allclose(ent[0], 8.3512, rtol=1e-3)
    assert ent[1] == -np.inf
    np.testing.assert_allclose(lp[0], [-2.5284, -2.3052, -2.0176], rtol=1e-3)
    np.testing.assert_allclose(lp[1, :2], [-2.7371, -5.4189], rtol=1e-3)
    assert np.isnan(lp[1, 2])


def test_observation_normalizer_known_answer():
    """src/observation_normalizer.cpp:108-129: three single-row updates, then normalise the last row."""
    rms = O.RunningMeanStd(5)
    rows = np.array([[-10., 0., 5., 3.2, 0.], [-5., 2., 4., 3.7, -3.], [1, 2, 3, 4, 5]], np.float32)
    for row in rows:
        rms.update(row)
    got = O.process_observation(rows[2], rms.mean, rms.variance, 10.0)
    np.testing.assert_allclose(got, [1.26008659, 0.70712887, -1.2240818, 1.10914509, 1.31322402], rtol=1e-5)


def test_observation_normalizer_clips():
    """src/observation_normalizer.cpp:95-106."""
    out = O.process_observation(np.array([-1000., -100, -10, 0, 10, 100, 1000], np.float32), np.zeros(7, np.float32),
                                np.ones(7, np.float32), 1.0)
    assert out.max() <= 1.0 and out.min() >= -1.0


def test_running_mean_std_matches_batch_moments():
    """src/running_mean_std.cpp:47-70: after several updates the running moments ~ the moments of everything seen."""
    rng = np.random.default_rng(0)
    rms = O.RunningMeanStd(4)
    xs = [rng.standard_normal((n, 4)).astype(np.float32) * 3 + 2 for n in (3, 100, 57)]
    for x in xs:
        rms.update(x)
    allx = np.concatenate(xs)
    np.testing.assert_allclose(rms.mean, allx.mean(0), atol=1e-3)
    np.testing.assert_allclose(rms.variance, allx.var(0), rtol=1e-3)


def test_generator_shapes_and_errors():
    """src/generators/feed_forward_generator.cpp:82-121, recurrent_generator.cpp:100-137, storage.cpp:146-218."""
    T, N = 3, 5
    ro = O.Rollout(np.zeros((T + 1, N, 4, 4), np.float32), np.zeros((T + 1, N, 3), np.float32),
                   np.zeros((T, N, 1), np.float32), np.zeros((T + 1, N, 1), np.float32), np.zeros((T + 1, N, 1), np.float32),
                   np.zeros((T, N, 1), np.float32), np.zeros((T, N, 1), np.int64), np.ones((T + 1, N, 1), np.float32))
    adv = np.zeros((T, N, 1), np.float32)
    mbs = O.feed_forward_minibatches(ro, adv, 5, np.random.default_rng(0).permutation(15))
    assert len(mbs) == 5 and mbs[0]["observations"].shape == (3, 4, 4) and mbs[0]["hidden_states"].shape == (3, 3)
    assert all(mb[k].shape == (3, 1) for mb in mbs for k in ("actions", "value_predictions", "returns", "masks",
                                                             "action_log_probs", "advantages"))
    with pytest.raises(RuntimeError, match="greater than or equal to the number of minibatches"):
        O.feed_forward_minibatches(ro, adv, 16, np.arange(15))
    rb = O.recurrent_minibatches(ro, adv, 5, np.arange(5))
    assert len(rb) == 5 and rb[0]["observations"].shape == (3, 4, 4) and rb[0]["hidden_states"].shape == (1, 3)
    with pytest.raises(RuntimeError):
        O.recurrent_minibatches(ro, adv, 6, np.arange(5))


# ---- 2. golden fixtures from the compiled reference -----------------------------------------------------------
def test_kats_fixture():
    d = np.load(G.GOLDEN / "kats.npz")
    logp_all, probs = O.categorical_from_logits(d["logits"])
    np.testing.assert_allclose(O.categorical_log_prob(logp_all, d["actions"]), d["c_logp"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(O.categorical_entropy(logp_all, probs), d["c_ent"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(probs, d["c_probs"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(O.normal_log_prob(d["loc"], d["scale"], d["value"]), d["n_logp"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(O.normal_entropy(d["loc"], np.broadcast_to(d["scale"], d["loc"].shape)), d["n_ent"], rtol=1e-6)
    D = d["rms_x"].shape[-1]
    rms = O.RunningMeanStd(D)
    for i in range(3):
        rms.update(d["rms_x"][i])
        np.testing.assert_allclose(np.concatenate([rms.count, rms.mean, rms.variance]), d["rms_states"][i], rtol=2e-5, atol=1e-6)
        np.testing.assert_allclose(O.process_observation(d["rms_y"], rms.mean, rms.variance, float(d["rms_clip"])),
                                   d["rms_y_out"][i], rtol=1e-5, atol=1e-5)


def run_oracle_case(w, d, dtype=np.float32):
    """Replays a fixture through the numpy restatement; yields (update index, stats, policy, rollout)."""
    r = G.rollout_arrays(d)
    pol = O.Policy(w.obs_dim, w.num_actions, w.head, w.hidden, w.base == "gru", w.normalize_obs, dtype,
                   cnn_shape=w.obs_shape if w.base == "cnn" else None)
    pol.set_params(G.initial_params(w, d))
    cast = lambda a: a.astype(dtype) if a.dtype == np.float32 else a
    ro = O.Rollout(**{k: cast(r[k]) for k in ("observations", "hidden_states", "rewards", "value_predictions", "returns",
                                              "action_log_probs", "actions", "masks")})
    if w.algo == "ppo":
        algo = O.PPO(pol, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef,
                     w.entropy_coef, w.learning_rate, w.epsilon, 0.5, w.kl_target)
    else:
        algo = O.A2C(pol, 1.0, w.value_loss_coef, w.entropy_coef, w.learning_rate, w.epsilon, w.alpha)
    for u in range(int(d["n_updates"])):
        O.compute_returns(ro.rewards, ro.value_predictions, ro.masks, ro.returns, cast(r["next_value"]), w.use_gae,
                          w.gamma, w.tau)
        stats = algo.update(ro, d[f"perms{u}"]) if w.algo == "ppo" else algo.update(ro)
        yield u, np.array([v for _, v in stats], np.float32), pol, ro


@pytest.mark.parametrize("name", G.CASES)
def test_golden_fixture(name):
    w, d = G.load(name)
    for u, stats, pol, ro in run_oracle_case(w, d):
        assert np.array_equal(ro.returns, d[f"returns{u}"]), "returns are bit-exact (same operator order, fp32)"
        want = d[f"stats{u}"]
        assert len(stats) == len(want)
        np.testing.assert_allclose(stats, want, rtol=1e-4, atol=5e-6, equal_nan=True)
        got_p, wp = G.expected_params(d, u, pol.flat)
        err = np.abs(got_p - wp) / np.maximum(np.abs(wp), w.learning_rate)
        # two fp32 implementations of a sign-like optimiser step: 1e-4 on all but a handful of near-zero gradients
        # (CnnBase: a fifth of the 866k weights sit behind nearly dead ReLU units, |gradient| ~ Adam's eps)
        tail = 0.5 if w.base == "cnn" else 5e-3
        assert np.quantile(err, 0.99) <= 1e-4 and err.max() <= tail, (u, err.max(), np.quantile(err, 0.99))
        if w.normalize_obs:
            got = np.concatenate([pol.normalizer.count, pol.normalizer.mean, pol.normalizer.variance])
            np.testing.assert_allclose(got, d[f"norm{u}"], rtol=2e-5, atol=1e-6)


# ---- 3. live cross-check against the compiled reference ----------------------------------------------------------
def test_live_reference_cross_check(ref):
    from pytorch_cpp_rl_b200.workloads import CONFIGS, synth_rollout
    from tests import helpers as H
    w = CONFIGS["cfg3"].scaled(T=16, N=32, num_epoch=2, num_mini_batch=4)
    r = synth_rollout(w, 11)
    sess = H.make_ref_session(ref, w, init_seed=11)
    H.fill_policy_outputs(sess, w, r)
    H.ref_load_rollout(sess, r)
    pol = O.Policy(w.obs_dim, w.num_actions, w.head, w.hidden, False, True)
    pol.set_params(sess.get_params())
    ro = O.Rollout(**{k: r[k].copy() for k in ("observations", "hidden_states", "rewards", "value_predictions", "returns",
                                               "action_log_probs", "actions", "masks")})
    algo = O.PPO(pol, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef, w.entropy_coef,
                 w.learning_rate, w.epsilon, 0.5, w.kl_target)
    sess.compute_returns(r["next_value"], True, w.gamma, w.tau)
    want, _ = sess.update(1.0, perm_seed=77)
    O.compute_returns(ro.rewards, ro.value_predictions, ro.masks, ro.returns, r["next_value"], True, w.gamma, w.tau)
    assert np.array_equal(ro.returns, sess.get_returns())
    got = np.array([v for _, v in algo.update(ro, H.draw_perms(77, w.rows, w.num_epoch).numpy())], np.float32)
    np.testing.assert_allclose(got, want, rtol=1e-4, atol=5e-6)
    wp = sess.get_params()
    err = np.abs(pol.flat - wp) / np.maximum(np.abs(wp), w.learning_rate)
    assert np.quantile(err, 0.99) <= 1e-4


# ---- the write side of the rollout buffer -----------------------------------------------
def test_rollout_insert_known_answer():
    """src/storage.cpp:298-339 ("insert() inserts values"): non-zero inputs land on rows 1 / 0, the mask row 1 is 0."""
    rng = np.random.default_rng(0)
    ro = O.new_rollout(3, 4, (5, 2), True, 1, 10)
    assert ro.actions.dtype == np.int64 and ro.masks.min() == 1.0 and ro.observations.shape == (4, 4, 5, 2)
    step = O.rollout_insert(ro, 0, rng.random((4, 5, 2), np.float32) + 1, rng.random((4, 10), np.float32) + 1,
                            rng.integers(1, 3, (4, 1)), rng.random((4, 1), np.float32) + 1,
                            rng.random((4, 1), np.float32) + 1, rng.random((4, 1), np.float32) + 1, np.zeros((4, 1)))
    assert step == 1
    assert ro.observations[1, 0, 0, 0] != 0 and ro.hidden_states[1, 0, 0] != 0 and ro.actions[0, 0, 0] != 0
    assert ro.action_log_probs[0, 0, 0] != 0 and ro.value_predictions[0, 0, 0] != 0 and ro.rewards[0, 0, 0] != 0
    assert ro.masks[1, 0, 0] == 0 and ro.observations[0].max() == 0 and ro.masks[0].min() == 1


def test_rollout_after_update_known_answer():
    """src/storage.cpp:438-488: three inserts, then after_update moves row T of observations / hidden states / masks
    to row 0 (expected 6, 2, 0 at the probed positions)."""
    ro = O.new_rollout(3, 2, (3,), True, 1, 2)
    z = np.zeros((2, 1), np.float32)
    step = 0
    for obs, hid, msk in (([0, 1, 2, 1, 2, 3], [0, 1, 0, 1], [0, 1]), ([0, 1, 2, 1, 2, 3], [0, 1, 0, 1], [0, 1]),
                          ([5, 6, 7, 7, 8, 9], [1, 2, 3, 4], [0, 0])):
        step = O.rollout_insert(ro, step, np.array(obs, np.float32).reshape(2, 3), np.array(hid, np.float32).reshape(2, 2),
                                z, z, z, z, np.array(msk, np.float32).reshape(2, 1))
    assert step == 0  # wrapped: (step + 1) % num_steps
    O.rollout_after_update(ro)
    assert ro.observations[0, 0, 1] == 6 and ro.hidden_states[0, 0, 1] == 2 and ro.masks[0, 0, 0] == 0


def test_rollout_concatenate_known_answer():
    """src/storage.cpp:504-581: five one-environment storages joined; column i of every field is storage i."""
    rng = np.random.default_rng(1)
    parts = []
    for _ in range(5):
        ro = O.new_rollout(3, 1, (4,), True, 1, 5)
        step = 0
        for v, m in ((1, 1), (2, 0), (3, 1)):
            step = O.rollout_insert(ro, step, rng.random((1, 4), np.float32), rng.random((1, 5), np.float32),
                                    rng.random((1, 1)), rng.random((1, 1), np.float32), np.float32(v), np.float32(v),
                                    np.float32(m))
        parts.append(ro)
    joined = O.rollout_concatenate(parts)
    assert joined.observations.shape == (4, 5, 4) and joined.hidden_states.shape == (4, 5, 5)
    assert joined.rewards.shape == (3, 5, 1) and joined.actions.shape == (3, 5, 1) and joined.masks.shape == (4, 5, 1)
    for i, p in enumerate(parts):
        for name in ("observations", "hidden_states", "rewards", "value_predictions", "returns", "action_log_probs",
                     "actions", "masks"):
            assert np.array_equal(getattr(joined, name)[:, i:i + 1], getattr(p, name)), name
    assert np.array_equal(joined.rewards[:, 0, 0], [1, 2, 3]) and np.array_equal(joined.masks[:, 0, 0], [1, 1, 0, 1])


def test_combine_normalizers_known_answer():
    """src/observation_normalizer.cpp:145-190: three normalisers after 1, 2 and 3 updates; the combination holds the
    unweighted means of their means / variances and a step count of 6."""
    rng = np.random.default_rng(2)
    others = []
    for i in range(3):
        rms = O.RunningMeanStd(3)
        for _ in range(i + 1):
            rms.update(rng.random(3, np.float32))
        others.append((10.0, rms))
    clip, combined = O.combine_normalizers(others)
    np.testing.assert_allclose(combined.mean, np.mean([r.mean for _, r in others], axis=0), rtol=1e-6)
    np.testing.assert_allclose(combined.variance, np.mean([r.variance for _, r in others], axis=0), rtol=1e-6)
    assert int(combined.count[0]) == 6 and clip == np.float32(10.0)
