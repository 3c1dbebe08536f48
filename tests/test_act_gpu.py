"""GPU checks of cpprl_b200_act (PolicyImpl::act, policy.cpp:52-81): forward + sample + log-prob.

Parity with the reference is distributional (it samples from the CPU mt19937 stream): value and log-prob must agree
with evaluate_actions — ours and the compiled reference's — on the SAMPLED actions, the empirical action
frequencies / moments must match the distribution the reference computes, and (seed, offset) must reproduce."""
import numpy as np
import pytest
import torch

from pytorch_cpp_rl_b200.workloads import CONFIGS, synth_rollout
from tests import helpers as H

pytestmark = pytest.mark.gpu


def _setup(ref, w, seed):
    r = synth_rollout(w, seed)
    sess = H.make_ref_session(ref, w, init_seed=seed)
    rng = np.random.default_rng(seed)
    params = sess.get_params() + (rng.standard_normal(sess.n_params) * 0.05).astype(np.float32)
    sess.set_params(params)
    policy, st = H.make_gpu_side(w, r, params, sess.get_normalizer() if w.normalize_obs else None)
    return r, sess, policy, st


def test_act_categorical_mlp(ref):
    w = CONFIGS["cfg1"].scaled(T=1, N=8192)
    r, sess, policy, st = _setup(ref, w, 3)
    obs = st.observations[0].reshape(w.N, -1).contiguous()
    v, a, lp, _ = policy.act(obs, None, None, seed=11)
    assert a.dtype == torch.int64 and a.shape == (w.N, 1) and int(a.min()) >= 0 and int(a.max()) < w.num_actions
    # same (seed, offset) -> same actions; another seed -> other actions
    assert torch.equal(policy.act(obs, None, None, seed=11)[1], a)
    assert not torch.equal(policy.act(obs, None, None, seed=12)[1], a)
    assert not torch.equal(policy.act(obs, None, None, seed=11, offset=1)[1], a)
    # value / log-prob on the sampled actions: our evaluate_actions, then the reference's
    ev = policy.evaluate_actions(obs, None, None, a.reshape(-1))
    assert torch.allclose(ev[0], v, rtol=0, atol=1e-6) and torch.allclose(ev[1], lp, rtol=0, atol=1e-6)
    obs_h = obs.cpu().numpy()
    rv, rlp, _ = sess.evaluate_actions(obs_h, np.zeros((w.N, 1), np.float32), np.ones((w.N, 1), np.float32), a.cpu().numpy())
    np.testing.assert_allclose(v.cpu().numpy(), rv, rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(lp.cpu().numpy(), rlp, rtol=1e-4, atol=2e-5)
    # frequencies: the mean of the reference's probabilities over the rows vs the empirical action histogram
    probs = np.stack([np.exp(sess.evaluate_actions(obs_h, np.zeros((w.N, 1), np.float32), np.ones((w.N, 1), np.float32),
                                                   np.full((w.N, 1), k))[1][:, 0]) for k in range(w.num_actions)], 1)
    np.testing.assert_allclose(probs.sum(1), 1.0, atol=1e-4)
    hist = np.bincount(a.cpu().numpy().ravel(), minlength=w.num_actions) / w.N
    sigma = np.sqrt((probs * (1 - probs)).sum(0)) / w.N
    assert np.all(np.abs(hist - probs.mean(0)) <= 5 * sigma + 1e-4), (hist, probs.mean(0), sigma)


def test_act_normal_mlp(ref):
    w = CONFIGS["cfg3"].scaled(T=1, N=512)
    r, sess, policy, st = _setup(ref, w, 4)
    obs = st.observations[0].reshape(w.N, -1).contiguous()
    v, a, lp, _ = policy.act(obs, None, None, seed=5)
    assert a.dtype == torch.float32 and a.shape == (w.N, w.num_actions)
    ev = policy.evaluate_actions(obs, None, None, a)
    assert torch.allclose(ev[0], v, atol=1e-6) and torch.allclose(ev[1], lp, rtol=1e-5, atol=1e-5)
    rv, rlp, _ = sess.evaluate_actions(obs.cpu().numpy(), np.zeros((w.N, 1), np.float32), np.ones((w.N, 1), np.float32),
                                       a.cpu().numpy())
    np.testing.assert_allclose(lp.cpu().numpy(), rlp, rtol=1e-4, atol=5e-5)
    # moments over 400 draws per row: std = exp(scale_log), mean independent of the seed
    draws = torch.stack([policy.act(obs, None, None, seed=100 + k)[1] for k in range(400)])
    named = dict(policy.named_parameters())
    sigma = torch.exp(next(p for n, p in named.items() if "logstd" in n or "scale" in n).reshape(-1))
    assert torch.allclose(draws.std(0).mean(0), sigma, rtol=0.03)
    z = (draws - draws.mean(0)) / sigma
    assert abs(float(z.mean())) < 0.01 and abs(float((z ** 3).mean())) < 0.05 and abs(float((z ** 4).mean()) - 3.0) < 0.15


def test_act_recurrent_and_cnn(ref):
    # GRU-fronted MlpBase, step mode (n_hxs == B): the new hidden state is the one evaluate_actions returns
    w = CONFIGS["cfg4"].scaled(T=1, N=64, num_mini_batch=64)
    r, sess, policy, st = _setup(ref, w, 6)
    obs = st.observations[0].reshape(w.N, -1).contiguous()
    hxs = torch.randn(w.N, 64, device="cuda") * 0.3
    masks = (torch.rand(w.N, 1, device="cuda") < 0.9).float()
    v, a, lp, h1 = policy.act(obs, hxs, masks, seed=1)
    ev = policy.evaluate_actions(obs, hxs, masks, a.reshape(-1))
    assert torch.allclose(ev[0], v, atol=1e-6) and torch.allclose(ev[1], lp, atol=1e-6) and torch.equal(ev[3], h1)
    # CnnBase
    w = CONFIGS["cfg5"].scaled(T=1, N=96)
    r, sess, policy, st = _setup(ref, w, 7)
    obs = st.observations[0].reshape(w.N, -1).contiguous()
    v, a, lp, _ = policy.act(obs, None, None, seed=2)
    ev = policy.evaluate_actions(obs, None, None, a.reshape(-1))
    assert torch.allclose(ev[0], v, atol=1e-6) and torch.allclose(ev[1], lp, atol=1e-6)
    rv, rlp, _ = sess.evaluate_actions(obs.cpu().numpy().reshape(w.N, *w.obs_shape), np.zeros((w.N, 1), np.float32),
                                       np.ones((w.N, 1), np.float32), a.cpu().numpy())
    np.testing.assert_allclose(lp.cpu().numpy(), rlp, rtol=1e-4, atol=2e-5)
    assert len(torch.unique(a)) > 1
