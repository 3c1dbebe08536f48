import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from pytorch_cpp_rl_b200 import api
from pytorch_cpp_rl_b200.workloads import CONFIGS, synth_rollout
from oracle import ref_binding as ref
from tests import helpers as H
for seed in (0, 1, 2):
    w = CONFIGS["cfg2"]
    r = synth_rollout(w, seed)
    sess = H.make_ref_session(ref, w, init_seed=seed)
    H.fill_policy_outputs(sess, w, r); H.ref_load_rollout(sess, r)
    policy, st = H.make_gpu_side(w, r, sess.get_params(), sess.get_normalizer())
    algo = api.A2C(policy, 1.0, w.value_loss_coef, w.entropy_coef, w.learning_rate, w.epsilon, w.alpha)
    nv = torch.from_numpy(r["next_value"]).cuda()
    for upd in range(3):
        sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
        sess.update(1.0)
        st.compute_returns(nv, w.use_gae, w.gamma, w.tau)
        algo.update(st, 1.0)
        want_p, got_p = sess.get_params(), policy.params.cpu().numpy()
        err = np.abs(got_p - want_p) / np.maximum(np.abs(want_p), w.learning_rate)
        i = int(np.argmax(err))
        print(f"TC={os.environ.get('CPPRL_B200_MLP_TC','1')} seed {seed} upd {upd}: max {err.max():.3e} at {i} (want {want_p[i]:.6e} got {got_p[i]:.6e}) p99.9 {np.quantile(err,0.999):.2e} median {np.median(err):.2e}")
