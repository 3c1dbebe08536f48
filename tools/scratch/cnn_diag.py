import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import numpy as np, torch
from oracle import ref_binding as ref
from pytorch_cpp_rl_b200 import api
from pytorch_cpp_rl_b200.workloads import CONFIGS, synth_rollout
from tests import helpers as H

w = CONFIGS["cfg5"].scaled(T=8, N=8, num_epoch=1, num_mini_batch=1)
r = synth_rollout(w, 0)
sess = H.make_ref_session(ref, w, init_seed=0)
H.fill_policy_outputs(sess, w, r)
H.ref_load_rollout(sess, r)
p0 = sess.get_params()
policy, st = H.make_gpu_side(w, r, p0)
algo = api.PPO(policy, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef, w.entropy_coef,
               w.learning_rate, w.epsilon, 0.5, w.kl_target)
nv = torch.from_numpy(r["next_value"]).cuda()
sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
want, _ = sess.update(1.0, perm_seed=300)
st.compute_returns(nv, w.use_gae, w.gamma, w.tau)
got = algo.update(st, 1.0, perms=H.draw_perms(300, w.rows, w.num_epoch))
print("stats", [d.value for d in got], want)
wp, gp = sess.get_params(), policy.params.cpu().numpy()
lr = w.learning_rate
o = 0
for name, n in sess.param_layout():
    a, b, z = gp[o:o + n], wp[o:o + n], p0[o:o + n]
    err = np.abs(a - b) / np.maximum(np.abs(b), lr)
    step_ref = (b - z) / lr
    step_gpu = (a - z) / lr
    bad = err > 1e-3
    print(f"{name:32s} n={n:7d} max={err.max():.2e} q999={np.quantile(err,0.999):.2e} bad={bad.sum():6d} "
          f"|step_ref|<0.9: {(np.abs(step_ref) < 0.9).sum():6d}  mean|step diff| on bad={np.abs(step_ref-step_gpu)[bad].mean() if bad.any() else 0:.3f}")
    if bad.any() and n <= 40000:
        idx = np.nonzero(bad)[0][:8]
        print("    idx", idx, "ref step", step_ref[idx], "gpu step", step_gpu[idx])
    o += n
