import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import numpy as np, torch
from oracle import ref_binding as ref, cpprl_oracle as O
from pytorch_cpp_rl_b200 import api
from pytorch_cpp_rl_b200.workloads import CONFIGS, synth_rollout
from tests import helpers as H

seed = int(sys.argv[1]); epochs = int(sys.argv[2]); nmb = int(sys.argv[3])
w = CONFIGS["cfg5"].scaled(T=8, N=8, num_epoch=epochs, num_mini_batch=nmb)
r = synth_rollout(w, seed)
sess = H.make_ref_session(ref, w, init_seed=seed)
H.fill_policy_outputs(sess, w, r)
H.ref_load_rollout(sess, r)
p0 = sess.get_params()
policy, st = H.make_gpu_side(w, r, p0)
algo = api.PPO(policy, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef, w.entropy_coef,
               w.learning_rate, w.epsilon, 0.5, w.kl_target)
pol64 = O.Policy(w.obs_dim, w.num_actions, w.head, w.hidden, False, False, np.float64, cnn_shape=w.obs_shape)
pol64.set_params(p0)
cast = lambda a: a.astype(np.float64) if a.dtype == np.float32 else a
ro64 = O.Rollout(**{k: cast(r[k].copy()) for k in ("observations", "hidden_states", "rewards", "value_predictions", "returns", "action_log_probs", "actions", "masks")})
algo64 = O.PPO(pol64, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef, w.value_loss_coef, w.entropy_coef, w.learning_rate, w.epsilon, 0.5, w.kl_target)
nv = torch.from_numpy(r["next_value"]).cuda()
perms = H.draw_perms(300 + seed, w.rows, w.num_epoch)
sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
want, _ = sess.update(1.0, perm_seed=300 + seed)
st.compute_returns(nv, w.use_gae, w.gamma, w.tau)
got = algo.update(st, 1.0, perms=perms)
O.compute_returns(ro64.rewards, ro64.value_predictions, ro64.masks, ro64.returns, cast(r["next_value"]), w.use_gae, w.gamma, w.tau)
s64 = algo64.update(ro64, perms.numpy())
print("stats gpu", [d.value for d in got]); print("stats ref", want); print("stats f64", [v for _, v in s64])
p64, wp, gp = pol64.flat, sess.get_params(), policy.params.cpu().numpy()
lr = w.learning_rate
o = 0
for name, n in sess.param_layout():
    f = np.maximum(np.abs(p64[o:o+n]), lr)
    er = np.abs(wp[o:o+n] - p64[o:o+n]) / f
    eg = np.abs(gp[o:o+n] - p64[o:o+n]) / f
    print(f"{name:30s} ref q50 {np.quantile(er,.5):.1e} q99 {np.quantile(er,.99):.1e} max {er.max():.1e} | gpu q50 {np.quantile(eg,.5):.1e} q99 {np.quantile(eg,.99):.1e} max {eg.max():.1e}")
    o += n
