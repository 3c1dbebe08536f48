import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import numpy as np
from oracle import ref_binding as ref
from pytorch_cpp_rl_b200.workloads import CONFIGS, synth_rollout
from tests import helpers as H
w = CONFIGS["cfg5"].scaled(T=8, N=8, num_epoch=2, num_mini_batch=2)
seed = int(sys.argv[1]) if len(sys.argv) > 1 else 0
outs = []
for threads in (1, 16):
    ref.load().ref_set_num_threads(threads)
    r = synth_rollout(w, seed)
    sess = H.make_ref_session(ref, w, init_seed=seed)
    H.fill_policy_outputs(sess, w, r)
    H.ref_load_rollout(sess, r)
    ps = []
    for upd in range(2):
        sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
        st, _ = sess.update(1.0, perm_seed=300 + seed + upd)
        ps.append((st, sess.get_params()))
    outs.append(ps)
for upd in range(2):
    a, b = outs[0][upd][1], outs[1][upd][1]
    err = np.abs(a - b) / np.maximum(np.abs(b), w.learning_rate)
    print(f"ref(1 thread) vs ref(16 threads) update {upd}: max {err.max():.2e} q99 {np.quantile(err,0.99):.2e} q999 {np.quantile(err,0.999):.2e}", outs[0][upd][0], outs[1][upd][0])
