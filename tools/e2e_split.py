#!/usr/bin/env python
"""Where the end-to-end step of a workload spends its time: host-clock medians (synchronised on both sides) of
(a) the rollout's host-to-device copies alone, (b) copies + compute_returns, (c) the update alone through the public API
(host permutations from the reference's random stream), (d) the whole e2e step, (e) the device-resident step."""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from bench import GpuSide
from pytorch_cpp_rl_b200.workloads import CONFIGS

cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
w = CONFIGS[cfg]
side = GpuSide(w, 0)


def copies():
    for n in side.fields:
        getattr(side.st, n).copy_(side.pinned[n], non_blocking=True)
    side.next_value.copy_(side.pinned_nv, non_blocking=True)


def copies_returns():
    copies()
    side.st.compute_returns(side.next_value, w.use_gae, w.gamma, w.tau)


def update_only():
    return side.algo.update(side.st)


def timed(fn, reps=7):
    out = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        out.append((time.perf_counter() - t0) * 1e3)
    return float(np.median(out[2:])), float(np.min(out[2:]))


rollout_bytes = sum(t.numel() * t.element_size() for t in side.pinned.values())
print(f"{cfg}: rollout {rollout_bytes / 1e6:.1f} MB in {len(side.fields)} copies")
for name, fn in [("copies", copies), ("copies + compute_returns", copies_returns), ("update (public API, host permutations)", update_only),
                 ("e2e step", side.step_e2e), ("device-resident step", side.step_device)]:
    med, mn = timed(fn)
    extra = f"  ({rollout_bytes / med / 1e6:.1f} GB/s)" if name == "copies" else ""
    print(f"  {name:42s} median {med:7.3f} ms  min {mn:7.3f} ms{extra}")
