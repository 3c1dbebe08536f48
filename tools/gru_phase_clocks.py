#!/usr/bin/env python
"""Per-phase cycle breakdown of one minibatch of the fused recurrent update kernel (clock64 markers)."""
import ctypes as C
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from bench import GpuSide
from pytorch_cpp_rl_b200 import cabi
from pytorch_cpp_rl_b200.workloads import CONFIGS

NAMES = ["weights+sequence stage", "GRU forward", "towers fwd+loss+bwd", "KL/stats + tower Adam", "BPTT", "GRU Adam"]
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 5
w = CONFIGS["cfg4"]
if len(sys.argv) > 2:
    w = w.scaled(N=int(sys.argv[2]), num_mini_batch=int(sys.argv[2]))
side = GpuSide(w, 0, 0, 1)
lib = cabi.load()
buf = torch.zeros(148 * 32, dtype=torch.int64, device="cuda")
side.step_device()
torch.cuda.synchronize()
lib.cpprl_b200_debug_phase_clocks(C.c_void_p(buf.data_ptr()), mb)
side.step_device()
torch.cuda.synchronize()
lib.cpprl_b200_debug_phase_clocks(None, 0)
c = buf.cpu().numpy().reshape(148, 32)[0]
d = np.diff(c[:7])
print(f"cfg4 (T={w.T}, N={w.N}) minibatch {mb}: cycles per phase")
for n, v in zip(NAMES, d):
    print(f"  {n:26s} {v:9d}  ({v / 1.965e3:7.1f} us)")
print(f"  {'total':26s} {c[6] - c[0]:9d}  ({(c[6] - c[0]) / 1.965e3:7.1f} us)")
