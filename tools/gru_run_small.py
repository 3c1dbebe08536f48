#!/usr/bin/env python
"""Small recurrent PPO run (cfg4 with 8 environments) for ncu captures of gru_update_kernel."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from bench import GpuSide
from pytorch_cpp_rl_b200.workloads import CONFIGS

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
w = CONFIGS["cfg4"].scaled(N=n, num_mini_batch=n)
side = GpuSide(w, 0, 0, 1)
for _ in range(2):
    side.step_device()
torch.cuda.synchronize()
print("ok")
