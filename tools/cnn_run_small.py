#!/usr/bin/env python
"""One CnnBase PPO update on 4096 rows (one pass per minibatch) for ncu captures of cnn_gemm_kernel."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from bench import GpuSide
from pytorch_cpp_rl_b200.workloads import CONFIGS

w = CONFIGS["cfg5"].scaled(N=32, num_epoch=1, num_mini_batch=1)
side = GpuSide(w, 0, 0, 1)
for _ in range(2):
    side.step_device()
torch.cuda.synchronize()
print("ok")
