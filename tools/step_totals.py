#!/usr/bin/env python
"""Cycle totals of several minibatch steps of the fused MLP update (clock64 markers): where in an epoch the time goes.

    python tools/step_totals.py cfg3            (or under torchrun for the env-sharded update)
"""
import ctypes as C
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from bench import GpuSide
from pytorch_cpp_rl_b200 import cabi
from pytorch_cpp_rl_b200.workloads import CONFIGS

cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
w = CONFIGS[cfg]
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
side = GpuSide(w, 0, rank, world)
lib = cabi.load()
buf = torch.zeros(148 * 32 + 128, dtype=torch.int64, device="cuda")
for _ in range(3):
    side.step_device()
torch.cuda.synchronize()
rows = []
for step in [0, 1, 2, 3, 8, 16, 24, 30, 31]:
    if step >= max(1, w.num_mini_batch):
        continue
    buf.zero_()
    lib.cpprl_b200_debug_phase_clocks(C.c_void_p(buf.data_ptr()), step)
    side.step_device()
    torch.cuda.synchronize()
    lib.cpprl_b200_debug_phase_clocks(None, 0)
    c = buf.cpu().numpy()[:148 * 32].reshape(148, 32)
    c = c[c[:, 0] != 0]
    last = 15 if step + 1 < w.num_mini_batch else 13
    tot = c[:, last] - c[:, 0]
    line = f"step {step:2d}: {len(c)} CTAs, total median {int(np.median(tot)):7d} max {tot.max():7d} cycles; barrier1 {int(np.median(c[:, 11] - c[:, 10])):6d}; reduce {int(np.median(c[:, 12] - c[:, 11])):6d}"
    if world > 1:
        line += f"; exchange poll median {int(np.median(c[:, 18] - c[:, 17])):6d} max {int((c[:, 18] - c[:, 17]).max()):6d}"
    if step == 0:
        line += f"\n         kernel: entry -> prologue done {int(np.median(c[:, 28] - c[:, 27]))} cycles, whole kernel {int(np.median(c[:, 29] - c[:, 27]))} cycles"
    rows.append(line)
# epoch launches timed with events (steady state): kernel cycles vs wall time shows the launch gap
# %globaltimer at entry / exit of CTA 0 of consecutive epoch launches: kernel durations and the gaps between them
buf.zero_()
lib.cpprl_b200_debug_phase_clocks(C.c_void_p(buf.data_ptr()), 10 ** 6)
side.step_device()
side.step_device()
torch.cuda.synchronize()
lib.cpprl_b200_debug_phase_clocks(None, 0)
g = buf.cpu().numpy()[148 * 32:].reshape(64, 2)
g = g[g[:, 0] != 0]
g = g[np.argsort(g[:, 0])]
if rank == 0:
    print("epoch launches (CTA 0, %globaltimer): duration us", [round((b - a) / 1e3, 1) for a, b in g],
          "gap to next us", [round((g[i + 1, 0] - g[i, 1]) / 1e3, 1) for i in range(len(g) - 1)])
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
for _ in range(5):
    side.step_device()
ev[1].record()
torch.cuda.synchronize()
if rank == 0:
    print("\n".join(rows))
    print(f"5 updates: {ev[0].elapsed_time(ev[1]) / 5:.3f} ms per update ({w.num_epoch} epoch launches each)")
torch.cuda.synchronize()
