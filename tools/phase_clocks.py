#!/usr/bin/env python
"""Per-phase cycle breakdown of the fused update kernel (clock64 markers, see cpprl_b200_debug_phase_clocks)."""
import ctypes as C
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from bench import GpuSide
from pytorch_cpp_rl_b200 import cabi
from pytorch_cpp_rl_b200.workloads import CONFIGS

NAMES = ["img+stage", "P1 L1", "P2 L2", "P3 heads", "P4 loss", "P5 head-bwd", "P6 dW2", "P7 dz1", "P8 dW1", "publish",
         "barrier1", "reduce", "optimiser", "arrive+stage", "barrier2 wait"]
import os
cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
step = int(sys.argv[2]) if len(sys.argv) > 2 else 2
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
lib.cpprl_b200_debug_phase_clocks(C.c_void_p(buf.data_ptr()), step)
side.step_device()
torch.cuda.synchronize()
lib.cpprl_b200_debug_phase_clocks(None, 0)
c = buf.cpu().numpy()[:148 * 32].reshape(148, 32)
used = c[:, 0] != 0
c = c[used]
if rank != 0:
    torch.cuda.synchronize()
    sys.exit(0)
# CTAs beyond the tile count only take part in phase 2 (their tile marks stay zero): statistics over the CTAs with tiles
has_tile = c[:, 3] != 0
print(f"{cfg}: {used.sum()} CTAs ({has_tile.sum()} with tiles), step {step}; cycles per phase (min / median / max over the CTAs with tiles); "
      "multi-tile CTAs show the LAST tile")
c = c[has_tile]
d = np.diff(c[:, :16], axis=1)
last = 15 if c[:, 15].all() else 13  # the last optimiser step of a launch has no second barrier (marks 14, 15 stay zero)
for i, n in enumerate(NAMES[:last]):
    print(f"  {n:12s} {d[:, i].min():8d} {int(np.median(d[:, i])):8d} {d[:, i].max():8d}")
if c[:, 20].any():
    # tensor-core variant: finer marks (slots 20..26) between the standard ones
    seq = [("step top -> image landed", 0, 26), ("weights -> tensor memory", 26, 1), ("finalize_stage", 1, 20), ("T1 issue+wait", 20, 2),
           ("E1 (tanh, operand planes)", 2, 21), ("T2 issue+wait", 21, 22), ("E2 (tanh, h2)", 22, 23), ("W2^T -> tensor memory", 23, 3),
           ("P3 heads", 3, 4), ("P4 loss", 4, 5), ("P5 head-bwd + dz2 planes", 5, 6), ("T3 issue+wait", 6, 7), ("E3 dz1", 7, 8),
           ("T4 chain 1 + drain", 8, 24), ("T4 chain 2 + drain", 24, 25), ("tile end", 25, 9)]
    print("  tensor-core variant, finer marks:")
    for name, i0, i1 in seq:
        dd = c[:, i1] - c[:, i0]
        print(f"    {name:28s} {dd.min():8d} {int(np.median(dd)):8d} {dd.max():8d}")
if c[:, 30].any():
    print("  barrier 1 in detail:")
    for name, i0, i1 in [("publish done -> arrived (CTA barrier + release)", 10, 30), ("next minibatch's first tile laid out", 30, 31),
                         ("poll + CTA barrier", 31, 11)]:
        dd = c[:, i1] - c[:, i0]
        print(f"    {name:48s} {dd.min():8d} {int(np.median(dd)):8d} {dd.max():8d}")
tot = c[:, last] - c[:, 0]
if world > 1:
    x = np.diff(c[:, 16:20], axis=1)
    for i, n in enumerate(["x: push to peers", "x: poll + add", "x: barrier"]):
        print(f"  {n:16s} {x[:, i].min():8d} {int(np.median(x[:, i])):8d} {x[:, i].max():8d}")
print(f"  total        {tot.min():8d} {int(np.median(tot)):8d} {tot.max():8d}  (= {np.median(tot) / 1.85e3:.1f} us at the 1.85 GHz the SMs run at under this kernel: tools/step_totals.py)")
