#!/usr/bin/env python
"""A few RolloutStorage.insert / after_update / concatenate calls on Atari-shaped frames for ncu captures of
storage_copy_kernel, plus a CUDA-event sweep of the insert bandwidth over the environment count."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from pytorch_cpp_rl_b200 import api


def insert_ms(N, T=2, iters=10):
    space = api.ActionSpace("Discrete", [6])
    st = api.RolloutStorage(T, N, (4, 84, 84), space, 1)
    rows = [torch.rand(N, 4, 84, 84, device="cuda") for _ in range(2)]
    small = [torch.zeros(N, 1, device="cuda") for _ in range(5)]
    act = torch.zeros(N, 1, dtype=torch.int64, device="cuda")
    for i in range(3):
        st.insert(rows[i % 2], small[0], act, small[1], small[2], small[3], small[4])
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(iters):
        st.insert(rows[i % 2], small[0], act, small[1], small[2], small[3], small[4])
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / iters
    nbytes = 2 * N * (4 * (4 * 84 * 84 + 5) + 8)
    st.after_update()
    return {"N": N, "ms": ms, "GBps": nbytes / ms / 1e6, "bytes": nbytes}, st


if __name__ == "__main__":
    out = []
    keep = []
    for N in (256, 1024, 4096):
        r, st = insert_ms(N)
        out.append(r)
        if N == 256:
            keep = [st, insert_ms(256)[1]]
    joined = api.RolloutStorage.concatenate(keep)
    torch.cuda.synchronize()
    print(json.dumps(out))
