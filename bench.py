#!/usr/bin/env python
"""bench.py — PPO/A2C-update transitions/s (and GAE HBM GB/s) on B200, next to the reference's CPU path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config cfg1] [--impl ours|reference]

A "step" is one pass of the hot path over one synthetic rollout: RolloutStorage::compute_returns followed by
PPO::update / A2C::update (all epochs, all minibatches, Adam/RMSprop steps, normaliser update).
  value : T*N / step time with every input already resident in HBM (CUDA events, L2 flushed between steps)
  e2e   : the same through the public API with HOST buffers: pinned H2D of the rollout, host-drawn permutations
          (torch.randperm, as the reference's generator does) and their H2D, one D2H of the statistics
Prints ONE JSON line on rank 0.  `--impl reference` times the reference itself (oracle/_ref, CPU libtorch).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import numpy as np  # noqa: E402

from pytorch_cpp_rl_b200.workloads import CONFIGS, Workload, flops_per_row, init_params, synth_rollout  # noqa: E402

METRIC = "ppo_update_transitions_per_sec"
UNIT = "transitions/s"


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.lines, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])), mx.append(float(f[1]))
            except ValueError:
                continue
            for name, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def make_ref_session(w: Workload, seed: int):
    from oracle import ref_binding as ref
    from tests import helpers as H
    sess = H.make_ref_session(ref, w, init_seed=seed)
    r = synth_rollout(w, seed)
    H.fill_policy_outputs(sess, w, r)
    H.ref_load_rollout(sess, r)
    return ref, sess, r


def cpu_sample_workload(w: Workload):
    """Bounded sample of the workload for the CPU arms -> (sample workload, fraction of the full update it covers).
    The recurrent and convolutional updates take tens of seconds to minutes per update on the host.  The GRU sample is
    fewer environments with the reference's own minibatch shape (one environment x T steps per minibatch either way);
    the CNN sample keeps the FULL rollout and the full 8192-row minibatches and runs ONE of the epochs (small conv
    batches would understate the CPU's per-row throughput), its time is scaled by the epoch count."""
    if w.base == "gru" and w.N > 32:
        return w.scaled(N=32, num_mini_batch=32), 1.0
    if w.base == "cnn" and w.algo == "ppo" and w.num_epoch > 1:
        return w.scaled(num_epoch=1), 1.0 / w.num_epoch
    return w, 1.0


def config_of(w: Workload, world: int) -> dict:
    """The `config` object of the JSON line: the SAME for the GPU arm and the reference arm of one workload."""
    return {"workload": f"{w.name}: {w.description}", "rows_per_step": w.rows, "optimizer_steps": w.optimizer_steps,
            "l2": "GPU arm: flushed between timed iterations (256 MiB write)",
            "parallelism": "single GPU" if world == 1 else
            f"{world} independent replicas (the recurrent path does not shard: SURVEY 8e)" if w.base == "gru" else
            f"dp{world}: {world * w.N} envs sharded by environment over {world} GPUs ({w.N} per GPU), global "
            f"minibatch {world * w.rows // max(1, w.num_mini_batch)} rows, gradient all-reduce "
            + ("kernel between backward and optimiser" if w.base == "cnn" else "inside the fused kernel") +
            " over NVLink peer memory"}


def sample_text(ws: Workload, full: Workload, frac: float) -> str:
    if frac != 1.0:
        return (f"{ws.num_epoch} of the {full.num_epoch} epochs of the full {full.name} rollout at the full minibatch size "
                f"({ws.rows // max(1, ws.num_mini_batch)} rows), time scaled by {1 / frac:.0f}")
    return f"{ws.N} of the {full.N} environments of {full.name} ({ws.rows} rows, same rows per minibatch)"


def time_reference(w: Workload, steps: int, warmup: int, threads: int, budget_s: float = 1e9):
    """The reference's own CPU implementation (oracle/_ref): compute_returns + update per step."""
    ref, sess, r = make_ref_session(w, 0)
    ref.load().ref_set_num_threads(threads)
    total, done = 0.0, 0
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        sess.compute_returns(r["next_value"], w.use_gae, w.gamma, w.tau)
        sess.update(1.0, perm_seed=1000 + i)
        dt = time.perf_counter() - t0
        if i >= warmup:
            total += dt
            done += 1
            if total > budget_s:
                break
    return total / max(done, 1), done, ref.load().ref_torch_version().decode()


def run_reference(args, w: Workload, rank: int, world: int):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    full = w
    w, frac = cpu_sample_workload(w)
    sampled = w is not full
    sec, done, tv = time_reference(w, args.steps, min(args.warmup, 1) if sampled else args.warmup, threads,
                                   budget_s=120.0 if sampled else 1e9)
    sec /= frac  # a sample that runs a fraction of the epochs stands for that fraction of the update's time
    value = w.rows / sec
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_of(full, args.gpus), "reference_build": f"oracle/_ref (libtorch {tv} CPU)",
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "reference",
                         "sample": f"{done} x (compute_returns + update) of " +
                                   (sample_text(w, full, frac) if sampled else f"the full {w.name} rollout")},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


class GpuSide:
    """Policy, storage and algorithm resident on one GPU, plus pinned host copies of the rollout for the e2e leg."""

    def __init__(self, w: Workload, seed: int, rank: int = 0, world: int = 1):
        """world > 1: this rank holds w.N environments of a global rollout of world * w.N (weak scaling); parameters
        are replicated (same init seed), the rollout shard is seeded per rank."""
        import torch
        from pytorch_cpp_rl_b200 import api
        self.torch, self.api, self.w = torch, api, w
        self.rank, self.world = rank, world
        param_seed, seed = seed, seed + rank
        space = api.ActionSpace("Discrete" if w.head == "categorical" else "Box", [w.num_actions])
        self.policy = api.Policy(space, w.obs_dim, normalize_observations=w.normalize_obs, hidden=w.hidden,
                                 recurrent=(w.base == "gru"), obs_shape=w.obs_shape if w.base == "cnn" else None)
        self.policy.params.copy_(torch.from_numpy(init_params(w, param_seed)))
        self.st = api.RolloutStorage(w.T, w.N, w.obs_shape, space, w.storage_hidden)
        r = synth_rollout(w, seed)
        self.fields = ["observations", "rewards", "masks", "actions", "value_predictions", "action_log_probs"]
        for name in self.fields + ["hidden_states"]:
            getattr(self.st, name).copy_(torch.from_numpy(r[name]))
        self.next_value = torch.from_numpy(r["next_value"]).cuda()
        # value_predictions[:T] / action_log_probs = this policy's own outputs at its initial weights (§8d)
        T, N = w.T, w.N
        gru = w.base == "gru"
        out = self.policy.evaluate_actions(self.st.observations[:T].reshape(T * N, -1),
                                           self.st.hidden_states[0].contiguous() if gru else None,
                                           self.st.masks[:T].reshape(T * N, 1).contiguous() if gru else None,
                                           self.st.actions.reshape(T * N, -1))
        self.st.value_predictions[:T].copy_(out[0].view(T, N, 1))
        self.st.action_log_probs.copy_(out[1].view(T, N, 1))
        torch.cuda.synchronize()
        self.pinned = {n: getattr(self.st, n).cpu().pin_memory() for n in self.fields}
        self.pinned_nv = self.next_value.cpu().pin_memory()
        if w.algo == "ppo":
            self.algo = api.PPO(self.policy, w.clip_param, w.num_epoch, w.num_mini_batch, w.actor_loss_coef,
                                w.value_loss_coef, w.entropy_coef, w.learning_rate, w.epsilon, 0.5, w.kl_target)
            g = torch.Generator().manual_seed(param_seed)  # the same GLOBAL permutations on every rank
            # recurrent policies draw randperm(N) per epoch (recurrent_generator.cpp:38), the others randperm(T*N)
            self.perm_len = w.N if gru else w.rows * world
            self.d_perms = [torch.randperm(self.perm_len, generator=g).cuda() for _ in range(w.num_epoch)]
        else:
            self.algo = api.A2C(self.policy, 1.0, w.value_loss_coef, w.entropy_coef, w.learning_rate, w.epsilon, w.alpha)
        self.comm = None
        if world > 1 and not gru:  # recurrent path: replicas only (one environment per minibatch: nothing to shard)
            # environments sharded over the GPUs of the box; gradient all-reduce inside the fused kernel (NVLink)
            self.comm = api.Communicator(self.policy.params.numel())
            if w.algo == "ppo":
                self.algo.set_communicator(self.comm, rank * w.N, world * w.N)
            else:
                self.algo.set_communicator(self.comm, world * w.N)
        self.e2e_calls = 0
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in self.pinned.values()) + self.pinned_nv.numel() * 4
        if w.algo == "ppo" and world == 1:  # (sharded updates draw the global permutations on the device)
            self.h2d_bytes += w.num_epoch * self.perm_len * 8
        self.d2h_bytes = 64

    def step_device(self):
        """Inputs resident in HBM, nothing synchronises."""
        w = self.w
        self.st.compute_returns(self.next_value, w.use_gae, w.gamma, w.tau)
        if w.algo == "ppo":
            self.algo.begin(self.st)
            for p in self.d_perms:
                self.algo.epoch_device(p)
            self.algo.end_async()
        else:
            self.algo.update_async(self.st)

    def step_e2e(self):
        """Public API with host buffers: H2D of the rollout, host permutations, D2H of the statistics."""
        w = self.w
        for n in self.fields:
            getattr(self.st, n).copy_(self.pinned[n], non_blocking=True)
        self.next_value.copy_(self.pinned_nv, non_blocking=True)
        self.st.compute_returns(self.next_value, w.use_gae, w.gamma, w.tau)
        if self.world > 1:
            # every rank must draw the same global permutations from its CPU generator
            self.e2e_calls += 1
            self.torch.manual_seed(12345 + self.e2e_calls)
        return self.algo.update(self.st)


def kernel_time(lib, family: str):
    import ctypes as C
    ms, n = C.c_double(), C.c_int64()
    lib.cpprl_b200_kernel_time_ms(family.encode(), C.byref(ms), C.byref(n))
    return ms.value, n.value


def gae_sweep(torch, api, hbm_peak, T=256, N=1 << 20, iters=5):
    """GAE HBM GB/s at a bandwidth-sweep size (inputs 4.3 GB >> L2): algorithmic 4*N*(4T+3) bytes per launch."""
    space = api.ActionSpace("Discrete", [4])
    st = api.RolloutStorage(T, N, [1], space, 1)
    st.rewards.normal_()
    st.value_predictions.normal_()
    nv = torch.randn(N, 1, device="cuda")
    for _ in range(3):
        st.compute_returns(nv, True, 0.99, 0.95)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        a.record()
        st.compute_returns(nv, True, 0.99, 0.95)
        b.record()
    torch.cuda.synchronize()
    ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    nbytes = 4 * N * (4 * T + 3)
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"kernel": "returns_scan_kernel", "bound": "hbm", "workload": f"T={T} N={N} GAE", "bytes_per_launch": nbytes,
            "ms_per_launch": ms, "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak}


def insert_sweep(torch, api, hbm_peak, T=2, N=4096, iters=8):
    """RolloutStorage::insert HBM GB/s with Atari frames (cpprl_b200_storage_insert, one launch for all seven
    fields): N x 4x84x84 fp32 = 462 MB read and 462 MB written per launch (>> L2)."""
    space = api.ActionSpace("Discrete", [6])
    st = api.RolloutStorage(T, N, (4, 84, 84), space, 1)
    rows = [torch.rand(N, 4, 84, 84, device="cuda") for _ in range(2)]
    small = [torch.zeros(N, 1, device="cuda") for _ in range(5)]
    act = torch.zeros(N, 1, dtype=torch.int64, device="cuda")

    def one(i):
        st.insert(rows[i % 2], small[0], act, small[1], small[2], small[3], small[4])
    for i in range(3):
        one(i)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for i, (a, b) in enumerate(ev):
        a.record()
        one(i)
        b.record()
    torch.cuda.synchronize()
    ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    row_bytes = 4 * (4 * 84 * 84 + 1 + 1 + 1 + 1 + 1) + 8  # observation, hidden, log-prob, value, reward, mask, int64 action
    nbytes = 2 * N * row_bytes
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"kernel": "storage_copy_kernel", "bound": "hbm", "workload": f"insert N={N} frames 4x84x84 (SURVEY 8f row 2)",
            "bytes_per_launch": nbytes, "ms_per_launch": ms, "achieved": gbs, "peak": hbm_peak, "unit": "GB/s",
            "frac": gbs / hbm_peak}


def run_ours(args, w: Workload, rank: int, local_rank: int, world: int):
    import torch
    import torch.distributed as dist
    from pytorch_cpp_rl_b200 import api, cabi
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = cabi.load()
    hbm_peak, peak_src, sm_max = measured_peaks()
    side = GpuSide(w, seed=0, rank=rank, world=world)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident leg ------------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        side.step_device()
    barrier()
    lib.cpprl_b200_kernel_time_enable(1)
    lib.cpprl_b200_kernel_time_reset()
    launches0 = lib.cpprl_b200_launch_count()
    events = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local_rank) as clocks:
        barrier()
        t_wall0 = time.perf_counter()
        for a, b in events:
            flush.zero_()  # L2 flush between timed iterations (outside the event pair)
            a.record()
            side.step_device()
            b.record()
        barrier()
        t_wall = time.perf_counter() - t_wall0
        launches = lib.cpprl_b200_launch_count() - launches0
        lib.cpprl_b200_kernel_time_enable(0)
        # keep the sampler alive long enough to see the load on very short runs
        if world > 1:
            for _ in range(20):  # the same count on every rank: the sharded kernels of all ranks must run in lock step
                side.step_device()
            torch.cuda.synchronize()
        elif t_wall < 1.0:
            t_end = time.perf_counter() + 1.0
            while time.perf_counter() < t_end:
                side.step_device()
            torch.cuda.synchronize()
    step_ms = [a.elapsed_time(b) for a, b in events]
    total_ms = float(sum(step_ms))
    upd_ms, upd_n = kernel_time(lib, "update")
    ret_ms, ret_n = kernel_time(lib, "returns")
    pro_ms, pro_n = kernel_time(lib, "prologue")
    rms_ms, rms_n = kernel_time(lib, "rms")
    side.algo.fetch_stats()

    # ---- e2e leg ----------------------------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        side.step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        side.step_e2e()
    barrier()
    e2e_s = (time.perf_counter() - t0) / args.steps

    # the same end-to-end call with the opt-in device-drawn permutations (api.PPO.device_permutations): what is left of
    # the gap between `value` and `e2e` once the host randperm (the reference's own CPU random stream) is out of the way
    e2e_dev_s = None
    if world == 1 and w.algo == "ppo" and w.base != "gru":
        side.algo.device_permutations = True
        side.step_e2e()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            side.step_e2e()
        barrier()
        e2e_dev_s = (time.perf_counter() - t0) / args.steps
        side.algo.device_permutations = False

    if world > 1:
        t = torch.tensor([total_ms, e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, e2e_s = float(t[0]), float(t[1])
    if rank != 0:
        return
    ms_per_step = total_ms / args.steps
    value = world * w.rows / (ms_per_step * 1e-3)
    e2e_value = world * w.rows / e2e_s

    # ---- roofline of the dominant kernel ---------------------------------------------------------------
    fp32_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12  # TFLOP/s, nominal (no fp32 entry in MEASURED_PEAKS.json)
    rows_per_launch = w.rows
    flops_per_launch = rows_per_launch * flops_per_row(w)
    upd_avg_ms = upd_ms / max(upd_n, 1)
    achieved = flops_per_launch / (upd_avg_ms * 1e-3) / 1e12 if upd_n else 0.0
    roofline = {
        "kernel": {"mlp": "mlp_update_kernel", "gru": "gru_update_kernel",
                   "cnn": "cnn_gemm_kernel (all instantiations of one epoch)"}[w.base], "bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s",
        "frac": achieved / fp32_peak, "traffic": None,
        "flops_per_launch": flops_per_launch, "ms_per_launch": upd_avg_ms, "launches_timed": upd_n,
        "peak_source": f"nominal 148 SM x 128 FMA/clk x 2 x {sm_max:.0f} MHz; fp32 CUDA-core bound, neither hbm nor tensor "
                       "(see DESIGN.md); share of step = " + f"{upd_ms / total_ms:.2f}",
    }
    tc_rows = lib.cpprl_b200_mlp_update_variant(w.rows // max(1, w.num_mini_batch), w.obs_dim, w.num_actions) if w.base == "mlp" else 0
    mlp_tc = tc_rows > 0
    if mlp_tc:
        # minibatches of more than 32 or at most 16 rows per SM train on the tensor-core variant (tcgen05 kind::tf32 as 3xTF32,
        # csrc/mlp_tc.cuh): three TF32 MMAs per algorithmic product at half the bf16 rate -> this path's own ceiling is the
        # measured bf16 peak / 6.  What bounds the kernel is NOT the tensor pipe (ncu: 13 % active) but the chain of dependent stages of each of
        # the update's sequential optimiser steps (two grid barriers per step): DESIGN.md section 4.
        pk = ROOT / "MEASURED_PEAKS.json"
        peaks = json.loads(pk.read_text()) if pk.exists() else {}
        tc_peak = peaks.get("bf16_tflops_sustained") or peaks.get("bf16_tflops") or 2250.0
        roofline.update({
            "kernel": f"mlp_update_kernel<{tc_rows}-row tiles, tensor cores> (tcgen05 3xTF32 towers + CUDA-core heads, loss, reduction, Adam)",
            "bound": "tensor", "peak": tc_peak, "frac": achieved / tc_peak,
            "ceiling_3xtf32": tc_peak / 6.0, "frac_of_3xtf32_ceiling": achieved / (tc_peak / 6.0),
            "frac_of_nominal_fp32_cuda_core_peak": achieved / fp32_peak,
            "peak_source": ("measured dense bf16 (MEASURED_PEAKS.json, sustained)" if peaks else "fallback 2250 nominal bf16") +
                           "; achieved counts ALGORITHMIC fp32 flops (2 x MACs of forward + both backward contractions); latency-bound: "
                           f"{w.optimizer_steps} dependent optimiser steps per update; share of step = {upd_ms / total_ms:.2f}"})
    if w.name == "cfg3" and world == 1:
        # dram__bytes_read.sum + dram__bytes_write.sum of one launch (= one epoch) from the committed ncu --set full capture;
        # the algorithmic bytes of an epoch are the gathered rollout rows, 262144 x 112 B = 29.4 MB: no wasted re-reads
        roofline["traffic"] = (29662464 + 775680) if mlp_tc else (29616128 + 353792)
        roofline["traffic_source"] = "profiles/r2c_mlp_update_tc_cfg3_raw.csv" if mlp_tc else "profiles/r1b_mlp_update_cfg3_raw.csv"
    if w.base == "cnn" and os.environ.get("CPPRL_B200_CNN_FP32") != "1":
        # the ten CnnBase contractions run on tcgen05 as 3xTF32 (hi*hi + hi*lo + lo*hi): three TF32 MMAs per
        # algorithmic product at half the bf16 rate, so this path's own ceiling is the measured bf16 peak / 6
        pk = ROOT / "MEASURED_PEAKS.json"
        peaks = json.loads(pk.read_text()) if pk.exists() else {}
        tc_peak = peaks.get("bf16_tflops_sustained") or peaks.get("bf16_tflops") or 2250.0
        roofline.update({
            "kernel": "tc::gemm_tc_kernel (tcgen05 3xTF32; all launches of one epoch: 10 contractions x passes, heads, "
                      "column sums, optimiser)",
            "bound": "tensor", "peak": tc_peak, "frac": achieved / tc_peak,
            "ceiling_3xtf32": tc_peak / 6.0, "frac_of_3xtf32_ceiling": achieved / (tc_peak / 6.0),
            "peak_source": ("measured dense bf16 (MEASURED_PEAKS.json, sustained)" if peaks else "fallback 2250 nominal bf16") +
                           "; achieved counts ALGORITHMIC fp32 flops (2 x MACs of forward + both backward contractions); "
                           "fp32-accurate results need 3 TF32 MMAs per product at bf16/2 rate -> ceiling = peak / 6; "
                           f"share of step = {upd_ms / total_ms:.2f}"})
    stages = [gae_sweep(torch, api, hbm_peak), insert_sweep(torch, api, hbm_peak)]
    for st in stages:
        st["peak_source"] = peak_src
    shares = {"update": upd_ms / total_ms, "returns": ret_ms / total_ms, "prologue": pro_ms / total_ms,
              "rms": rms_ms / total_ms}

    # ---- CPU baseline: the reference itself on this box's host cores ----------------------------------
    cpu = None
    if world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        ws, frac = cpu_sample_workload(w)
        sampled = ws is not w
        wu = 1 if sampled else 2
        secN, doneN, tv = time_reference(ws, 50, wu, threads, budget_s=20.0 if sampled else 8.0)
        secN /= frac
        # the single-thread leg only where it can win (small MLP rollouts); the heavy bases scale with threads
        sec1, done1 = (secN * threads, 0) if sampled else time_reference(ws, 50, wu, 1, budget_s=8.0)[:2]
        best, cores, done = (sec1, 1, done1) if sec1 <= secN else (secN, threads, doneN)
        what = sample_text(ws, w, frac) if sampled else f"the full {w.name} rollout"
        cpu = {"value": ws.rows / best, "unit": UNIT, "cores": cores, "kind": "reference",
               "sample": f"{done} x (compute_returns + update) of {what}, oracle/_ref libtorch {tv}",
               "value_1_thread": (ws.rows / sec1) if done1 else None, "value_all_threads": ws.rows / secN,
               "host_threads": threads}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": config_of(w, world),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": side.h2d_bytes, "d2h_bytes_per_step": side.d2h_bytes,
                "ms_per_step": e2e_s * 1e3,
                "permutations": "device (sharded update)" if world > 1 and w.base != "gru" else
                "the reference's CPU random stream (torch's CPU generator state in / out), drawn by the library's host threads "
                "while the device runs: bit-identical to torch.randperm",
                "with_device_permutations": None if e2e_dev_s is None else
                {"value": world * w.rows / e2e_dev_s, "ms_per_step": e2e_dev_s * 1e3,
                 "h2d_bytes_per_step": side.h2d_bytes - w.num_epoch * side.perm_len * 8}},
        "gpu_launches": int(launches), "roofline": roofline, "stages": stages, "kernel_time_share": shares,
        "cpu_baseline": cpu, "clocks": clocks.summary(), "wall_ms_per_step_incl_flush": t_wall * 1e3 / args.steps,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg3", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    w = CONFIGS[args.config]
    if args.impl == "reference":
        run_reference(args, w, rank, world)
    else:
        run_ours(args, w, rank, local_rank, world)
        if world > 1:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.destroy_process_group()


if __name__ == "__main__":
    main()
