#!/usr/bin/env python
"""Headline benchmark: MNIST-CNN training throughput (images/s) at batch 200 per rank.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Metric and config are BASELINE.json's: the reference's 20-epoch run at batch 200/rank
(README.md:42,48) — reported as whole-job images/s (20*60000 / time) so that 1/2/4/8-GPU runs are
comparable; ``epoch20_time_s`` in the JSON line is the same number expressed as the reference's
wall-time metric.  Synthetic MNIST-shaped data (60k x 28 x 28 uint8, random labels), random-init
weights (seed 1), fp32 parameters with TF32 tensor-core math where the engine uses tcgen05.

Timed regions (both bracketed by barrier + synchronize, max over ranks):
  value : K optimizer steps on the HBM-resident dataset, each step timed by its own CUDA-event
          pair with an L2 flush (256 MiB memset) between steps.
  e2e   : K steps through the public streaming API ``NativeEngine.train_step_host`` — every
          step copies its uint8 batch from pinned host memory and the step's loss is copied
          back to the host — timed as one region with CUDA events.
The reference arm (--impl reference) is baseline/reference_arm.py: the unmodified scripts
installed under baseline/_ref, driven through their own train() function.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PUBLISHED_IMG_S = {1: 20 * 60000 / 242.3, 2: 20 * 60000 / 137.1, 4: 20 * 60000 / 73.6}  # README.md:57-59
BATCH = 200


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int) -> None:
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        # A timed region of a few tens of milliseconds is shorter than nvidia-smi's polling period, so the same
        # quantities are also polled through NVML every 5 ms from a thread (best effort: any failure -> skipped)
        self.nvml_rows, self._nvml_stop = [], threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))

            def poll():
                while not self._nvml_stop.is_set():
                    try:
                        sm = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                        try:
                            mask = int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(h))
                        except Exception:
                            mask = int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                        try:
                            power = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                        except Exception:
                            power = None
                        self.nvml_rows.append((sm, mx, power, mask))
                    except Exception:
                        return
                    time.sleep(0.005)

            self.nvml_thread = threading.Thread(target=poll, daemon=True)
            self.nvml_thread.start()
        except Exception:
            self.nvml_thread = None

    def n_samples(self) -> int:
        return len(self.rows) + len(getattr(self, "nvml_rows", []))

    def _pump(self) -> None:
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self) -> dict:
        if getattr(self, "nvml_thread", None) is not None:
            self._nvml_stop.set()
            self.nvml_thread.join(timeout=1)
        nvml_rows = list(getattr(self, "nvml_rows", []))
        if self.proc is None and not nvml_rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi and NVML unavailable"]}
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        smi_samples = len(sm)
        bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        for s_, m_, p_, mask in nvml_rows:
            sm.append(s_); mx.append(m_)
            if p_ is not None:
                power.append(p_)
            for name, bit in bits.items():
                if mask & bit:
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm),
                "samples_nvidia_smi": smi_samples, "samples_nvml": len(nvml_rows), "reasons": sorted(reasons)}


def _dist_env():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    return rank, world, local


def run_ours(args) -> dict:
    import torch
    import torch.distributed as dist
    from pytorch_mnist_ddp_b200 import _native
    from pytorch_mnist_ddp_b200.engine import B200Backend, LOSS_RING
    from pytorch_mnist_ddp_b200.parallel.dist import DistInfo
    from pytorch_mnist_ddp_b200.parallel.sampler import distributed_indices
    from pytorch_mnist_ddp_b200.utils.data import synthetic_split

    rank, world, local = _dist_env()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local)
    info = DistInfo(False, 0, 1, local, "")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", init_method="env://", world_size=world, rank=rank,
                                device_id=torch.device("cuda", local))
        info = DistInfo(True, rank, world, local, "nccl")
    torch.manual_seed(1)
    train, test = synthetic_split(True), synthetic_split(False, 1000)
    backend = B200Backend(train, test, info, lr=1.0, seed=1, comm=args.comm, max_batch=max(args.batch, 1000))
    eng = backend.engine
    torch.cuda.set_stream(eng.stream)  # events, flushes and copies below are ordered with the engine
    B, K, W = args.batch, args.steps, args.warmup
    per_rank = len(train) // world
    steps_per_epoch = per_rank // B
    state = {"epoch": 0, "i": steps_per_epoch}

    def next_step_setup():
        """Start a new (shuffled, DistributedSampler-exact) epoch when the cursor hits the end."""
        if state["i"] >= steps_per_epoch:
            state["epoch"] += 1
            state["i"] = 0
            state["idx"] = distributed_indices(len(train), world, rank, state["epoch"])
            eng.begin_epoch(state["idx"])
        state["i"] += 1

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    backend.use_graph = not args.no_graph
    for _ in range(max(W, 3)):
        next_step_setup()
        backend.step(B)
    barrier()

    # ---- device-timed steps, L2 flushed between steps
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = _native.launch_count()
    barrier()
    t_wall0 = time.perf_counter()
    for k in range(K):
        next_step_setup()
        flush.zero_()
        starts[k].record()
        backend.step(B)
        ends[k].record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = _native.launch_count() - launches0
    step_ms = [s.elapsed_time(e) for s, e in zip(starts, ends)]
    total_ms = sum(step_ms)

    # ---- steady state (no flush), for information
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(K):
        next_step_setup()
        backend.step(B)
    e1.record()
    barrier()
    steady_ms = e0.elapsed_time(e1)

    # ---- end to end through the streaming API: pinned H2D of every batch + loss D2H every step
    idx = distributed_indices(len(train), world, rank, 99).long()
    n_host = min(per_rank, (K + W) * B)
    host_img = train.images[idx[:n_host]].reshape(-1, 784).contiguous().pin_memory()
    host_lbl = train.labels[idx[:n_host]].contiguous().pin_memory()
    eng.begin_epoch(None)
    nb = n_host // B

    steps_e2e = []

    def e2e_step(j, slot):
        o = (j % nb) * B
        # H2D: this step's uint8 batch from pinned host memory (copy stream, overlaps the previous step);
        # D2H: the step's head kernel stores its mean loss into mapped pinned host memory (4 bytes)
        steps_e2e.append(backend.step_host(host_img[o: o + B], host_lbl[o: o + B]))

    for j in range(max(W, 4)):
        e2e_step(j, K + j % 8)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for j in range(K):
        e2e_step(W + j, j)
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    # A very short run can end before the first clock sample arrives: keep the same workload running (untimed)
    # until there are a few samples, and say so in the record.
    extended, j_ext = False, W + K

    def too_few_samples() -> bool:
        # every rank must run the same number of steps (the fused all-reduce kernels wait for their peers),
        # so the decision is taken on the minimum over ranks
        n = torch.tensor([sampler.n_samples()], device="cuda", dtype=torch.int32)
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(n, op=dist.ReduceOp.MIN)
        return int(n.item()) < 3

    for _ in range(12):  # at most ~12 x 64 steps (~0.15 s of GPU time at batch 200)
        if not too_few_samples():
            break
        extended = True
        for _ in range(64):
            e2e_step(j_ext, 0)
            j_ext += 1
        torch.cuda.synchronize()
        time.sleep(0.02)
    clocks = sampler.stop()
    if extended:
        clocks["note"] = "timed region shorter than the sampling period: samples include untimed steps of the same workload run right after it"
    host_loss = torch.stack([eng.host_loss_ring[s % LOSS_RING] for s in steps_e2e[-min(K, LOSS_RING):]]).clone()
    losses_ok = bool(torch.isfinite(host_loss).all()) and float(host_loss.abs().sum()) > 0

    def reduce_max(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    total_ms, steady_ms, e2e_ms = reduce_max(total_ms), reduce_max(steady_ms), reduce_max(e2e_ms)
    imgs = world * B * K
    value = imgs / (total_ms / 1e3)
    out = {
        "metric": "mnist_cnn_train_images_per_sec (20-epoch batch-200/rank run, whole job)",
        "value": value, "unit": "images/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": (value / PUBLISHED_IMG_S[world]) if world in PUBLISHED_IMG_S else None,
        "dtype": "fp32 params/accumulate; tf32 tensor-core math where tcgen05 is used", "data": "synthetic",
        "impl": "ours",
        "config": {"model": "mnist_cnn Net (conv1-32, conv2-64, fc1 9216-128, fc2 128-10), Adadelta lr=1.0",
                   "global_batch": world * B, "batch_per_rank": B, "seq_len": None, "parallelism": f"dp{world}",
                   "l2": "flushed between timed steps (256 MiB memset), each step timed by its own event pair",
                   "comm": getattr(backend.ddp, "mode", "none"), "use_tc_mask": eng.use_tc,
                   "cuda_graph": not args.no_graph},
        "epoch20_time_s": 20 * 60000 / value,
        "steady_state": {"images_per_s": imgs / (steady_ms / 1e3), "ms_per_step": steady_ms / K,
                         "note": "same K steps back to back, no L2 flush"},
        "e2e": {"value": imgs / (e2e_ms / 1e3), "unit": "images/s", "ms_per_step": e2e_ms / K,
                "h2d_bytes_per_step": B * 784 + B, "d2h_bytes_per_step": 4,
                "api": "NativeEngine.train_step_host(pinned uint8 images, pinned uint8 labels); the loss of every "
                       "step is stored by the head kernel into mapped pinned host memory (zero-copy D2H)",
                "losses_finite": losses_ok},
        "gpu_launches": int(launches),
        "launches_per_step": eng.launches_per_step(B),
        "host_wall_s_timed_region": t_wall,
        "clocks": clocks,
        "last_loss": float(host_loss[-1]),
    }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out if rank == 0 else {}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--comm", default="auto")
    ap.add_argument("--no-graph", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        from baseline.reference_arm import run_reference
        out = run_reference(args)
    else:
        out = run_ours(args)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
