#!/usr/bin/env python
"""Headline benchmark: MNIST-CNN data-parallel training, batch 200 per rank (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B] [--epochs E]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

BASELINE.json's metric is "MNIST-CNN 20-epoch time & images/s at 1/2/4/8 B200 vs ref's own NCCL DDP".  Both halves
are measured, on synthetic MNIST-shaped data (60k x 28 x 28 uint8, random labels), random-init weights (seed 1):

  value      images/s over K optimizer steps (whole job, all ranks), HBM-resident dataset, every step timed on the
             device by its own CUDA-event pair with an L2 flush (256 MiB memset) between steps, max over ranks.
  e2e        the same K steps through the public data-loader API (``B200Backend.step_gather``): every step gathers
             its batch by sampler index from the HOST-resident uint8 dataset into pinned memory, copies it to the
             GPU, trains, and the step's loss comes back to host memory -- one timed region, CUDA events.
  epoch20_*  the reference's own protocol (README.md:42,48; mnist_ddp.py:179-203) run for real, not extrapolated:
             E epochs x ceil(60000/W/B) steps with the ragged last batch, per-epoch DistributedSampler order,
             rank-0 evaluation of the 10k test images, StepLR, loss read-back every 10 steps -- through the same
             ``trainer.fit`` the CLI uses.  Device-timed (CUDA events around the whole job, max over ranks) and
             wall-clocked.

``vs_baseline`` divides by the reference README's published wall times (unknown hardware, data loading included);
the like-for-like comparison is the reference arm (--impl reference, baseline/reference_arm.py: the unmodified
scripts installed under baseline/_ref, driven through their own train()/test()) on the same box.
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
GOLDEN = os.path.join(ROOT, "benchmarks", "golden_losses.json")


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


METRIC = "mnist_cnn_train_images_per_sec (batch 200/rank data-parallel steps; 20-epoch job in epoch20_*)"
L2_POLICY = "flushed between timed steps (256 MiB memset), each step timed by its own event pair"


def shared_config(world: int, batch: int) -> dict:
    """The `config` entry of the JSON line.  Both arms print exactly this dict (arm-specific facts go to `impl_config`),
    so that the driver's same-config check compares like with like."""
    return {"model": "mnist_cnn Net (conv1 1-32 3x3, conv2 32-64 3x3, maxpool, dropout, fc1 9216-128, fc2 128-10), "
                     "Adadelta lr=1.0 StepLR gamma=0.7",
            "global_batch": world * batch, "batch_per_rank": batch, "seq_len": None, "parallelism": f"dp{world}",
            "l2": L2_POLICY, "dataset": "synthetic 60000x1x28x28 uint8 + 10000 test, random labels"}


def b200_env() -> dict:
    """Every B200_* switch that is set: a run with a non-default engine must say so in its own record."""
    return {k: v for k, v in sorted(os.environ.items()) if k.startswith("B200_")}


def check_golden(world: int, batch: int, dtype_key: str, losses) -> dict:
    """First losses of the run (seed 1, synthetic data, dropout on) against the stored trajectory of this
    configuration: a step that skipped work, dropped a gradient or mixed up a batch does not reproduce it."""
    key = f"w{world}_b{batch}_{dtype_key}"
    try:
        with open(GOLDEN) as f:
            golden = json.load(f)
    except Exception:
        golden = {}
    ref = golden.get(key)
    out = {"key": key, "first_losses": [round(float(x), 6) for x in losses]}
    if ref is None:
        out["status"] = "absent (no stored trajectory for this configuration)"
        return out
    n = min(len(ref), len(losses))
    err = max(abs(float(a) - float(b)) for a, b in zip(ref[:n], losses[:n])) if n else float("inf")
    out.update(status="ok" if err < 2e-3 else "MISMATCH", max_abs_err=err, compared=n)
    return out


class _JobArgs:
    """The CLI arguments of `mnist_ddp.py --batch-size B --epochs E` (README.md:42,48), for trainer.fit()."""

    def __init__(self, batch: int, epochs: int) -> None:
        self.batch_size, self.test_batch_size, self.epochs = batch, 1000, epochs
        self.lr, self.gamma, self.dry_run, self.log_interval = 1.0, 0.7, False, 10


def run_ours(args) -> dict:
    import contextlib
    import io
    import torch
    import torch.distributed as dist
    from pytorch_mnist_ddp_b200 import _native, trainer
    from pytorch_mnist_ddp_b200.engine import B200Backend, LOSS_RING, TC_FC1_FP32
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
    train, test = synthetic_split(True), synthetic_split(False)
    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    backend = B200Backend(train, test, info, lr=1.0, seed=1, comm=args.comm, max_batch=max(B, 1000))
    eng = backend.engine
    torch.cuda.set_stream(eng.stream)  # events, flushes and copies below are ordered with the engine
    p_init = eng.params.clone()
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

    def reduce_max(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    backend.use_graph = not args.no_graph
    first_steps = []
    for _ in range(W):
        next_step_setup()
        first_steps.append(backend.step(B))
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
        s = backend.step(B)
        ends[k].record()
        if len(first_steps) < 20:
            first_steps.append(s)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = _native.launch_count() - launches0
    first_losses = [float(eng.host_loss_ring[s % LOSS_RING]) for s in first_steps[:20]]
    step_ms = [s.elapsed_time(e) for s, e in zip(starts, ends)]
    total_ms = reduce_max(sum(step_ms))

    # ---- steady state (no flush), for information
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(K):
        next_step_setup()
        backend.step(B)
    e1.record()
    barrier()
    steady_ms = reduce_max(e0.elapsed_time(e1))

    # ---- end to end through the data-loader API: host gather by sampler index -> pinned -> H2D -> step -> loss on host
    idx_e2e = distributed_indices(len(train), world, rank, 99).to(torch.int32).contiguous()
    host_img, host_lbl = train.images.contiguous(), train.labels.contiguous()  # the dataset stays in host memory
    eng.begin_epoch(None)
    nb = per_rank // B
    steps_e2e = []

    def e2e_step(j):
        o = (j % nb) * B
        steps_e2e.append(backend.step_gather(host_img, host_lbl, idx_e2e[o: o + B]))

    for j in range(max(W, 4)):
        e2e_step(j)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for j in range(K):
        e2e_step(W + j)
    e1.record()
    barrier()
    e2e_ms = reduce_max(e0.elapsed_time(e1))
    # A very short run can end before the first clock sample arrives: keep the same workload running (untimed)
    # until there are a few samples, and say so in the record.
    extended, j_ext = False, W + K

    def too_few_samples() -> bool:
        # every rank must run the same number of steps (the fused all-reduce kernels wait for their peers),
        # so the decision is taken on the minimum over ranks
        n = torch.tensor([sampler.n_samples()], device="cuda", dtype=torch.int32)
        if world > 1:
            dist.all_reduce(n, op=dist.ReduceOp.MIN)
        return int(n.item()) < 3

    for _ in range(12):  # at most ~12 x 64 steps (~0.15 s of GPU time at batch 200)
        if not too_few_samples():
            break
        extended = True
        for _ in range(64):
            e2e_step(j_ext)
            j_ext += 1
        torch.cuda.synchronize()
        time.sleep(0.02)
    clocks = sampler.stop()
    if extended:
        clocks["note"] = ("timed region shorter than the sampling period: samples include untimed steps of the same "
                          "workload run right after it")
    host_loss = torch.stack([eng.host_loss_ring[s % LOSS_RING] for s in steps_e2e[-min(K, LOSS_RING):]]).clone()
    losses_ok = bool(torch.isfinite(host_loss).all()) and float(host_loss.abs().sum()) > 0

    # ---- the reference's protocol, run for real: E epochs through the CLI's own epoch driver
    job = None
    if args.epochs > 0:
        jargs = _JobArgs(B, args.epochs)
        eng.set_lr(1.0)
        buf = io.StringIO()
        barrier()
        j0, j1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        j0.record()
        tw = time.perf_counter()
        with contextlib.redirect_stdout(buf):
            stats = trainer.fit(jargs, backend, train, test, info, sampler_mode="distributed")
        j1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - tw
        barrier()
        dev_s = reduce_max(j0.elapsed_time(j1)) / 1e3
        wall = reduce_max(wall)
        per_rank_n = -(-len(train) // world)
        spe = -(-per_rank_n // B)
        imgs_job = args.epochs * len(train)
        tests = [l for l in buf.getvalue().splitlines() if l.startswith("Test set:")]
        job = {"epochs": args.epochs, "steps_per_epoch_per_rank": spe, "last_batch": per_rank_n - (spe - 1) * B,
               "optimizer_steps": int(stats["steps"]), "device_s": dev_s, "wall_s": wall,
               "images_per_s": imgs_job / dev_s,
               "includes": "per-epoch sampler order upload, ragged last batch, rank-0 eval of 10000 test images per "
                           "epoch, StepLR, loss read-back every 10 steps (trainer.fit, the CLI's epoch driver)",
               "last_test_line": tests[-1] if tests else None}

    # ---- sanity of what was timed
    p_now = eng.params.clone()
    moved = float((p_now - p_init).norm())
    drift = backend.ddp.replica_drift() if (world > 1 and backend.ddp is not None and args.comm != "nccl") else 0.0
    fp32fc = bool(eng.use_tc & TC_FC1_FP32)
    golden = check_golden(world, B, "fc3xtf32" if fp32fc else f"tc{eng.use_tc}", first_losses)

    imgs = world * B * K
    value = imgs / (total_ms / 1e3)
    out = {
        "metric": METRIC,
        "value": value, "unit": "images/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": (value / PUBLISHED_IMG_S[world]) if world in PUBLISHED_IMG_S else None,
        "vs_baseline_note": "published README wall times (unknown GPU, data loading and start-up included) turned into "
                            "images/s; the same-box comparison is the reference arm (--impl reference)",
        "dtype": ("fp32 parameters/accumulation; conv2 TF32 on tcgen05 (as the reference's cuDNN convs); fc1 "
                  + ("3xTF32 split-operand (fp32-class, as the reference's fp32 SGEMM)" if fp32fc else "single-pass TF32")
                  + "; conv1, fc2, loss, optimizer fp32"),
        "data": "synthetic", "impl": "ours",
        "config": shared_config(world, B),
        "impl_config": {"comm": getattr(backend.ddp, "mode", "none"), "use_tc_mask": eng.use_tc,
                        "cuda_graph": not args.no_graph, "env": b200_env()},
        "steady_state": {"images_per_s": imgs / (steady_ms / 1e3), "ms_per_step": steady_ms / K,
                         "note": "same K steps back to back, no L2 flush"},
        "e2e": {"value": imgs / (e2e_ms / 1e3), "unit": "images/s", "ms_per_step": e2e_ms / K,
                "h2d_bytes_per_step": B * 784 + B, "d2h_bytes_per_step": 4,
                "api": "B200Backend.step_gather(host uint8 dataset, labels, this step's sampler indices): the native "
                       "library gathers the batch into pinned memory (784-byte rows), copies it to the GPU on a copy "
                       "stream, trains; the head kernel stores the step's mean loss into mapped pinned host memory",
                "losses_finite": losses_ok},
        "gpu_launches": int(launches),
        "launches_per_step": eng.launches_per_step(B),
        "host_wall_s_timed_region": t_wall,
        "clocks": clocks,
        "checks": {"golden_losses": golden, "param_update_norm": moved, "replica_drift": drift,
                   "ok": bool(losses_ok and moved > 0 and drift == 0.0 and golden.get("status") != "MISMATCH")},
        "last_loss": float(host_loss[-1]),
    }
    if job is not None:
        out["epoch20_measured_s"] = job["device_s"] * (20.0 / job["epochs"])
        out["epoch20_images_per_s"] = job["images_per_s"]
        out["epoch20_job"] = job
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out if rank == 0 else {}


def _self_launch_cmd(gpus: int, argv) -> list:
    port = 29500 + os.getpid() % 2000
    return [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={gpus}", "--master-addr",
            "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__), *argv]


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--epochs", type=int, default=20, help="epochs of the whole-job measurement (0 = skip it)")
    ap.add_argument("--comm", default="auto")
    ap.add_argument("--no-graph", action="store_true")
    args = ap.parse_args()
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        # started by hand without a launcher: become the N-rank job the flag asks for (one process per GPU)
        raise SystemExit(subprocess.call(_self_launch_cmd(args.gpus, sys.argv[1:])))
    if args.impl == "reference":
        from baseline.reference_arm import run_reference
        out = run_reference(args)
    else:
        out = run_ours(args)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
