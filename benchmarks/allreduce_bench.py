#!/usr/bin/env python
"""Gradient all-reduce microbenchmark (BASELINE.json config #5): the hand-written peer-to-peer kernels
(one-shot, two-shot) against the reference's path, ncclAllReduce (what DistributedDataParallel issues,
SURVEY N17), for 4 KB ... 256 MB plus the two real bucket sizes of this model.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 \
        benchmarks/allreduce_bench.py [--max-mb 256] [--iters 20]

Timing: CUDA events on the launching stream, 5 warm-up + `iters` timed calls per size, barrier +
synchronize around the region, MAX over ranks.  bus GB/s = 2 (W-1)/W * bytes / time; the NVLink 5
reference is 900 GB/s per direction per GPU nominal (770 GB/s measured peer copy, B200_PROFILING.md).
Also reports the fused all-reduce + 1/W + Adadelta kernel on the two buckets with its roofline:
max(NVLink bytes / 770 GB/s, HBM bytes / measured copy bandwidth).
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

BUCKETS = {"bucket1 (conv, 75,264 B)": 19200 * 4, "bucket0 (fc, 4,724,264 B)": 1181184 * 4}
NVLINK_GBS = 770.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=256)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    from pytorch_mnist_ddp_b200.ops import native_ops as ops
    from pytorch_mnist_ddp_b200.parallel.dist import DistInfo
    from pytorch_mnist_ddp_b200.parallel.fused_ddp import CommPlan, SymmetricArena

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", init_method="env://", world_size=world, rank=rank,
                            device_id=torch.device("cuda", local))
    info = DistInfo(True, rank, world, local, "nccl")
    hbm_gbs = 6545.0
    try:
        hbm_gbs = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass

    sizes = []
    b = 4096
    while b <= args.max_mb << 20:
        sizes.append((f"{b} B", b))
        b *= 4
    sizes += list(BUCKETS.items())
    nmax = max(s for _, s in sizes) // 4
    try:  # torch symmetric memory gives the NVLS multicast alias on top of the peer mapping
        arena = SymmetricArena.create_torch(info, 4 * nmax)  # [out/params | grads | v | u]
    except Exception as e:
        if rank == 0:
            print(f"# torch symmetric memory unavailable ({e!r}); falling back to the CUDA-IPC arena (no NVLS)")
        arena = SymmetricArena.create(info, 4 * nmax)
    has_mc = bool(getattr(arena, "has_multicast", False))
    view = arena.view()
    view[nmax:2 * nmax].normal_()
    ctrl = ops.make_ctrl(step=0, lr=1.0)
    nccl_buf = torch.randn(nmax, device="cuda")

    def timed(fn):
        for _ in range(5):
            fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.iters], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        return float(t.item()) * 1e-3  # seconds

    rows = []
    for name, nbytes in sizes:
        n = nbytes // 4
        plan = CommPlan.create(arena, nmax, 0, 2 * nmax, 3 * nmax, n, channel=2, ctrl=ctrl)
        t_nccl = timed(lambda: dist.all_reduce(nccl_buf[:n]))
        # large messages get more CTAs than the 64 the training path uses (that cap keeps the reducer
        # small enough to share SMs with the conv backward)
        nb = 0 if nbytes <= (4 << 20) else min(256, nbytes // (64 << 10))  # 0 = the library's own choice
        t_one = timed(lambda: plan.run(0, opt=False, nblocks=nb)) if nbytes * world <= (64 << 20) * 8 else None
        t_two = timed(lambda: plan.run(1, opt=False, nblocks=nb))
        t_nvls = timed(lambda: plan.run(2, opt=False, nblocks=nb)) if has_mc else None
        bus = lambda t: None if t is None else 2 * (world - 1) / world * nbytes / t / 1e9
        row = {"size": name, "bytes": nbytes, "nccl_us": t_nccl * 1e6, "oneshot_us": None if t_one is None else t_one * 1e6,
               "twoshot_us": t_two * 1e6, "nvls_us": None if t_nvls is None else t_nvls * 1e6,
               "nccl_busGBs": bus(t_nccl), "oneshot_busGBs": bus(t_one), "twoshot_busGBs": bus(t_two),
               "nvls_busGBs": bus(t_nvls)}
        if name in BUCKETS:  # fused all-reduce + 1/W + Adadelta, with its roofline
            for mode, key in ((0, "fused_oneshot"), (1, "fused_twoshot")) + (((2, "fused_nvls"),) if has_mc else ()):
                t = timed(lambda: plan.run(mode, opt=True))  # explicit tokens, lr from ctrl
                if mode == 0:
                    nvl, hbm = (world - 1) * nbytes, (world + 6) * nbytes      # read W grads; r/w p,v,u
                elif mode == 2:  # per-GPU link traffic: its slice in (reduced by the switch) + its slice out
                    nvl = 2 * nbytes / world
                    hbm = 8 / world * nbytes + nbytes
                else:
                    nvl = 2 * (world - 1) / world * nbytes                     # in: grads slices, out: params
                    hbm = (world + 6) / world * nbytes + (world - 1) / world * nbytes
                roof = max(nvl / (NVLINK_GBS * 1e9), hbm / (hbm_gbs * 1e9))
                row[key + "_us"] = t * 1e6
                row[key + "_roofline_us"] = roof * 1e6
                row[key + "_fraction_of_roofline"] = roof / t
        rows.append(row)
        plan.destroy()
        if rank == 0:
            print(json.dumps(row), flush=True)
    if rank == 0 and args.out:
        with open(args.out, "w") as f:
            json.dump({"world": world, "iters": args.iters, "nvlink_ref_GBs": NVLINK_GBS, "hbm_ref_GBs": hbm_gbs,
                       "rows": rows}, f, indent=1)
    dist.barrier()
    arena.destroy()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
