"""Reference arm of bench.py: the UNMODIFIED reference scripts, driven through their own API.

``baseline/_ref`` holds a pip-installed copy of /root/reference (mnist.py, mnist_ddp.py are
byte-identical; only a packaging shim was added in a /tmp copy because the upstream repo has no
setup.py/pyproject — see DESIGN.md "Reference install").  Nothing from pytorch_mnist_ddp_b200's
models, kernels or engine is on this path: the model is ``mnist_ddp.Net``, the step is
``mnist_ddp.train()`` (mnist_ddp.py:65-86), distribution is ``mnist_ddp.init_distributed_mode``
+ ``DistributedDataParallel`` over NCCL, the optimizer ``optim.Adadelta`` — constructed exactly
as ``mnist_ddp.main()`` does (mnist_ddp.py:142-178).  Only two harness pieces come from this
repo: the synthetic idx files (there is no network to download MNIST) and the timing wrapper
around the loader iterator.

Two numbers, same rules as the ours arm:
  value : train() fed with batches that are already on the GPU (the "R-resident" baseline of
          BASELINE.md: data loading removed, NCCL DDP + cuDNN/cuBLAS + foreach Adadelta kept);
          per-step CUDA events, L2 flushed between steps.
  e2e   : train() fed by the reference's own DataLoader configuration (torchvision MNIST,
          ToTensor+Normalize per sample, num_workers=1, pin_memory=True, DistributedSampler),
          --log-interval 1 so the loss is read back every step.
"""
from __future__ import annotations

import argparse
import contextlib
import io
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
PUBLISHED_IMG_S = {1: 20 * 60000 / 242.3, 2: 20 * 60000 / 137.1, 4: 20 * 60000 / 73.6}


def _unavailable(why: str) -> dict:
    return {"impl": "reference", "unavailable": why}


class _TimedLoader:
    """Iterable handed to the reference's train(): yields W warm-up + K timed batches and brackets
    every timed step with its own CUDA event pair (optionally flushing L2 in between)."""

    def __init__(self, source, dataset, n_batches_per_epoch, warmup, steps, flush, sync):
        self.source, self.dataset = source, dataset
        self.n, self.warmup, self.steps, self.flush, self.sync = n_batches_per_epoch, warmup, steps, flush, sync
        import torch
        self.starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        self.ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        self.region = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

    def __len__(self):
        return self.n

    def __iter__(self):
        it = iter(self.source)
        for i in range(self.warmup + self.steps):
            try:
                batch = next(it)
            except StopIteration:
                it = iter(self.source)
                batch = next(it)
            k = i - self.warmup
            if k == 0:
                self.sync()
                self.region[0].record()
            if k > 0:
                self.ends[k - 1].record()
            if k >= 0:
                if self.flush is not None:
                    self.flush.zero_()
                self.starts[k].record()
            yield batch
        self.ends[self.steps - 1].record()
        self.region[1].record()


def run_reference(cli) -> dict:
    if not os.path.isfile(os.path.join(REF_DIR, "mnist_ddp.py")):
        return _unavailable("baseline/_ref is missing (pip install of the reference copy was not run)")
    try:
        import torch
        import torch.distributed as dist
        import torch.optim as optim
        from torchvision import datasets, transforms
    except Exception as e:  # pragma: no cover
        return _unavailable(f"torch/torchvision import failed: {e}")
    if not torch.cuda.is_available():
        return _unavailable("no CUDA device: the reference's DDP path hard-codes nccl (mnist_ddp.py:33)")
    sys.path.insert(0, REF_DIR)
    import mnist_ddp as ref  # the unmodified reference module

    sys.path.insert(0, os.path.dirname(HERE))
    from pytorch_mnist_ddp_b200.utils.data import synthetic_split, write_idx_files, idx_files_present
    from bench import ClockSampler, METRIC, shared_config

    K, W, B = cli.steps, max(cli.warmup, 3), cli.batch
    # ---- what mnist_ddp.main() does up to the epoch loop (mnist_ddp.py:110-178)
    args = argparse.Namespace(batch_size=B, test_batch_size=1000, epochs=1, lr=1.0, gamma=0.7, no_cuda=False,
                              dry_run=False, seed=1, log_interval=10 ** 9, save_model=False, local_rank=None,
                              world_size=1, dist_url="env://")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    with contextlib.redirect_stdout(io.StringIO()):
        ref.init_distributed_mode(args)
    rank = dist.get_rank() if args.distributed else 0
    world = dist.get_world_size() if args.distributed else 1
    if world != cli.gpus:
        raise SystemExit(f"--gpus {cli.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    torch.manual_seed(args.seed)
    device = torch.device("cuda")

    data_root = os.environ.get("B200_REF_DATA", "/tmp/b200_ref_data")
    if rank == 0 and not idx_files_present(data_root):
        write_idx_files(data_root, synthetic_split(True), synthetic_split(False))
    if args.distributed:
        dist.barrier()
    transform = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))])
    dataset_train = datasets.MNIST(data_root, train=True, download=True, transform=transform)
    if args.distributed:
        train_sampler = torch.utils.data.distributed.DistributedSampler(dataset_train, shuffle=True)
    else:
        train_sampler = torch.utils.data.RandomSampler(dataset_train)
    train_loader = torch.utils.data.DataLoader(dataset_train, sampler=train_sampler, batch_size=B,
                                               num_workers=1, pin_memory=True)
    model = ref.Net().to(device)
    if args.distributed:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[args.gpu])
    optimizer = optim.Adadelta(model.parameters(), lr=args.lr)

    def sync():
        if args.distributed:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- value: GPU-resident batches through ref.train()
    g = torch.Generator().manual_seed(1 + rank)
    n_res = 32
    resident = [(torch.randn(B, 1, 28, 28, generator=g).to(device),
                 torch.randint(0, 10, (B,), generator=g).to(device)) for _ in range(n_res)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
    sampler = ClockSampler(local)
    sampler.start()
    timed = _TimedLoader(resident, dataset_train, len(train_loader), W, K, flush, sync)
    with contextlib.redirect_stdout(io.StringIO()):
        ref.train(args, model, device, timed, optimizer, 1)
    sync()
    total_ms = reduce_max(sum(s.elapsed_time(e) for s, e in zip(timed.starts, timed.ends)))
    timed = _TimedLoader(resident, dataset_train, len(train_loader), W, K, None, sync)
    with contextlib.redirect_stdout(io.StringIO()):
        ref.train(args, model, device, timed, optimizer, 1)
    sync()
    steady_ms = reduce_max(timed.region[0].elapsed_time(timed.region[1]))

    # ---- e2e: the reference's own input pipeline, loss read back every step
    args.log_interval = 1
    if args.distributed:
        train_sampler.set_epoch(1)
    k_e2e = min(K, max(10, len(train_loader) - W - 1))
    timed = _TimedLoader(train_loader, dataset_train, len(train_loader), W, k_e2e, None, sync)
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        ref.train(args, model, device, timed, optimizer, 1)
    sync()
    e2e_ms = reduce_max(timed.region[0].elapsed_time(timed.region[1]))
    clocks = sampler.stop()
    del train_loader

    # ---- the reference's protocol for real (mnist_ddp.py:179-189): E epochs of train() + rank-0 test() + StepLR.
    # "R-resident": the dataset sits normalised in HBM and every batch is gathered there with the indices of the
    # reference's own DistributedSampler, so the number is the reference's NCCL DDP + cuDNN/cuBLAS + foreach-Adadelta
    # stack without its Python data loader (that one is in `e2e`).
    job = None
    if getattr(cli, "epochs", 0) > 0:
        from torch.optim.lr_scheduler import StepLR
        E = cli.epochs
        tr, te = synthetic_split(True), synthetic_split(False)
        norm = lambda u8: ((u8.to(device).float() / 255) - 0.1307) / 0.3081
        xtr, ytr = norm(tr.images).unsqueeze(1), tr.labels.to(device).long()
        xte, yte = norm(te.images).unsqueeze(1), te.labels.to(device).long()
        args.log_interval, args.lr = 10, 1.0
        for g_ in optimizer.param_groups:
            g_["lr"] = 1.0
        scheduler = StepLR(optimizer, step_size=1, gamma=args.gamma)
        job_sampler = (torch.utils.data.distributed.DistributedSampler(dataset_train, shuffle=True) if args.distributed
                       else torch.utils.data.RandomSampler(dataset_train))
        class _TestBatches(list):  # ref.test() reads len(test_loader.dataset) (mnist_ddp.py:101)
            dataset = te
        test_batches = _TestBatches((xte[i: i + 1000], yte[i: i + 1000]) for i in range(0, len(te), 1000))

        class _Resident:
            def __init__(self, idx):
                self.idx, self.dataset = idx, dataset_train
            def __len__(self):
                return -(-len(self.idx) // B)
            def __iter__(self):
                for o in range(0, len(self.idx), B):
                    sel = self.idx[o: o + B]
                    yield xtr[sel], ytr[sel]

        sync()
        j0, j1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        j0.record()
        tw = time.perf_counter()
        steps = 0
        out_buf = io.StringIO()
        with contextlib.redirect_stdout(out_buf):
            for epoch in range(1, E + 1):
                if args.distributed:
                    job_sampler.set_epoch(epoch)
                idx_e = torch.tensor(list(iter(job_sampler)), device=device)
                loader_e = _Resident(idx_e)
                ref.train(args, model, device, loader_e, optimizer, epoch)
                steps += len(loader_e)
                if rank == 0:
                    ref.test(model.module if args.distributed else model, device, test_batches)
                scheduler.step()
        j1.record()
        torch.cuda.synchronize()
        wall = reduce_max(time.perf_counter() - tw)
        sync()
        dev_s = reduce_max(j0.elapsed_time(j1)) / 1e3
        tests = [l for l in out_buf.getvalue().splitlines() if l.startswith("Test set:")]
        job = {"epochs": E, "steps_per_epoch_per_rank": len(loader_e), "optimizer_steps": steps, "device_s": dev_s,
               "wall_s": wall, "images_per_s": E * 60000 / dev_s,
               "includes": "mnist_ddp.train() on HBM-resident batches gathered with the reference's DistributedSampler "
                           "order, rank-0 mnist_ddp.test() of 10000 images per epoch, StepLR, loss.item() every 10 steps",
               "last_test_line": tests[-1] if tests else None}

    imgs = world * B * K
    value = imgs / (total_ms / 1e3)
    out = {
        "metric": METRIC,
        "value": value, "unit": "images/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": (value / PUBLISHED_IMG_S[world]) if world in PUBLISHED_IMG_S else None,
        "vs_baseline_note": "published README wall times (unknown GPU, data loading and start-up included) turned into "
                            "images/s",
        "dtype": "fp32 (cuDNN TF32 conv, fp32 matmul: torch defaults, as the reference trains)",
        "data": "synthetic", "impl": "reference",
        "config": shared_config(world, B),
        "impl_config": {"model_class": "mnist_ddp.Net (unmodified reference)",
                        "parallelism": "torch DistributedDataParallel + NCCL" if world > 1 else "single process",
                        "value_path": "mnist_ddp.train() on GPU-resident batches (R-resident, BASELINE.md)"},
        "steady_state": {"images_per_s": imgs / (steady_ms / 1e3), "ms_per_step": steady_ms / K},
        "e2e": {"value": world * B * k_e2e / (e2e_ms / 1e3), "unit": "images/s", "ms_per_step": e2e_ms / k_e2e,
                "steps": k_e2e, "h2d_bytes_per_step": B * 784 * 4 + B * 8, "d2h_bytes_per_step": 4,
                "api": "mnist_ddp.train() over the reference DataLoader (torchvision MNIST, num_workers=1, "
                       "pin_memory) with --log-interval 1"},
        "gpu_launches": 0,
        "clocks": clocks,
    }
    if job is not None:
        out["epoch20_measured_s"] = job["device_s"] * (20.0 / job["epochs"])
        out["epoch20_images_per_s"] = job["images_per_s"]
        out["epoch20_job"] = job
    if args.distributed:
        dist.barrier()
        dist.destroy_process_group()
    return out if rank == 0 else {}
