"""Epoch driver shared by both backends, plus the plain-PyTorch backend.

`fit()` reproduces the control flow of the reference's ``main()`` after argument parsing
(mnist_ddp.py:140-197): seed, data, model, optimizer, StepLR, then per epoch
``set_epoch -> train -> (rank 0) test -> scheduler.step`` and an optional final save.  The
three printed line formats and their cadence are those of ``train()``/``test()``
(mnist_ddp.py:65-105).

Backends
--------
`TorchBackend`  plain PyTorch ops on the CPU (``--no-cuda``, the no-GPU CI path, BASELINE
                config #1).  Same op order as the reference so, for the same seed and data, the
                loss trajectory is bit-identical to ``mnist.py --no-cuda``.  Multi-process runs
                all-reduce two flat gradient buckets over gloo.
`B200Backend`   (engine.py) the native sm_100a engine; one CUDA graph per step.

A backend exposes: ``begin_epoch(indices)``, ``train_step(i, batch) -> token``,
``loss(token) -> float``, ``evaluate(split, batch) -> (nll_sum, correct)``, ``set_lr(lr)``,
``state_dict()``.
"""
from __future__ import annotations

import time
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist
import torch.nn.functional as F

from .models.net import Net, ParamLayout
from .parallel import sampler as samplers
from .parallel.dist import DistInfo
from .utils import logfmt
from .utils.data import MnistSplit, normalize_batch


class TorchBackend:
    """Reference semantics with stock PyTorch ops; the oracle for everything else."""

    name = "torch"

    def __init__(self, train: MnistSplit, test: MnistSplit, info: DistInfo, lr: float,
                 device: torch.device | str = "cpu") -> None:
        self.device = torch.device(device)
        self.info = info
        self.train_split, self.test_split = train, test
        self.model = Net().to(self.device)  # built on the CPU first: seed-determined init
        if info.distributed:
            self._broadcast_params()
        self.opt = torch.optim.Adadelta(self.model.parameters(), lr=lr)
        self.layout = ParamLayout()
        self._idx: Optional[torch.Tensor] = None

    # -- distributed helpers (gloo on CPU): two flat buckets, sum then divide
    def _broadcast_params(self) -> None:
        for p in self.model.parameters():
            dist.broadcast(p.data, src=0)

    def _allreduce_grads(self) -> None:
        named = dict(self.model.named_parameters())
        w = self.info.world_size
        for names in ParamLayout.BUCKETS:
            flat = torch.cat([named[n].grad.reshape(-1) for n in names])
            dist.all_reduce(flat, op=dist.ReduceOp.SUM)
            flat.div_(w)
            off = 0
            for n in names:
                g = named[n].grad
                g.copy_(flat[off: off + g.numel()].view_as(g))
                off += g.numel()

    # -- backend interface
    def begin_epoch(self, indices: torch.Tensor) -> None:
        self._idx = indices.long()
        self.model.train()

    def train_step(self, step: int, batch: int, want_loss: bool = True):
        sel = self._idx[step * batch: (step + 1) * batch]
        data = normalize_batch(self.train_split.images[sel]).to(self.device)
        target = self.train_split.labels[sel].long().to(self.device)
        self.opt.zero_grad()
        loss = F.nll_loss(self.model(data), target)
        loss.backward()
        if self.info.distributed:
            self._allreduce_grads()
        self.opt.step()
        return loss

    def loss(self, token) -> float:
        return float(token.item())

    @torch.no_grad()
    def evaluate(self, split: MnistSplit, batch: int) -> Tuple[float, int]:
        self.model.eval()
        nll, correct = 0.0, 0
        for s in range(0, len(split), batch):
            data = normalize_batch(split.images[s: s + batch]).to(self.device)
            target = split.labels[s: s + batch].long().to(self.device)
            out = self.model(data)
            nll += F.nll_loss(out, target, reduction="sum").item()
            correct += out.argmax(dim=1).eq(target).sum().item()
        self.model.train()
        return nll, int(correct)

    def set_lr(self, lr: float) -> None:
        for g in self.opt.param_groups:
            g["lr"] = lr

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return {k: v.detach().cpu() for k, v in self.model.state_dict().items()}

    def load_state_dict(self, state: Dict[str, torch.Tensor]) -> None:
        state = {k[len("module."):] if k.startswith("module.") else k: v for k, v in state.items()}
        self.model.load_state_dict(state)

    def finish(self) -> None:
        pass


def fit(args, backend, train: MnistSplit, test: MnistSplit, info: DistInfo, *,
        sampler_mode: str, eval_all_ranks: bool = False) -> Dict[str, float]:
    """Run ``args.epochs`` epochs.  ``sampler_mode``: 'distributed' | 'random' | 'sequential'.

    Returns timing/metric facts for callers that want them (bench, tests).
    """
    n = len(train)
    lr = float(args.lr)
    world = info.world_size if info.distributed else 1
    stats: Dict[str, float] = {"steps": 0}
    t_train = 0.0
    last_eval: Tuple[float, int] = (float("nan"), 0)
    # NVTX ranges around every epoch, its training loop and its evaluation pass: they show up in
    # nsys / ncu timelines; a no-op without CUDA
    nvtx = getattr(backend, "name", "") == "b200" and torch.cuda.is_available()
    push = torch.cuda.nvtx.range_push if nvtx else (lambda name: None)
    pop = torch.cuda.nvtx.range_pop if nvtx else (lambda: None)
    for epoch in range(1, args.epochs + 1):
        push(f"epoch {epoch}")
        push("train")
        samplers.draw_loader_seed()  # iter(train_loader)
        if sampler_mode == "distributed":
            idx = samplers.distributed_indices(n, world, info.rank, epoch)  # set_epoch(epoch)
        elif sampler_mode == "random":
            idx = samplers.random_indices(n)
        else:
            idx = samplers.sequential_indices(n)
        per_rank = int(idx.numel())
        n_batches = -(-per_rank // args.batch_size)
        backend.begin_epoch(idx)
        t0 = time.perf_counter()
        pending: List[Tuple[int, int, object]] = []
        for batch_idx in range(n_batches):
            bsz = min(args.batch_size, per_rank - batch_idx * args.batch_size)
            log_now = info.is_main and batch_idx % args.log_interval == 0
            token = backend.train_step(batch_idx, args.batch_size, want_loss=log_now)
            stats["steps"] += 1
            if log_now:
                pending.append((batch_idx, bsz, token))
                # print one log point late so the device never waits for the host
                while len(pending) > 1:
                    _emit(args, epoch, pending.pop(0), world, n, n_batches, backend, info)
            if args.dry_run:
                break
        while pending:
            _emit(args, epoch, pending.pop(0), world, n, n_batches, backend, info)
        t_train += time.perf_counter() - t0
        pop()
        if info.is_main or eval_all_ranks:
            push("test")
            samplers.draw_loader_seed()  # iter(test_loader)
            nll, correct = backend.evaluate(test, args.test_batch_size)
            last_eval = (nll / len(test), correct)
            if info.is_main:
                print(logfmt.test_line(nll / len(test), correct, len(test)))
            pop()
        lr = lr * args.gamma  # StepLR(step_size=1), stepped after eval (mnist_ddp.py:189)
        backend.set_lr(lr)
        pop()
    backend.finish()
    stats.update(train_seconds=t_train, test_loss=last_eval[0], test_correct=last_eval[1])
    return stats


def _emit(args, epoch, item, world, n, n_batches, backend, info) -> None:
    batch_idx, bsz, token = item
    seen = (world if info.distributed else 1) * batch_idx * bsz
    print(logfmt.train_line(epoch, seen, n, batch_idx, n_batches, backend.loss(token)))
