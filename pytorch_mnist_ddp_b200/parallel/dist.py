"""Process topology discovery and process-group bootstrap.

Mirrors ``init_distributed_mode`` (mnist_ddp.py:13-37): torchrun env contract first
(RANK / WORLD_SIZE / LOCAL_RANK), then SLURM (SLURM_PROCID), then a pre-set ``args.rank``,
else "Not using distributed mode".  Differences, all deliberate:

* the backend is ``nccl`` when CUDA is used and ``gloo`` otherwise, so the ``--no-cuda`` path
  also works multi-process (the reference hard-codes nccl, mnist_ddp.py:33);
* ``torch.cuda.set_device`` is only called when CUDA is used;
* the process group is used for bootstrap only (exchange of IPC handles, barriers at start-up,
  the initial parameter broadcast).  Per-step gradient traffic goes through the fused
  peer-to-peer kernel in ``parallel/fused_ddp.py``, not through this group.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch
import torch.distributed as dist

from ..utils import logfmt


@dataclass
class DistInfo:
    distributed: bool = False
    rank: int = 0
    world_size: int = 1
    gpu: int = 0
    backend: str = ""

    @property
    def is_main(self) -> bool:
        return self.rank == 0


def init_distributed_mode(args, use_cuda: bool) -> DistInfo:
    env = os.environ
    if "RANK" in env and "WORLD_SIZE" in env:
        args.rank = int(env["RANK"])
        args.world_size = int(env["WORLD_SIZE"])
        args.gpu = int(env.get("LOCAL_RANK", 0))
    elif "SLURM_PROCID" in env:
        args.rank = int(env["SLURM_PROCID"])
        ndev = torch.cuda.device_count() if use_cuda else 1
        args.gpu = args.rank % max(ndev, 1)
    elif getattr(args, "rank", None) is not None:
        args.gpu = getattr(args, "gpu", 0)
    else:
        print(logfmt.NOT_DISTRIBUTED)
        args.distributed = False
        return DistInfo(False, 0, 1, 0, "")

    args.distributed = True
    if use_cuda:
        torch.cuda.set_device(args.gpu)
    args.dist_backend = "nccl" if use_cuda else "gloo"
    print(logfmt.banner(args.rank, args.dist_url, args.gpu, args.world_size), flush=True)
    kwargs = {}
    if use_cuda:
        kwargs["device_id"] = torch.device("cuda", args.gpu)
    dist.init_process_group(backend=args.dist_backend, init_method=args.dist_url,
                            world_size=args.world_size, rank=args.rank, **kwargs)
    return DistInfo(True, args.rank, args.world_size, args.gpu, args.dist_backend)


def shutdown() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
