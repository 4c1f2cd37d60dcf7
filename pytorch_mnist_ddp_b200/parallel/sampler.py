"""Index streams equal to the samplers the reference uses.

* distributed: ``DistributedSampler(dataset, shuffle=True)`` + ``set_epoch(epoch)`` with
  1-based epochs (mnist_ddp.py:162,179-181): ``perm = randperm(n, seed + epoch)``, pad by
  wrapping to ``ceil(n/W)*W``, rank r takes ``perm[r::W]``.
* single process: ``RandomSampler`` (mnist_ddp.py:164): a fresh generator seeded from ONE
  int64 drawn from the global RNG, then ``randperm(n)``.  Drawing the seed the same way keeps
  the global RNG stream (and therefore CPU dropout masks) in step with the reference.
* sequential for the test split (mnist_ddp.py:165).

The whole epoch's indices are produced at once (int32) because the B200 backend uploads them
to the GPU once per epoch and gathers batches on the device.
"""
from __future__ import annotations

import math

import torch


def distributed_indices(n: int, world: int, rank: int, epoch: int, seed: int = 0,
                        shuffle: bool = True) -> torch.Tensor:
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        perm = torch.randperm(n, generator=g)
    else:
        perm = torch.arange(n)
    per_rank = math.ceil(n / world)
    total = per_rank * world
    pad = total - n
    if pad > 0:
        reps = math.ceil(pad / n)
        perm = torch.cat([perm, perm.repeat(reps)[:pad]])
    return perm[rank:total:world].to(torch.int32).contiguous()


def random_indices(n: int) -> torch.Tensor:
    seed = int(torch.empty((), dtype=torch.int64).random_().item())
    g = torch.Generator()
    g.manual_seed(seed)
    return torch.randperm(n, generator=g).to(torch.int32).contiguous()


def draw_loader_seed() -> int:
    """Every ``iter(DataLoader)`` draws one int64 'base seed' from the global RNG
    (LIB:torch/utils/data/dataloader.py:706) before the sampler draws anything.  The epoch
    driver calls this once per train epoch and once per eval pass so that the global CPU RNG
    stream stays aligned with the reference's."""
    return int(torch.empty((), dtype=torch.int64).random_().item())


def sequential_indices(n: int) -> torch.Tensor:
    return torch.arange(n, dtype=torch.int32)
