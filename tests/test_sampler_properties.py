"""Property tests (hypothesis) of the index streams against torch's own samplers: any dataset length, world size,
rank, epoch and seed -- including worlds larger than the dataset, where DistributedSampler's wrap-around padding
repeats the permutation more than once (LIB:torch/utils/data/distributed.py:110-140; mnist_ddp.py:162,181)."""
import torch
from hypothesis import given, settings, strategies as st
from torch.utils.data.distributed import DistributedSampler

from pytorch_mnist_ddp_b200.parallel.sampler import distributed_indices, sequential_indices
from pytorch_mnist_ddp_b200.utils.data import split_sizes


@settings(max_examples=150, deadline=None)
@given(n=st.integers(1, 700), world=st.integers(1, 16), epoch=st.integers(0, 50), seed=st.integers(0, 3),
       shuffle=st.booleans(), data=st.data())
def test_distributed_indices_match_torch_for_any_shape(n, world, epoch, seed, shuffle, data):
    rank = data.draw(st.integers(0, world - 1))
    s = DistributedSampler(range(n), num_replicas=world, rank=rank, shuffle=shuffle, seed=seed)
    s.set_epoch(epoch)
    got = distributed_indices(n, world, rank, epoch, seed=seed, shuffle=shuffle)
    assert got.dtype == torch.int32 and got.is_contiguous()
    assert got.tolist() == list(iter(s)) and got.numel() == len(s)


@settings(max_examples=60, deadline=None)
@given(n=st.integers(1, 500), world=st.integers(1, 9), epoch=st.integers(1, 20))
def test_ranks_partition_the_padded_permutation(n, world, epoch):
    """Every sample is seen at least once per epoch, shards have equal length, and the union is the permutation plus
    its wrapped-around head."""
    shards = [distributed_indices(n, world, r, epoch) for r in range(world)]
    per = -(-n // world)
    assert all(s.numel() == per for s in shards)
    merged = torch.stack(shards, 1).reshape(-1)  # rank r took perm[r::world]
    assert sorted(set(merged.tolist())) == list(range(n))
    g = torch.Generator()
    g.manual_seed(epoch)
    perm = torch.randperm(n, generator=g)
    assert merged[:n].tolist() == perm.tolist()
    pad = per * world - n
    assert merged[n:].tolist() == perm.repeat(-(-pad // n) if pad else 1)[:pad].tolist()


@settings(max_examples=60, deadline=None)
@given(n=st.integers(1, 100000), world=st.integers(1, 8), batch=st.integers(1, 5000))
def test_split_sizes_equal_loader_lengths(n, world, batch):
    """Per-rank sample count and steps per epoch as len(DistributedSampler) / len(DataLoader) report them
    (drop_last=False: the ragged last batch is a step of its own)."""
    per_rank, steps = split_sizes(n, world, batch)
    assert per_rank == -(-n // world) and steps == -(-per_rank // batch)
    assert sequential_indices(7).tolist() == list(range(7))
