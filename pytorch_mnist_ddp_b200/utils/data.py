"""MNIST-shaped data: idx files on disk, or a deterministic synthetic stand-in.

The reference loads ``torchvision.datasets.MNIST('./data', download=True)`` and applies
``ToTensor`` + ``Normalize((0.1307,), (0.3081,))`` per sample in a worker process
(mnist_ddp.py:153-160).  Here the raw uint8 images stay in one tensor (on the host for the CPU
backend, resident in HBM for the B200 backend) and the normalisation is applied to a whole
batch at once — on the GPU it is fused into the conv1 kernel (csrc/conv1.cu).

There is no network in the build/bench environment, so when the four uncompressed idx files
are not found under ``<root>/MNIST/raw`` a synthetic dataset of the same shape (60k/10k
28x28 uint8 images, uniform-random pixels and labels) is generated from a fixed seed
(BASELINE.md §2).  `write_idx_files` writes the same bytes in the idx format torchvision
parses, so the unmodified reference can be run on identical data.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass
from typing import Tuple

import numpy as np
import torch

MEAN, STD = 0.1307, 0.3081
TRAIN_N, TEST_N = 60000, 10000
_FILES = {
    (True, "images"): "train-images-idx3-ubyte",
    (True, "labels"): "train-labels-idx1-ubyte",
    (False, "images"): "t10k-images-idx3-ubyte",
    (False, "labels"): "t10k-labels-idx1-ubyte",
}


@dataclass
class MnistSplit:
    images: torch.Tensor  # uint8 [N, 28, 28]
    labels: torch.Tensor  # uint8 [N]
    synthetic: bool

    def __len__(self) -> int:
        return int(self.images.shape[0])


def synthetic_split(train: bool, n: int | None = None, seed: int = 20181) -> MnistSplit:
    """Uniform-random pixels and labels; the stream depends only on (seed, train)."""
    n = (TRAIN_N if train else TEST_N) if n is None else n
    rng = np.random.Generator(np.random.PCG64(seed + (0 if train else 1)))
    images = rng.integers(0, 256, size=(n, 28, 28), dtype=np.uint8)
    labels = rng.integers(0, 10, size=(n,), dtype=np.uint8)
    return MnistSplit(torch.from_numpy(images), torch.from_numpy(labels), True)


def _read_idx(path: str) -> np.ndarray:
    """An idx file, plain or gzip-compressed (the form MNIST is distributed in: *-ubyte.gz)."""
    if path.endswith(".gz"):
        import gzip
        with gzip.open(path, "rb") as f:
            data = f.read()
    else:
        with open(path, "rb") as f:
            data = f.read()
    magic = struct.unpack(">I", data[:4])[0]
    ndim = magic & 0xFF
    if (magic >> 8) != 0x08:
        raise ValueError(f"{path}: not an unsigned-byte idx file (magic {magic:#x})")
    dims = struct.unpack(">" + "I" * ndim, data[4: 4 + 4 * ndim])
    arr = np.frombuffer(data, dtype=np.uint8, offset=4 + 4 * ndim)
    return arr.reshape(dims).copy()


def _write_idx(path: str, arr: np.ndarray) -> None:
    arr = np.ascontiguousarray(arr, dtype=np.uint8)
    magic = 0x0800 | arr.ndim
    with open(path, "wb") as f:
        f.write(struct.pack(">I", magic))
        f.write(struct.pack(">" + "I" * arr.ndim, *arr.shape))
        f.write(arr.tobytes())


def raw_dir(root: str) -> str:
    return os.path.join(root, "MNIST", "raw")


def _find(root: str, name: str):
    """The idx file `name` under <root>/MNIST/raw: uncompressed (what torchvision leaves after download), else .gz."""
    for cand in (os.path.join(raw_dir(root), name), os.path.join(raw_dir(root), name + ".gz")):
        if os.path.isfile(cand):
            return cand
    return None


def idx_files_present(root: str) -> bool:
    return all(_find(root, f) is not None for f in _FILES.values())


def write_idx_files(root: str, train: MnistSplit, test: MnistSplit) -> None:
    os.makedirs(raw_dir(root), exist_ok=True)
    for is_train, split in ((True, train), (False, test)):
        _write_idx(os.path.join(raw_dir(root), _FILES[(is_train, "images")]), split.images.numpy())
        _write_idx(os.path.join(raw_dir(root), _FILES[(is_train, "labels")]), split.labels.numpy())


def load_split(root: str, train: bool, announce: bool = False, require_real: bool = False) -> MnistSplit:
    """idx files if present (same location torchvision uses; .gz accepted), else the synthetic split.

    The reference downloads MNIST (``download=True``, mnist_ddp.py:157); there is no network here, so a missing
    dataset falls back to synthetic data -- loudly on rank 0, or not at all with ``require_real`` (--require-real-data)."""
    if idx_files_present(root):
        images = _read_idx(_find(root, _FILES[(train, "images")]))
        labels = _read_idx(_find(root, _FILES[(train, "labels")]))
        if images.shape[0] != labels.shape[0] or images.shape[1:] != (28, 28):
            raise ValueError(f"{raw_dir(root)}: idx files disagree (images {images.shape}, labels {labels.shape})")
        return MnistSplit(torch.from_numpy(images), torch.from_numpy(labels), False)
    if require_real:
        raise FileNotFoundError(
            f"--require-real-data: no MNIST idx files ({', '.join(_FILES.values())}, optionally .gz) under "
            f"{raw_dir(root)}; this build cannot download them (no network)")
    if announce:
        print(f"[data] no idx files under {raw_dir(root)} and no network: using synthetic "
              f"MNIST-shaped data (random pixels/labels)", flush=True)
    return synthetic_split(train)


def normalize_batch(images_u8: torch.Tensor) -> torch.Tensor:
    """uint8 [B,28,28] -> float32 [B,1,28,28]; same op order as ToTensor+Normalize, so the
    result is bit-identical to the reference's per-sample transform (mnist_ddp.py:153-156)."""
    x = images_u8.to(torch.float32).div(255)
    return x.sub_(MEAN).div_(STD).unsqueeze(1)


def split_sizes(n: int, world: int, batch: int) -> Tuple[int, int]:
    """(samples per rank, steps per epoch per rank) under DistributedSampler + drop_last=False."""
    per_rank = -(-n // world)
    return per_rank, -(-per_rank // batch)
