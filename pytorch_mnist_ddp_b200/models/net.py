"""The MNIST CNN and the flat parameter layout used by the native engine.

`Net` has the architecture of the reference model (mnist_ddp.py:39-62 == mnist.py:11-34):
conv 1->32 3x3, ReLU, conv 32->64 3x3, ReLU, 2x2 max-pool, dropout .25, flatten,
fc 9216->128, ReLU, dropout .5, fc 128->10, log_softmax.  It is a plain ``nn.Module`` so that
(a) the default PyTorch initialisation (and therefore the seed-1 weights of SURVEY A.1) is
inherited, (b) ``state_dict()`` has exactly the reference's keys/shapes (mnist_ddp.py:195), and
(c) the ``--no-cuda`` path and every numerics test have an fp32 oracle.

The native engine does not run this module.  It keeps all parameters in ONE flat fp32 buffer
whose layout (`ParamLayout`) is chosen for the sm_100a kernels:

* activations are NHWC, so ``fc1.weight``'s 9216 axis is stored in (h, w, c) order instead of
  the reference's NCHW-flatten order (c, h, w)  (mnist_ddp.py:56);
* ``conv2.weight`` is stored tap-major ``[kh*3+kw][out][in]`` so every tap is a K-major
  64x32 (128-byte rows) tcgen05 B operand;
* ``conv1.weight`` is stored ``[tap][out]``;
* the buffer is split in the two gradient buckets that DistributedDataParallel builds for this
  model (bucket 0 = fc2.bias, fc2.weight, fc1.bias, fc1.weight; bucket 1 = conv2.*, conv1.* —
  SURVEY §2.B N17), because the fused all-reduce + Adadelta kernel is launched per bucket.

`ParamLayout.pack/unpack` convert between the two representations, so checkpoints stay
interchangeable with the reference.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

# name -> shape, in state_dict order
PARAM_SPECS: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict(
    [
        ("conv1.weight", (32, 1, 3, 3)),
        ("conv1.bias", (32,)),
        ("conv2.weight", (64, 32, 3, 3)),
        ("conv2.bias", (64,)),
        ("fc1.weight", (128, 9216)),
        ("fc1.bias", (128,)),
        ("fc2.weight", (10, 128)),
        ("fc2.bias", (10,)),
    ]
)
NUM_PARAMS = 1_199_882


class Net(nn.Module):
    """Reference-architecture CNN (mnist_ddp.py:39-62); x[N,1,28,28] -> log-probs [N,10]."""

    def __init__(self) -> None:
        super().__init__()
        # construction order matters: it fixes the RNG stream of the default init
        self.conv1 = nn.Conv2d(1, 32, 3, 1)
        self.conv2 = nn.Conv2d(32, 64, 3, 1)
        self.dropout1 = nn.Dropout(0.25)
        self.dropout2 = nn.Dropout(0.5)
        self.fc1 = nn.Linear(9216, 128)
        self.fc2 = nn.Linear(128, 10)

    def features(self, x: torch.Tensor) -> torch.Tensor:
        x = F.relu(self.conv1(x))
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        return torch.flatten(self.dropout1(x), 1)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.features(x)
        x = self.dropout2(F.relu(self.fc1(x)))
        return F.log_softmax(self.fc2(x), dim=1)


def _round_up(v: int, m: int) -> int:
    return (v + m - 1) // m * m


@dataclass(frozen=True)
class Segment:
    name: str
    offset: int  # in floats, from the start of the flat buffer
    numel: int
    bucket: int


class ParamLayout:
    """Flat fp32 layout shared by parameters, gradients and both Adadelta state tensors.

    Every segment starts on a 256-float (1 KiB) boundary so TMA descriptors and 128-bit
    vector accesses are always aligned; padding floats are zero and stay zero under Adadelta
    (g = 0 -> delta = 0).
    """

    ALIGN = 256
    # bucket membership and in-bucket order follow DDP's rebuilt buckets
    BUCKETS: List[List[str]] = [
        ["fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"],
        ["conv2.weight", "conv2.bias", "conv1.weight", "conv1.bias"],
    ]

    def __init__(self) -> None:
        self.segments: Dict[str, Segment] = OrderedDict()
        self.bucket_ranges: List[Tuple[int, int]] = []
        off = 0
        for b, names in enumerate(self.BUCKETS):
            start = off
            for n in names:
                numel = 1
                for d in PARAM_SPECS[n]:
                    numel *= d
                self.segments[n] = Segment(n, off, numel, b)
                off = _round_up(off + numel, self.ALIGN)
            self.bucket_ranges.append((start, off))
        self.total = off

    # ---- offsets used by the native engine (order must match b200_layout in engine.cu)
    def offsets(self) -> List[int]:
        order = ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
                 "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"]
        return [self.segments[n].offset for n in order]

    def real_bucket_bytes(self) -> List[int]:
        return [4 * sum(self.segments[n].numel for n in names) for names in self.BUCKETS]

    # ---- torch layout <-> internal layout, per tensor
    @staticmethod
    def to_internal(name: str, t: torch.Tensor) -> torch.Tensor:
        t = t.detach().to(torch.float32)
        if name == "fc1.weight":  # [128, c*144+hw] -> [128, hw*64+c]
            return t.reshape(128, 64, 144).permute(0, 2, 1).reshape(-1)
        if name == "conv2.weight":  # [o,i,kh,kw] -> [tap][o][i]
            return t.permute(2, 3, 0, 1).reshape(-1)
        if name == "conv1.weight":  # [o,1,kh,kw] -> [tap][o]
            return t.reshape(32, 9).t().reshape(-1)
        return t.reshape(-1)

    @staticmethod
    def from_internal(name: str, flat: torch.Tensor) -> torch.Tensor:
        if name == "fc1.weight":
            return flat.reshape(128, 144, 64).permute(0, 2, 1).reshape(128, 9216).contiguous()
        if name == "conv2.weight":
            return flat.reshape(3, 3, 64, 32).permute(2, 3, 0, 1).contiguous()
        if name == "conv1.weight":
            return flat.reshape(9, 32).t().reshape(32, 1, 3, 3).contiguous()
        return flat.reshape(PARAM_SPECS[name]).clone()

    def pack(self, state: Dict[str, torch.Tensor], out: torch.Tensor | None = None) -> torch.Tensor:
        """state_dict (reference layout; an optional ``module.`` prefix is accepted) -> flat."""
        state = {k[len("module."):] if k.startswith("module.") else k: v for k, v in state.items()}
        flat = torch.zeros(self.total, dtype=torch.float32) if out is None else out
        for n, seg in self.segments.items():
            src = self.to_internal(n, state[n].cpu())
            flat[seg.offset: seg.offset + seg.numel].copy_(src)
        return flat

    def unpack(self, flat: torch.Tensor, prefix: str = "") -> "OrderedDict[str, torch.Tensor]":
        """flat -> state_dict with the reference's key order, shapes and dtype."""
        flat = flat.detach().cpu()
        out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        for n in PARAM_SPECS:
            seg = self.segments[n]
            out[prefix + n] = self.from_internal(n, flat[seg.offset: seg.offset + seg.numel])
        return out
