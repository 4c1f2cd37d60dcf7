"""Weights-only checkpoints in the reference's format.

The reference writes ``torch.save(model.state_dict(), "mnist_cnn.pt")`` once at the end
(mnist_ddp.py:191-197, mnist.py:132-133): an OrderedDict of 8 fp32 tensors, keys prefixed with
``module.`` when the DDP wrapper was saved.  `save` writes exactly that; `load` (which the
reference lacks — it has no resume path) accepts both key styles and validates
shapes, and always uses ``weights_only=True``.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict

import torch

from ..models.net import PARAM_SPECS


def save(state: Dict[str, torch.Tensor], path: str, prefix: str = "") -> None:
    out = OrderedDict()
    for name in PARAM_SPECS:
        out[prefix + name] = state[name].detach().to("cpu", torch.float32).contiguous()
    torch.save(out, path)


def load(path: str) -> "OrderedDict[str, torch.Tensor]":
    raw = torch.load(path, map_location="cpu", weights_only=True)
    out = OrderedDict()
    for name, shape in PARAM_SPECS.items():
        t = raw.get(name, raw.get("module." + name))
        if t is None:
            raise KeyError(f"{path}: missing parameter {name}")
        if tuple(t.shape) != shape:
            raise ValueError(f"{path}: {name} has shape {tuple(t.shape)}, expected {shape}")
        out[name] = t.to(torch.float32)
    return out
