#!/usr/bin/env python
"""In-kernel timeline of a persistent kernel (debug build: B200_EXTRA_NVCC_FLAGS=-DB200_STAMPS).

    B200_EXTRA_NVCC_FLAGS=-DB200_STAMPS python scripts/kernel_timeline.py fwd|bwd_dgrad [batch]

Runs a few training steps, then prints, for a few CTAs, the %globaltimer milestones of the last launch relative to
the earliest kernel entry of that launch (us)."""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_mnist_ddp_b200 import _native
from pytorch_mnist_ddp_b200.engine import NativeEngine
from pytorch_mnist_ddp_b200.models.net import Net
from pytorch_mnist_ddp_b200.utils.data import synthetic_split

which = sys.argv[1] if len(sys.argv) > 1 else "fwd"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 200
train, test = synthetic_split(True, 4096), synthetic_split(False, 256)
torch.manual_seed(1)
eng = NativeEngine(max(batch, 256), len(train), len(test), seed=1)
eng.load_data(0, train); eng.load_data(1, test)
eng.set_state_dict(Net().state_dict())
eng.begin_epoch(torch.randperm(4096).to(torch.int32))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for i in range(6):
    with eng.on_stream():
        flush.zero_()
    eng.train_step(batch)
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * (160 * 48))()
fn = getattr(_native.lib(), "b200_debug_stamps_" + which)
assert fn(buf) == 0
import numpy as np
a = np.frombuffer(buf, dtype=np.uint64).reshape(160, 48).astype(np.int64)
live = a[:, 0] > 0
t0 = a[live, 0].min()
rel = np.where(a > 0, (a - t0) / 1000.0, np.nan)
np.set_printoptions(precision=1, suppress=True, linewidth=250)
print("slots: columns = stamp slot; rows = CTA; us since the first CTA's entry")
for cta in (0, 1, 73, 103, 104, 147):
    if live[cta]:
        print(f"CTA {cta:3d}:", " ".join(f"{s}:{rel[cta, s]:.1f}" for s in range(48) if not np.isnan(rel[cta, s])))
print("per-slot mean over CTAs:", " ".join(f"{s}:{np.nanmean(rel[live, s]):.1f}" for s in range(48) if live.any() and not np.isnan(rel[live, s]).all()))
print("per-slot max  over CTAs:", " ".join(f"{s}:{np.nanmax(rel[live, s]):.1f}" for s in range(48) if live.any() and not np.isnan(rel[live, s]).all()))
