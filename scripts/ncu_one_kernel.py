"""Smallest program that runs the training step eagerly at batch 200 (for `ncu -k regex:<kernel> -s 2 -c 1`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_mnist_ddp_b200.engine import NativeEngine
from pytorch_mnist_ddp_b200.models.net import Net
from pytorch_mnist_ddp_b200.utils.data import synthetic_split

train, test = synthetic_split(True, 1024), synthetic_split(False, 256)
torch.manual_seed(1)
eng = NativeEngine(200, len(train), len(test), seed=1)
eng.load_data(0, train); eng.load_data(1, test)
eng.set_state_dict(Net().state_dict())
eng.begin_epoch(None)
for _ in range(4):
    eng.train_step(200, use_graph=False)
eng.synchronize()
print("ok", eng.read_loss(3))
