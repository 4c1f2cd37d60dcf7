"""pytorch_mnist_ddp_b200 — a Blackwell-native data-parallel CNN training framework.

Capabilities and CLI of FlyingAnt2018/pytorch_mnist_ddp (mnist.py / mnist_ddp.py), rebuilt
B200-first: one process per GPU, an HBM-resident dataset, hand-written sm_100a kernels for
every hot op (csrc/), and a fused peer-to-peer all-reduce + Adadelta kernel over NVLink in
place of DistributedDataParallel's NCCL path.

Layout
------
models/    Net (the reference's architecture, mnist_ddp.py:39-62) and parameter layouts
ops/       thin Python wrappers over the C-ABI kernels in libb200mnist.so (ctypes)
parallel/  process-group bootstrap, sampler, symmetric memory, fused data-parallel reducer
utils/     dataset (idx files / synthetic), logging formats, checkpoint, timers, clocks
engine.py  the native training engine (one CUDA graph per step)
trainer.py epoch loop shared by the CPU (plain torch) and B200 (native) backends
cli.py     argparse front-ends equivalent to mnist.py and mnist_ddp.py
"""
__version__ = "0.1.0"

from .models.net import Net  # noqa: F401
