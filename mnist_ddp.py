#!/usr/bin/env python
"""Data-parallel MNIST training — same CLI and launch contract as the reference's
mnist_ddp.py (mnist_ddp.py:108-203): `torchrun --nproc_per_node=N mnist_ddp.py ...`,
`python -m torch.distributed.launch ...` (both --local_rank spellings), SLURM, or plain
`python mnist_ddp.py` for one process."""
from pytorch_mnist_ddp_b200.cli import main_ddp

if __name__ == "__main__":
    main_ddp()
