#!/usr/bin/env python
"""Single-process MNIST training — same CLI as the reference's mnist.py (mnist.py:73-133)."""
from pytorch_mnist_ddp_b200.cli import main_single

if __name__ == "__main__":
    main_single()
