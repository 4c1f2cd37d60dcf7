"""Command-line front ends equivalent to the reference's two scripts.

``main_single`` ~ mnist.py (mnist.py:73-133) and ``main_ddp`` ~ mnist_ddp.py
(mnist_ddp.py:108-203).  All 10 shared flags and the 3 DDP flags keep their names, types and
defaults; both spellings ``--local_rank`` / ``--local-rank`` are accepted because
current launchers send the latter.  Extra, clearly separated flags select the
backend and the data root; their defaults reproduce the reference behaviour.

Behavioural quirks kept on purpose: mnist.py trains un-shuffled on the CPU and
shuffled on CUDA; mnist_ddp.py always shuffles; same seed on every rank; 1-based ``set_epoch``;
the "ms" label on a value in seconds; checkpoint names ``mnist_cnn.pt`` / ``mnist_cnn_.pt``.
"""
from __future__ import annotations

import argparse
import time
from typing import List, Optional

import torch

from .parallel.dist import DistInfo, init_distributed_mode, shutdown
from .trainer import TorchBackend, fit
from .utils import checkpoint, logfmt
from .utils.data import load_split


def build_parser(ddp: bool) -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="PyTorch MNIST Example (B200-native engine)")
    p.add_argument("--batch-size", type=int, default=64, metavar="N",
                   help="input batch size for training, per rank (default: 64)")
    p.add_argument("--test-batch-size", type=int, default=1000, metavar="N",
                   help="input batch size for testing (default: 1000)")
    p.add_argument("--epochs", type=int, default=14, metavar="N",
                   help="number of epochs to train (default: 14)")
    p.add_argument("--lr", type=float, default=1.0, metavar="LR", help="learning rate (default: 1.0)")
    p.add_argument("--gamma", type=float, default=0.7, metavar="M",
                   help="Learning rate step gamma (default: 0.7)")
    p.add_argument("--no-cuda", action="store_true", default=False, help="disables CUDA training")
    p.add_argument("--dry-run", action="store_true", default=False, help="quickly check a single pass")
    p.add_argument("--seed", type=int, default=1, metavar="S", help="random seed (default: 1)")
    p.add_argument("--log-interval", type=int, default=10, metavar="N",
                   help="how many batches to wait before logging training status")
    p.add_argument("--save-model", action="store_true", default=False, help="For Saving the current Model")
    if ddp:
        p.add_argument("--local_rank", "--local-rank", dest="local_rank", type=int, default=None,
                       help="local rank, passed by the launcher (unused: LOCAL_RANK env wins)")
        p.add_argument("--world-size", default=1, type=int, help="number of distributed processes")
        p.add_argument("--dist-url", default="env://", type=str,
                       help="url used to set up distributed training")
    g = p.add_argument_group("b200 engine (not in the reference)")
    g.add_argument("--backend", choices=["auto", "b200", "torch"], default="auto",
                   help="auto: native sm_100a engine on CUDA, plain torch ops on the CPU")
    g.add_argument("--data-root", default="./data", help="where MNIST/raw/*-ubyte is looked for")
    g.add_argument("--load-model", default=None, metavar="PATH",
                   help="initialise weights from a reference-format checkpoint")
    g.add_argument("--comm", choices=["auto", "oneshot", "twoshot", "nvls", "nccl"], default="auto",
                   help="gradient reduction path of the native engine (nccl = baseline, unfused)")
    g.add_argument("--no-graph", action="store_true", default=False,
                   help="launch kernels eagerly instead of replaying the step CUDA graph")
    g.add_argument("--eval-all-ranks", action="store_true", default=False,
                   help="run the eval pass on every rank (the reference runs it on rank 0 only)")
    g.add_argument("--require-real-data", action="store_true", default=False,
                   help="fail if the MNIST idx files are missing instead of training on synthetic data")
    g.add_argument("--verify-replicas", type=int, default=0, metavar="N",
                   help="every N optimizer steps compare all ranks' parameters bit for bit (0 = off)")
    return p


def _make_backend(args, use_cuda: bool, train, test, info: DistInfo):
    want = args.backend
    if want == "auto":
        want = "b200" if use_cuda else "torch"
    if want == "b200":
        if not use_cuda:
            raise SystemExit("--backend b200 needs CUDA (drop --no-cuda)")
        from .engine import B200Backend  # imports (and if needed builds) the native library
        return B200Backend(train, test, info, lr=args.lr, seed=args.seed, comm=args.comm,
                           use_graph=not args.no_graph,
                           max_batch=max(args.batch_size, args.test_batch_size),
                           verify_replicas=args.verify_replicas)
    return TorchBackend(train, test, info, lr=args.lr, device="cuda" if use_cuda else "cpu")


def _run(args, ddp: bool) -> None:
    use_cuda = not args.no_cuda and torch.cuda.is_available()
    if ddp:
        info = init_distributed_mode(args, use_cuda)
    else:
        info = DistInfo()
    torch.manual_seed(args.seed)  # same seed on every rank (mnist_ddp.py:140)

    train = load_split(args.data_root, True, announce=info.is_main, require_real=args.require_real_data)
    test = load_split(args.data_root, False, require_real=args.require_real_data)
    backend = _make_backend(args, use_cuda, train, test, info)
    if args.load_model:
        backend.load_state_dict(checkpoint.load(args.load_model))

    if ddp:
        mode = "distributed" if info.distributed else "random"
    else:
        mode = "random" if use_cuda else "sequential"  # mnist.py:105-110 quirk
    fit(args, backend, train, test, info, sampler_mode=mode, eval_all_ranks=args.eval_all_ranks)

    if args.save_model:
        if not ddp:
            checkpoint.save(backend.state_dict(), "mnist_cnn.pt")  # mnist.py:133
        elif info.distributed:
            if info.is_main:  # keys carry the DDP wrapper's "module." prefix (mnist_ddp.py:195)
                checkpoint.save(backend.state_dict(), "mnist_cnn.pt", prefix="module.")
        else:
            checkpoint.save(backend.state_dict(), "mnist_cnn_.pt")  # mnist_ddp.py:197
    if ddp and info.distributed:
        shutdown()


def main_single(argv: Optional[List[str]] = None) -> None:
    _run(build_parser(False).parse_args(argv), ddp=False)


def main_ddp(argv: Optional[List[str]] = None) -> None:
    start = time.time()
    _run(build_parser(True).parse_args(argv), ddp=True)
    print(logfmt.total_time_line(time.time() - start))
