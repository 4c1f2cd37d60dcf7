"""The reference's three log lines, byte for byte."""
from __future__ import annotations


def banner(rank: int, url: str, gpu: int, world: int) -> str:
    # mnist_ddp.py:34
    return f"| distributed init (rank {rank}): {url}, local rank:{gpu}, world size:{world}"


NOT_DISTRIBUTED = "Not using distributed mode"  # mnist_ddp.py:26


def train_line(epoch: int, seen: int, total: int, batch_idx: int, n_batches: int, loss: float) -> str:
    # mnist_ddp.py:77-79 / 82-84; `seen` already includes the world-size factor in DDP mode
    return "Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}".format(
        epoch, seen, total, 100.0 * batch_idx / n_batches, loss)


def test_line(avg_loss: float, correct: int, total: int) -> str:
    # mnist_ddp.py:103-105
    return "\nTest set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n".format(
        avg_loss, correct, total, 100.0 * correct / total)


def total_time_line(seconds: float) -> str:
    # mnist_ddp.py:203 — the value is seconds although the label says ms; kept for parity
    return f"Total cost time:{seconds} ms"
