"""Python face of the native training engine (csrc/engine.cu) and the B200 trainer backend.

`NativeEngine` is a thin ctypes wrapper: it owns an engine handle, exposes the engine's device
buffers as zero-copy ``torch`` tensors (``__cuda_array_interface__``), and forwards
``train_step`` / ``eval`` calls.  All engine work is enqueued on ONE dedicated CUDA stream
(``engine.stream``; the legacy default stream cannot be graph-captured); helper methods that mix
torch ops with engine buffers run those ops on the same stream, and every method that hands data
back to the caller synchronises that stream first.  `B200Backend` adapts the engine to the epoch
driver in trainer.py, i.e. it is what ``mnist.py`` / ``mnist_ddp.py`` run on a GPU.

Reference call sites replaced: train() body mnist_ddp.py:67-73, test() body mnist_ddp.py:93-99,
``Net().to(device)`` :170, ``optim.Adadelta`` :176, the DDP wrap :173 (see parallel/fused_ddp.py).
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, Optional, Tuple

import torch

from . import _native as N
from .models.net import Net, ParamLayout
from .parallel.dist import DistInfo
from .utils.data import MnistSplit

# buffer ids (include/b200mnist.h)
(BUF_PARAMS, BUF_GRADS0, BUF_GRADS1, BUF_SQ_AVG, BUF_ACC_DELTA, BUF_A1, BUF_POOLED, BUF_AMAX, BUF_H,
 BUF_DLOGITS, BUF_DH, BUF_DP, BUF_DA1, BUF_LOSS_I, BUF_LOSS_RING, BUF_CTRL, BUF_LOGP, BUF_STAGE_IMG,
 BUF_STAGE_LBL, BUF_FC1_PART) = range(20)
LOSS_RING = 1024
# B200EngineConfig.use_tc bits (csrc/engine.h): which layers run on the tcgen05 kernels; TC_FC1_FP32 makes the fc1
# GEMMs 3xTF32 split-operand products (fp32-class, what the reference's fp32 SGEMM delivers)
TC_CONV2_FWD, TC_FC1_FWD, TC_FC1_BWD, TC_CONV2_BWD, TC_FC1_FP32 = 1, 2, 4, 8, 16
TC_ALL = 15              # every layer on tensor cores, fc1 in single-pass TF32
DEFAULT_TC = TC_ALL | TC_FC1_FP32
_NVTX_STEPS = os.environ.get("B200_NVTX") == "1"


def default_tc_mask() -> int:
    """Which layers run on tcgen05 tensor cores (bit mask); B200_TC=0 forces CUDA cores, 15 = TF32 fc1."""
    env = os.environ.get("B200_TC")
    if env is not None:
        return int(env, 0)
    return DEFAULT_TC if N.lib().b200_has_tc() else 0


class _DevArray:
    def __init__(self, ptr: int, shape: Tuple[int, ...], typestr: str) -> None:
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}


_TYPESTR = {torch.float32: ("<f4", 4), torch.uint8: ("|u1", 1), torch.int32: ("<i4", 4), torch.float64: ("<f8", 8)}


class NativeEngine:
    def __init__(self, max_batch: int, train_n: int, test_n: int, seed: int = 1, world: int = 1, rank: int = 0,
                 use_tc: Optional[int] = None, device: Optional[int] = None,
                 external: Optional[Dict[str, int]] = None) -> None:
        if not torch.cuda.is_available():
            raise N.NativeError("the native engine needs a CUDA device")
        self.lib = N.lib()
        self.layout = ParamLayout()
        self.device = torch.cuda.current_device() if device is None else device
        # highest priority: the engine's branch streams (csrc/engine.cu) run at the lowest, so that critical-path
        # kernels get the SMs first wherever the two compete
        self.stream = torch.cuda.Stream(device=self.device, priority=-5)
        self.max_batch, self.world, self.rank = max_batch, world, rank
        cfg = N.EngineConfig()
        cfg.device, cfg.max_batch, cfg.world, cfg.rank, cfg.seed = self.device, max_batch, world, rank, seed
        cfg.use_tc = default_tc_mask() if use_tc is None else use_tc
        cfg.train_n, cfg.test_n, cfg.param_floats = train_n, test_n, self.layout.total
        for i, o in enumerate(self.layout.offsets()):
            cfg.offsets[i] = o
        for b, (s, e) in enumerate(self.layout.bucket_ranges):
            cfg.bucket_begin[b], cfg.bucket_end[b] = s, e
        ext = external or {}
        cfg.params = ext.get("params")
        cfg.grads[0], cfg.grads[1] = ext.get("grads0"), ext.get("grads1")
        cfg.sq_avg, cfg.acc_delta = ext.get("sq_avg"), ext.get("acc_delta")
        self.use_tc = cfg.use_tc
        self._cfg = cfg
        h = ctypes.c_void_p()
        torch.cuda.synchronize(self.device)
        N.check(self.lib.b200_engine_create(ctypes.byref(cfg), ctypes.byref(h)))
        self.handle = h
        self._keep = []
        self.params = self.buffer(BUF_PARAMS, torch.float32)
        self.grads = self.buffer(BUF_GRADS0, torch.float32)
        self.sq_avg = self.buffer(BUF_SQ_AVG, torch.float32)
        self.acc_delta = self.buffer(BUF_ACC_DELTA, torch.float32)
        self.loss_ring = self.buffer(BUF_LOSS_RING, torch.float32)
        ring = ctypes.POINTER(ctypes.c_float)()
        N.check(self.lib.b200_engine_host_loss_ring(self.handle, ctypes.byref(ring)))
        # mapped pinned host memory that the head kernel writes every step (zero-copy D2H)
        self.host_loss_ring = torch.frombuffer((ctypes.c_float * LOSS_RING).from_address(
            ctypes.addressof(ring.contents)), dtype=torch.float32)
        self.steps = 0

    # ---- plumbing
    def _s(self) -> int:
        return self.stream.cuda_stream

    def on_stream(self):
        """Context manager: torch ops inside are ordered with the engine's work."""
        return torch.cuda.stream(self.stream)

    def synchronize(self) -> None:
        try:
            self.stream.synchronize()
        except RuntimeError as e:
            # a cross-GPU barrier that gave up on a lost peer traps; say who waited for whom instead of only
            # "unspecified launch failure"
            fault = self.lib.b200_comm_fault_message().decode()
            if fault:
                raise N.NativeError(f"{fault} [{e}]") from e
            raise

    def buffer(self, which: int, dtype: torch.dtype, shape: Optional[Tuple[int, ...]] = None) -> torch.Tensor:
        p, nb = ctypes.c_void_p(), ctypes.c_size_t()
        N.check(self.lib.b200_engine_buffer(self.handle, which, ctypes.byref(p), ctypes.byref(nb)))
        ts, isz = _TYPESTR[dtype]
        if shape is None:
            shape = (nb.value // isz,)
        return torch.as_tensor(_DevArray(p.value, shape, ts), device=f"cuda:{self.device}")

    def close(self) -> None:
        if getattr(self, "handle", None):
            self.lib.b200_engine_destroy(self.handle)
            self.handle = None

    def __del__(self) -> None:  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    # ---- data / parameters
    def load_data(self, split: int, data: MnistSplit) -> None:
        img = data.images.contiguous()
        lbl = data.labels.contiguous()
        N.check(self.lib.b200_engine_load_data(self.handle, split, img.data_ptr(), lbl.data_ptr(), len(data)))

    def set_state_dict(self, state: Dict[str, torch.Tensor]) -> None:
        flat = self.layout.pack(state).pin_memory()
        with self.on_stream():
            self.params.copy_(flat, non_blocking=True)
        self.synchronize()

    def state_dict(self, prefix: str = "") -> Dict[str, torch.Tensor]:
        self.synchronize()
        return self.layout.unpack(self.params, prefix)

    def grad_dict(self) -> Dict[str, torch.Tensor]:
        self.synchronize()
        return self.layout.unpack(self.grads)

    # ---- training
    def begin_epoch(self, idx: Optional[torch.Tensor]) -> None:
        if idx is None:
            N.check(self.lib.b200_engine_begin_epoch(self.handle, None, self._cfg.train_n, self._s()))
            return
        pinned = idx.to(torch.int32).contiguous().pin_memory()
        self._keep = (self._keep + [pinned])[-2:]  # must outlive the async copy
        N.check(self.lib.b200_engine_begin_epoch(self.handle, pinned.data_ptr(), pinned.numel(), self._s()))

    def set_lr(self, lr: float) -> None:
        N.check(self.lib.b200_engine_set_lr(self.handle, float(lr), self._s()))

    def set_dropout(self, on: bool) -> None:
        N.check(self.lib.b200_engine_set_dropout(self.handle, int(on), self._s()))

    def train_step(self, batch: int, use_graph: bool = True) -> int:
        if _NVTX_STEPS:  # B200_NVTX=1: one NVTX range per optimizer step (epoch / train / test ranges are always on)
            torch.cuda.nvtx.range_push(f"step {self.steps}")
        N.check(self.lib.b200_engine_train_step(self.handle, batch, int(use_graph), self._s()))
        if _NVTX_STEPS:
            torch.cuda.nvtx.range_pop()
        self.steps += 1
        return self.steps - 1

    def train_step_host(self, images_u8: torch.Tensor, labels_u8: torch.Tensor, use_graph: bool = True) -> int:
        """Public streaming API: the batch is copied from (pinned) host memory every step."""
        b = int(images_u8.shape[0])
        N.check(self.lib.b200_engine_train_step_host(self.handle, images_u8.data_ptr(), labels_u8.data_ptr(), b,
                                                     int(use_graph), self._s()))
        self.steps += 1
        return self.steps - 1

    def train_step_gather(self, images_u8: torch.Tensor, labels_u8: torch.Tensor, idx_i32: torch.Tensor,
                          use_graph: bool = True) -> int:
        """The data-loader path: `idx_i32` (host int32) selects this step's samples from a HOST-resident uint8
        dataset; they are gathered into pinned staging memory by the native library, copied to the GPU on the copy
        stream (overlapping the previous step) and trained on."""
        if not (images_u8.dtype == torch.uint8 and labels_u8.dtype == torch.uint8 and idx_i32.dtype == torch.int32
                and images_u8.device.type == labels_u8.device.type == idx_i32.device.type == "cpu"
                and images_u8.is_contiguous() and labels_u8.is_contiguous() and idx_i32.is_contiguous()
                and images_u8.shape[0] == labels_u8.shape[0] and images_u8[0].numel() == 784):
            raise ValueError("train_step_gather: expected contiguous host tensors uint8 [n,28,28] / uint8 [n] / int32 [batch]")
        # indices are range-checked by the native side against the dataset size
        N.check(self.lib.b200_engine_train_step_gather(self.handle, images_u8.data_ptr(), labels_u8.data_ptr(),
                                                       int(images_u8.shape[0]), idx_i32.data_ptr(),
                                                       int(idx_i32.numel()), int(use_graph), self._s()))
        self.steps += 1
        return self.steps - 1

    def read_loss(self, step: int) -> float:
        out = ctypes.c_float()
        N.check(self.lib.b200_engine_read_loss(self.handle, step, ctypes.byref(out), self._s()))
        return float(out.value)

    def launches_per_step(self, batch: int) -> int:
        n = ctypes.c_int()
        N.check(self.lib.b200_engine_launches_per_step(self.handle, batch, ctypes.byref(n)))
        return int(n.value)

    # ---- evaluation
    def evaluate(self, split: int, n: int, batch: int) -> Tuple[float, int]:
        s = self._s()
        N.check(self.lib.b200_engine_eval_reset(self.handle, s))
        for first in range(0, n, batch):
            N.check(self.lib.b200_engine_eval_batch(self.handle, split, first, min(batch, n - first), s))
        nll, correct = ctypes.c_double(), ctypes.c_longlong()
        N.check(self.lib.b200_engine_eval_result(self.handle, ctypes.byref(nll), ctypes.byref(correct), s))
        return float(nll.value), int(correct.value)

    def forward_logp(self, split: int, first: int, batch: int) -> torch.Tensor:
        """Eval-mode log-probabilities of `batch` samples of a split (for tests / inference)."""
        N.check(self.lib.b200_engine_eval_batch(self.handle, split, first, batch, self._s()))
        with self.on_stream():
            out = self.buffer(BUF_LOGP, torch.float32)[: batch * 10].view(batch, 10).clone()
        self.synchronize()
        return out


class B200Backend:
    """trainer.fit() backend running on the native engine (one process = one GPU)."""

    name = "b200"

    def __init__(self, train: MnistSplit, test: MnistSplit, info: DistInfo, lr: float, seed: int = 1,
                 comm: str = "auto", use_graph: bool = True, max_batch: int = 4096, verify_replicas: int = 0) -> None:
        self.info = info
        self.verify_replicas = int(verify_replicas)
        self.use_graph = use_graph
        self.train_split, self.test_split = train, test
        world = info.world_size if info.distributed else 1
        model = Net()  # CPU construction right after manual_seed: reference init stream
        self.ddp = None
        external = None
        # B200_FORCE_DDP=1 runs the data-parallel code path even with one rank (overhead measurements)
        if (info.distributed and world > 1) or os.environ.get("B200_FORCE_DDP") == "1":
            from .parallel.fused_ddp import FusedDataParallel
            self.ddp = FusedDataParallel(info, comm=comm)
            external = self.ddp.allocate(ParamLayout().total)
        self.engine = NativeEngine(max_batch, len(train), len(test), seed=seed, world=world, rank=info.rank,
                                   external=external)
        self.engine.load_data(0, train)
        self.engine.load_data(1, test)
        self.engine.set_state_dict(model.state_dict())
        self._nccl_baseline = False
        if self.ddp is not None:
            self.ddp.attach(self.engine)  # broadcasts rank 0's parameters, installs the bucket hooks
            if self.ddp.comm == "nccl":   # in-tree baseline: NCCL all-reduce + separate Adadelta kernel
                self._nccl_baseline = True
                N.check(self.engine.lib.b200_engine_set_defer_update(self.engine.handle, 1))
        self.engine.set_lr(lr)
        self._pinned_loss = torch.empty(64, dtype=torch.float32).pin_memory()
        self._slot = 0
        self._epoch_n = len(train)

    def begin_epoch(self, indices: torch.Tensor) -> None:
        self._epoch_n = int(indices.numel())
        self.engine.begin_epoch(indices)

    def train_step(self, step: int, batch: int, want_loss: bool = True):
        bsz = min(batch, self._epoch_n - step * batch)
        s = self.step(bsz)
        if not want_loss:
            return None
        slot = self._slot
        self._slot = (self._slot + 1) % 64
        with self.engine.on_stream():
            self._pinned_loss[slot: slot + 1].copy_(self.engine.loss_ring[s % LOSS_RING: s % LOSS_RING + 1],
                                                    non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        return (slot, ev)

    def step(self, bsz: int) -> int:
        """One full optimizer step on the next `bsz` samples of the epoch (whatever the comm mode)."""
        s = self.engine.train_step(bsz, self.use_graph)
        if self._nccl_baseline:
            self._nccl_reduce_and_update(bsz)
        if self.verify_replicas and self.ddp is not None and (s + 1) % self.verify_replicas == 0:
            drift = self.ddp.replica_drift()  # collective: every rank gets here at the same step
            if drift != 0.0:
                raise N.NativeError(f"replicas diverged at step {s}: max |p_rank - p_rank0| = {drift} "
                                    f"(--comm {self.ddp.mode}); the fused reduction must keep them bit-identical")
        return s

    def step_host(self, images_u8: torch.Tensor, labels_u8: torch.Tensor) -> int:
        """Same as `step`, but the batch is streamed from pinned host memory (uint8 images, uint8 labels)."""
        s = self.engine.train_step_host(images_u8, labels_u8, self.use_graph)
        if self._nccl_baseline:
            self._nccl_reduce_and_update(0)
        return s

    def step_gather(self, images_u8: torch.Tensor, labels_u8: torch.Tensor, idx_i32: torch.Tensor) -> int:
        """Same as `step_host`, with the batch gathered by index from a host-resident uint8 dataset."""
        s = self.engine.train_step_gather(images_u8, labels_u8, idx_i32, self.use_graph)
        if self._nccl_baseline:
            self._nccl_reduce_and_update(0)
        return s

    def _nccl_reduce_and_update(self, bsz: int) -> None:
        """Reference-style reduction (mnist_ddp.py:173 via ncclAllReduce per bucket), then Adadelta."""
        import torch.distributed as dist
        eng = self.engine
        with eng.on_stream():
            for lo, hi in eng.layout.bucket_ranges:
                dist.all_reduce(eng.grads[lo:hi], op=dist.ReduceOp.SUM)
            eng.grads.mul_(1.0 / self.info.world_size)
        N.check(eng.lib.b200_engine_apply_update(eng.handle, bsz, eng._s()))

    def loss(self, token) -> float:
        slot, ev = token
        ev.synchronize()
        return float(self._pinned_loss[slot])

    def evaluate(self, split: MnistSplit, batch: int) -> Tuple[float, int]:
        which = 1 if split is self.test_split else 0
        return self.engine.evaluate(which, len(split), min(batch, self.engine.max_batch))

    def set_lr(self, lr: float) -> None:
        self.engine.set_lr(lr)

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return self.engine.state_dict()

    def load_state_dict(self, state: Dict[str, torch.Tensor]) -> None:
        self.engine.set_state_dict(state)

    def finish(self) -> None:
        self.engine.synchronize()
