"""Data parallelism without DistributedDataParallel: symmetric buffers + fused P2P kernels.

What ``DistributedDataParallel(model, device_ids=[gpu])`` (mnist_ddp.py:173) sets up — parameter
broadcast from rank 0, gradient buckets, an all-reduce per bucket overlapped with backward —
is rebuilt here on top of csrc/symm.cu and csrc/allreduce.cu:

* `SymmetricArena`: one device allocation per rank holding ``[params | grads | sq_avg | acc_delta | reduced conv
  grads]`` plus barrier flags, exported with CUDA IPC and mapped into every peer.  ``torch.distributed`` is
  used for bootstrap only (all-gather of the 64-byte handles, barriers).
* `FusedDataParallel.allocate()` hands the arena's buffers to the native engine, so backward
  kernels write gradients straight into the peer-visible bucket (no bucket copy-in / copy-out).
* `FusedDataParallel.attach()` pulls rank 0's parameters over NVLink (DDP's ``_sync_module_states``)
  and installs the engine hooks that launch the fused all-reduce + 1/W + Adadelta kernels: the fc bucket
  on a branch stream while the conv backward still runs, conv2's half of the conv bucket on the wgrad branch,
  conv1's 1.3 KB and the conv-bucket update in the step's last kernel (SURVEY N17 bucket layout; section 8
  rows 11, 12, 15).

Reduction modes (``--comm``): ``oneshot`` (every rank reduces the whole bucket), ``twoshot``
(reduce-scatter + sharded Adadelta + parameter all-gather), ``nvls`` = two-shot through the NVSwitch
(``multimem.ld_reduce`` + ``multimem.st`` on a multicast mapping from torch symmetric memory), ``auto`` =
``twoshot`` at 2 ranks, ``nvls`` from 4 ranks on when such a mapping exists (measured), ``twoshot`` otherwise.  The mode applies to the fc bucket and to the
conv2 half of the conv bucket (reduced on the wgrad branch, under the rest of the backward pass); conv1's
1.3 KB is exchanged one-shot by the step's last kernel.  ``nccl`` keeps the reference's unfused structure
(ncclAllReduce + separate Adadelta) as the in-tree baseline, and is what multi-host jobs fall back to.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from .. import _native as N
from .dist import DistInfo

ONESHOT, TWOSHOT, NVLS = 0, 1, 2


class SymmetricArena:
    """Peer-mapped device memory: ``payload_floats`` fp32 per rank + the barrier flag block."""

    def __init__(self, handle: ctypes.c_void_p, world: int, rank: int, payload_floats: int) -> None:
        self.handle, self.world, self.rank, self.payload_floats = handle, world, rank, payload_floats
        self._views: Dict[int, torch.Tensor] = {}
        # barrier tokens must keep growing per channel across ALL plans that share the flag words
        self.tokens: Dict[int, int] = {}

    @classmethod
    def create(cls, info: DistInfo, payload_floats: int, exchange: bool = True) -> "SymmetricArena":
        """exchange=False: a purely local arena (no IPC handles exchanged, peers unmapped) for the NCCL baseline."""
        lib = N.lib()
        h = ctypes.c_void_p()
        N.check(lib.b200_symm_create(info.world_size, info.rank, payload_floats * 4, ctypes.byref(h)))
        arena = cls(h, info.world_size, info.rank, payload_floats)
        if info.world_size > 1 and exchange:
            nb = lib.b200_symm_handle_bytes()
            mine = (ctypes.c_char * nb)()
            N.check(lib.b200_symm_export(h, mine))
            dev = torch.device("cuda", torch.cuda.current_device())
            local = torch.frombuffer(bytearray(mine.raw), dtype=torch.uint8).to(dev)
            everyone = torch.empty(info.world_size * nb, dtype=torch.uint8, device=dev)
            dist.all_gather_into_tensor(everyone, local)
            blob = bytes(everyone.cpu().numpy().tobytes())
            N.check(lib.b200_symm_import(h, blob))
            dist.barrier()
        return arena

    @classmethod
    def create_torch(cls, info: DistInfo, payload_floats: int) -> "SymmetricArena":
        """Arena backed by ``torch.distributed._symmetric_memory``: same peer mapping, plus the NVLS
        multicast alias (``multimem.ld_reduce`` / ``multimem.st`` through the NVSwitch) when the driver and
        fabric allow it.  Allocation and handle exchange are torch's; the kernels are ours."""
        import torch.distributed._symmetric_memory as symm_mem
        lib = N.lib()
        payload_floats = (payload_floats + 63) // 64 * 64
        flag_floats = lib.b200_symm_flag_bytes() // 4
        dev = torch.device("cuda", torch.cuda.current_device())
        buf = symm_mem.empty(payload_floats + flag_floats, dtype=torch.float32, device=dev)
        buf.zero_()
        hdl = symm_mem.rendezvous(buf, group=dist.group.WORLD)
        torch.cuda.synchronize()
        dist.barrier()
        ptrs = (ctypes.c_uint64 * info.world_size)(*[int(p) for p in hdl.buffer_ptrs])
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        h = ctypes.c_void_p()
        N.check(lib.b200_symm_wrap(info.world_size, info.rank, payload_floats * 4, ptrs, mc, ctypes.byref(h)))
        arena = cls(h, info.world_size, info.rank, payload_floats)
        arena._torch_keepalive = (buf, hdl)
        arena.has_multicast = mc != 0
        return arena

    @classmethod
    def create_emulated(cls, world: int, payload_floats: int) -> List["SymmetricArena"]:
        """`world` arenas on the current GPU, cross-linked like real ranks (single-GPU tests)."""
        lib = N.lib()
        hs = (ctypes.c_void_p * world)()
        N.check(lib.b200_symm_create_emulated(world, payload_floats * 4, hs))
        return [cls(ctypes.c_void_p(hs[r]), world, r, payload_floats) for r in range(world)]

    def base_ptr(self, rank: Optional[int] = None) -> int:
        p = ctypes.c_void_p()
        N.check(N.lib().b200_symm_peer(self.handle, self.rank if rank is None else rank, ctypes.byref(p)))
        return p.value

    def view(self, rank: Optional[int] = None) -> torch.Tensor:
        """The payload of a rank's arena as a flat fp32 tensor (peer views are real NVLink mappings)."""
        from ..engine import _DevArray
        r = self.rank if rank is None else rank
        if r not in self._views:
            arr = _DevArray(self.base_ptr(r), (self.payload_floats,), "<f4")
            self._views[r] = torch.as_tensor(arr, device=f"cuda:{torch.cuda.current_device()}")
        return self._views[r]

    def destroy(self) -> None:
        if self.handle:
            N.lib().b200_symm_destroy(self.handle)
            self.handle = None


class CommPlan:
    """One bucket bound to a barrier channel; `run` launches the fused kernel."""

    def __init__(self, handle: ctypes.c_void_p, arena: "SymmetricArena", channel: int) -> None:
        self.handle, self.arena, self.channel = handle, arena, channel

    @classmethod
    def create(cls, arena: SymmetricArena, grad_off: int, param_off: int, v_off: int, u_off: int, n: int,
               channel: int, ctrl: Optional[torch.Tensor] = None) -> "CommPlan":
        h = ctypes.c_void_p()
        N.check(N.lib().b200_comm_plan_create(arena.handle, grad_off, param_off, v_off, u_off, n, channel,
                                              None if ctrl is None else ctrl.data_ptr(), ctypes.byref(h)))
        return cls(h, arena, channel)

    @classmethod
    def create_emulated(cls, arenas: List[SymmetricArena], grad_off: int, param_off: int, v_off: int, u_off: int,
                        n: int, channel: int, ctrls: Optional[List[torch.Tensor]] = None) -> "CommPlan":
        w = len(arenas)
        hs = (ctypes.c_void_p * w)(*[a.handle for a in arenas])
        cs = None
        if ctrls is not None:
            cs = (ctypes.c_void_p * w)(*[c.data_ptr() for c in ctrls])
        h = ctypes.c_void_p()
        N.check(N.lib().b200_comm_plan_create_emulated(hs, w, grad_off, param_off, v_off, u_off, n, channel, cs,
                                                       ctypes.byref(h)))
        return cls(h, arenas[0], channel)

    def run(self, mode: int, opt: bool, nblocks: int = 0, token: Optional[int] = None) -> None:
        """token=None: use a private, monotonically increasing token; 0: derive from ctrl->step."""
        if token is None:
            token = self.arena.tokens.get(self.channel, 1)
            self.arena.tokens[self.channel] = token + 2
        N.check(N.lib().b200_comm_run(self.handle, mode, int(opt), token, nblocks,
                                      torch.cuda.current_stream().cuda_stream))

    def destroy(self) -> None:
        if self.handle:
            N.lib().b200_comm_plan_destroy(self.handle)
            self.handle = None


class FusedDataParallel:
    """Arena layout per rank, in floats: ``[params T | grads T | sq_avg T | acc_delta T | reduced conv grads T]``."""

    REGIONS = 5

    def __init__(self, info: DistInfo, comm: str = "auto") -> None:
        if comm not in ("auto", "oneshot", "twoshot", "nccl", "nvls"):
            raise ValueError(f"unknown comm mode {comm}")
        self.info, self.comm = info, comm
        # (fc bucket, conv2 half of the conv bucket); conv1's 1.3 KB is always exchanged one-shot in the last kernel
        self.modes = {"oneshot": (ONESHOT, ONESHOT), "twoshot": (TWOSHOT, TWOSHOT), "nvls": (NVLS, NVLS),
                      "auto": (TWOSHOT, TWOSHOT), "nccl": (ONESHOT, ONESHOT)}[comm]
        self.mode = comm
        self.arena: Optional[SymmetricArena] = None
        self.total = 0

    def _single_host(self) -> bool:
        """CUDA IPC and multicast mappings only exist between GPUs of one host: compare host names."""
        if not (dist.is_initialized() and self.info.world_size > 1):
            return True
        import socket
        names = [None] * self.info.world_size
        dist.all_gather_object(names, socket.gethostname())
        return len(set(names)) == 1

    def allocate(self, total_floats: int) -> Dict[str, int]:
        """Create the arena and return the raw pointers the engine must use for its flat buffers."""
        self.total = total_floats
        n = self.REGIONS * total_floats
        if self.comm != "nccl" and not self._single_host():
            # the reference's env contract reaches across nodes (mnist_ddp.py:16-22,134); peer-mapped memory does not
            if self.info.rank == 0:
                print(f"b200: ranks are on different hosts -- NVLink peer memory is unavailable, falling back from "
                      f"--comm {self.comm} to --comm nccl (ncclAllReduce per bucket + separate Adadelta)", flush=True)
            self.comm = self.mode = "nccl"
        if self.comm == "nvls":
            self.arena = SymmetricArena.create_torch(self.info, n)
            if not getattr(self.arena, "has_multicast", False):
                raise N.NativeError("--comm nvls: this host/driver gave no multicast mapping (NVLS unavailable)")
        elif self.comm == "auto":
            # NVLS (in-switch reduction, constant bytes per GPU) when torch symmetric memory hands out a multicast
            # mapping, two-shot over CUDA IPC otherwise.  Every rank must take the same branch, so the outcome of
            # the probe is agreed on with an all-reduce.
            # Measured (us/step, batch 200): at 2 GPUs two-shot 149.0 vs NVLS 160.2 (with one peer the switch adds a
            # hop and saves no traffic); at 4 GPUs 157.7 vs 157.0, at 8 GPUs 163.7 vs 160.7.  So NVLS from 4
            # ranks on, where it moves 2*S/W bytes per GPU instead of 2*(W-1)/W*S.
            arena, ok = None, 0
            if os.environ.get("B200_NO_NVLS") != "1" and self.info.world_size >= 4:
                try:
                    arena = SymmetricArena.create_torch(self.info, n)
                    ok = 1 if getattr(arena, "has_multicast", False) else 0
                except Exception:
                    ok = 0
            flag = torch.tensor([ok], device="cuda", dtype=torch.int32)
            if dist.is_initialized():
                dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) == 1:
                self.arena = arena
                self.modes = (NVLS, NVLS)
                self.mode = "auto(nvls)"
            else:
                if arena is not None:
                    arena.destroy()
                self.arena = SymmetricArena.create(self.info, n)
                self.mode = "auto(twoshot)"
        else:
            self.arena = SymmetricArena.create(self.info, n, exchange=self.comm != "nccl")
        env = os.environ.get("B200_DDP_MODES")  # diagnostics: "fc,conv2" with 0 one-shot / 1 two-shot / 2 NVLS
        if env:
            self.modes = tuple(int(x) for x in env.split(","))[:2]
            self.mode += f"[modes {self.modes[0]},{self.modes[1]}]"
        base = self.arena.base_ptr()
        t = total_floats * 4
        return {"params": base, "grads0": base + t, "grads1": base + t, "sq_avg": base + 2 * t,
                "acc_delta": base + 3 * t}

    def attach(self, engine) -> None:
        lib = N.lib()
        # DDP's _sync_module_states: every rank takes rank 0's parameters, pulled over NVLink
        barrier = dist.barrier if dist.is_initialized() else (lambda: None)
        engine.synchronize()
        barrier()
        if self.comm == "nccl":
            if dist.is_initialized():
                with engine.on_stream():
                    dist.broadcast(engine.params, src=0)
        else:
            N.check(lib.b200_symm_pull(self.arena.handle, 0, 0, self.total, engine._s()))
        engine.synchronize()
        barrier()
        if self.comm != "nccl":
            N.check(lib.b200_ddp_attach(engine.handle, self.arena.handle, self.modes[0], self.modes[1]))
        barrier()

    def replica_drift(self) -> float:
        """max |p_rank - p_rank0| over all parameters, reduced over ranks.  Must be exactly 0: every element is
        summed by one owner (two-shot / NVLS) or in a fixed rank order (one-shot), so replicas stay bit-identical.
        Collective: every rank must call it at the same step."""
        torch.cuda.synchronize()
        if dist.is_initialized():
            dist.barrier()
        mine = self.arena.view()[: self.total]
        if self.comm == "nccl":
            ref = mine.clone()
            if dist.is_initialized():
                dist.broadcast(ref, src=0)
        else:
            ref = self.arena.view(0)[: self.total]
        d = (mine - ref).abs().max()
        if dist.is_initialized():
            dist.all_reduce(d, op=dist.ReduceOp.MAX)
        return float(d.item())

    verify_replicas = replica_drift
