"""torch-tensor wrappers over the individual C-ABI kernels (include/b200mnist.h).

These are what the numerics tests call: each wrapper allocates outputs with torch, passes raw
device pointers and torch's current stream to the kernel, and returns tensors in the engine's
internal layouts (activations NHWC, conv2 weight [tap][o][i], conv1 weight [tap][o]).  The
training engine does not go through here — it sequences the same kernels in C++.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from .. import _native as N

DROP1_SCALE = 1.0 / 0.75


def _s() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def make_ctrl(step: int = 0, seed: int = 1, pos: int = 0, train: bool = True, lr: float = 1.0,
              inv_world: float = 1.0) -> torch.Tensor:
    L = N.lib()
    ctrl = torch.zeros(L.b200_ctrl_bytes(), dtype=torch.uint8, device="cuda")
    # train: False -> 0 (eval), True -> 3 (backward + dropout), or an explicit bit mask (1 = no dropout)
    flag = (3 if train else 0) if isinstance(train, bool) else int(train)
    N.check(L.b200_ctrl_set(ctrl.data_ptr(), step, seed, pos, flag, lr, inv_world, _s()))
    return ctrl


def read_ctrl(ctrl: torch.Tensor) -> dict:
    raw = ctrl.cpu().numpy().tobytes()
    import struct
    step, seed, pos, train, lr, inv_world = struct.unpack_from("<QQiiff", raw, 0)
    return dict(step=step, seed=seed, pos=pos, train=train, lr=lr, inv_world=inv_world)


def conv1_fwd(images_u8: torch.Tensor, idx: Optional[torch.Tensor], ctrl: torch.Tensor, w: torch.Tensor,
              b: torch.Tensor, batch: int) -> torch.Tensor:
    a1 = torch.empty(batch, 26, 26, 32, device="cuda")
    N.check(N.lib().b200_conv1_fwd(_p(images_u8), _p(idx), _p(ctrl), _p(w), _p(b), _p(a1), batch, _s()))
    return a1


def conv12_fwd(images_u8: torch.Tensor, idx: Optional[torch.Tensor], ctrl: torch.Tensor, w1: torch.Tensor,
               b1: torch.Tensor, w2: torch.Tensor, b2: torch.Tensor, batch: int):
    """conv1 + conv2 forward in one launch (csrc/tc_conv2_fwd.cu): returns (a1, pooled, amax)."""
    a1 = torch.empty(batch, 26, 26, 32, device="cuda")
    pooled = torch.empty(batch, 12, 12, 64, device="cuda")
    amax = torch.empty(batch, 12, 12, 64, dtype=torch.uint8, device="cuda")
    N.check(N.lib().b200_conv12_fwd_tc(_p(images_u8), _p(idx), _p(ctrl), _p(w1), _p(b1), _p(a1), _p(w2), _p(b2),
                                       _p(pooled), _p(amax), batch, _s()))
    return a1, pooled, amax


def conv1_wgrad(images_u8, idx, ctrl, da1: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    batch = da1.shape[0]
    part = torch.empty(2 * batch, 320, device="cuda")
    dw, db = torch.empty(9, 32, device="cuda"), torch.empty(32, device="cuda")
    N.check(N.lib().b200_conv1_wgrad(_p(images_u8), _p(idx), _p(ctrl), _p(da1), _p(part), _p(dw), _p(db), batch, _s()))
    return dw, db


def conv2_fwd(a1: torch.Tensor, w: torch.Tensor, b: torch.Tensor, ctrl: torch.Tensor, tc: bool = False):
    batch = a1.shape[0]
    pooled = torch.empty(batch, 12, 12, 64, device="cuda")
    amax = torch.empty(batch, 12, 12, 64, dtype=torch.uint8, device="cuda")
    L = N.lib()
    fn = L.b200_conv2_fwd_tc if tc else L.b200_conv2_fwd_simt
    N.check(fn(_p(a1), _p(w), _p(b), _p(ctrl), _p(pooled), _p(amax), batch, _s()))
    return pooled, amax


def conv2_dgrad(dp, amax, w, a1, tc: bool = False) -> torch.Tensor:
    batch = dp.shape[0]
    da1 = torch.empty(batch, 26, 26, 32, device="cuda")
    fn = N.lib().b200_conv2_dgrad_tc if tc else N.lib().b200_conv2_dgrad_simt
    N.check(fn(_p(dp), _p(amax), _p(w), _p(a1), _p(da1), batch, _s()))
    return da1


def conv2_wgrad(dp, amax, a1, tc: bool = False) -> Tuple[torch.Tensor, torch.Tensor]:
    batch = dp.shape[0]
    L = N.lib()
    parts = L.b200_conv2_wgrad_tc_parts(batch) if tc else L.b200_conv2_wgrad_parts()
    part = torch.empty(parts * (18432 + 64), device="cuda")
    dw, db = torch.empty(9, 64, 32, device="cuda"), torch.empty(64, device="cuda")
    fn = L.b200_conv2_wgrad_tc if tc else L.b200_conv2_wgrad_simt
    N.check(fn(_p(dp), _p(amax), _p(a1), _p(part), _p(dw), _p(db), batch, _s()))
    return dw, db


def gemm(A: torch.Tensor, B: torch.Tensor, a_mn: bool, b_mn: bool, M: int, N_: int, K: int, splits: int = 1,
         tc: bool = False, aux: Optional[torch.Tensor] = None, scale: float = 0.0, fp32: bool = False) -> torch.Tensor:
    """D[M][N] = A.B.  a_mn False: A is [M][K]; True: A is [K][M].  b_mn False: B is [N][K]; True: [K][N].
    tc: tcgen05 kernel (TF32; fp32=True: 3xTF32 split-operand product), else the fp32 CUDA-core GEMM."""
    D = torch.empty(splits, M, N_, device="cuda")
    epi = 0 if aux is None else 1
    L = N.lib()
    if tc:
        N.check(L.b200_tc_gemm_ex(_p(A), _p(B), _p(D), M, N_, K, int(a_mn), int(b_mn), splits, epi, _p(aux), scale,
                                  int(fp32), _s()))
    else:
        sam, sak = (1, M) if a_mn else (K, 1)
        sbk, sbn = (N_, 1) if b_mn else (1, K)
        N.check(L.b200_sgemm(_p(A), sam, sak, _p(B), sbk, sbn, _p(D), M, N_, K, splits, epi, _p(aux), scale, _s()))
    return D


def head_fused(part: torch.Tensor, b1, w2, b2, labels_u8, idx, ctrl, train: bool):
    splits, batch = part.shape[0], part.shape[1]
    h = torch.empty(batch, 128, device="cuda")
    dlogits = torch.zeros(batch, 16, device="cuda")
    dh = torch.zeros(batch, 128, device="cuda")
    loss_i = torch.empty(batch, device="cuda")
    correct = torch.zeros(batch, dtype=torch.int32, device="cuda")
    logp = torch.empty(batch, 10, device="cuda")
    N.check(N.lib().b200_head_fused(_p(part), splits, batch, _p(b1), _p(w2), _p(b2), _p(labels_u8), _p(idx), _p(ctrl),
                                    _p(h), _p(dlogits) if train else None, _p(dh) if train else None, _p(loss_i),
                                    _p(correct), _p(logp), _s()))
    return h, dlogits, dh, loss_i, correct, logp


def head_wgrad(h, dlogits, dh, loss_i, ctrl):
    batch = h.shape[0]
    dw2, db2, db1 = torch.empty(10, 128, device="cuda"), torch.empty(10, device="cuda"), torch.empty(128, device="cuda")
    loss = torch.empty(4, device="cuda")
    N.check(N.lib().b200_head_wgrad(_p(h), _p(dlogits), _p(dh), _p(loss_i), batch, _p(dw2), _p(db2), _p(db1), _p(loss),
                                    _p(ctrl), None, 0, None, _s()))
    return dw2, db2, db1, loss[0]


def adadelta(p, g, v, u, ctrl, rho: float = 0.9, eps: float = 1e-6, advance: int = -1) -> None:
    counter = torch.zeros(4, dtype=torch.int32, device="cuda")
    N.check(N.lib().b200_adadelta(_p(p), _p(g), _p(v), _p(u), p.numel(), _p(ctrl), rho, eps, advance, _p(counter), _s()))
