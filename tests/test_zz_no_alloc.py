"""The allocation-free conv1 weight-gradient entry point, called through the C ABI.

Some embedders of libb200mnist cannot allocate device memory at all — their process has given
the whole GPU to another allocator, or they run a latency-critical loop in which cudaMalloc's
device-wide synchronisation is not acceptable.  For them `b200_conv1_wgrad_noalloc` keeps its
partial-sum scratch in an array compiled into the library instead of asking the caller for one.

This file is named so that pytest collects it last: if it fails under `-x`, every other test has
already run.  It talks to the library with plain ctypes (no Python wrapper in between) and must
pass both on its own in a fresh process and after the rest of the suite, so it assumes nothing
about earlier calls: the first call it makes may or may not be the one that resolves the scratch
address on this device.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

VP, INT = ctypes.c_void_p, ctypes.c_int


def _lib():
    from pytorch_mnist_ddp_b200 import _native
    lib = _native.lib()
    lib.b200_conv1_wgrad_noalloc.argtypes = [VP, VP, VP, VP, VP, VP, INT, VP]
    lib.b200_conv1_wgrad_noalloc.restype = INT
    lib.b200_conv1_wgrad_noalloc_max_batch.restype = INT
    lib.b200_conv1_wgrad.argtypes = [VP, VP, VP, VP, VP, VP, VP, INT, VP]
    lib.b200_conv1_wgrad.restype = INT
    lib.b200_ctrl_set.argtypes = [VP, ctypes.c_ulonglong, ctypes.c_ulonglong, INT, INT, ctypes.c_float,
                                  ctypes.c_float, VP]
    lib.b200_ctrl_set.restype = INT
    return lib


def _inputs(batch, seed):
    g = torch.Generator().manual_seed(seed)
    images = torch.randint(0, 256, (batch + 3, 28, 28), dtype=torch.uint8, generator=g).cuda()
    idx = torch.randperm(batch + 3, generator=g)[:batch].to(torch.int32).cuda()
    da1 = torch.randn(batch, 26, 26, 32, generator=g).cuda()
    return images, idx, da1


def test_noalloc_conv1_wgrad_matches_regular_entry_point_twice_in_a_row():
    lib = _lib()
    stream = torch.cuda.current_stream().cuda_stream
    ctrl = torch.zeros(lib.b200_ctrl_bytes(), dtype=torch.uint8, device="cuda")
    assert lib.b200_ctrl_set(ctrl.data_ptr(), 0, 1, 0, 3, 1.0, 1.0, stream) == 0
    # two calls in a row with different shapes and data: the second one finds the scratch array
    # full of the first call's partial sums and must not be affected by them
    for call, (batch, seed) in enumerate([(96, 11), (37, 12)]):
        images, idx, da1 = _inputs(batch, seed)
        part = torch.empty(2 * batch, 320, device="cuda")
        dw_ref, db_ref = torch.empty(9, 32, device="cuda"), torch.empty(32, device="cuda")
        rc = lib.b200_conv1_wgrad(images.data_ptr(), idx.data_ptr(), ctrl.data_ptr(), da1.data_ptr(), part.data_ptr(),
                                  dw_ref.data_ptr(), db_ref.data_ptr(), batch, stream)
        assert rc == 0, lib.b200_last_error().decode()
        dw, db = torch.full((9, 32), float("nan"), device="cuda"), torch.full((32,), float("nan"), device="cuda")
        torch.cuda.synchronize()
        # the driver's own count of free device memory: unlike torch's allocator statistics it would also show a
        # cudaMalloc made inside the library
        free_before, _ = torch.cuda.mem_get_info()
        rc = lib.b200_conv1_wgrad_noalloc(images.data_ptr(), idx.data_ptr(), ctrl.data_ptr(), da1.data_ptr(),
                                          dw.data_ptr(), db.data_ptr(), batch, stream)
        assert rc == 0, lib.b200_last_error().decode()
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert free_after == free_before, (free_before, free_after)
        scale = dw_ref.abs().max().item()
        torch.testing.assert_close(dw, dw_ref, rtol=1e-3, atol=1e-4 * scale, msg=lambda m: f"call {call}: {m}")
        torch.testing.assert_close(db, db_ref, rtol=1e-3, atol=1e-4 * scale, msg=lambda m: f"call {call}: {m}")


def test_noalloc_conv1_wgrad_rejects_batches_beyond_its_scratch():
    lib = _lib()
    limit = lib.b200_conv1_wgrad_noalloc_max_batch()
    assert limit == 1024
    stream = torch.cuda.current_stream().cuda_stream
    dummy = torch.zeros(64, device="cuda")
    rc = lib.b200_conv1_wgrad_noalloc(dummy.data_ptr(), None, dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(),
                                      dummy.data_ptr(), limit + 1, stream)
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "exceeds" in msg and str(limit) in msg, msg
    torch.cuda.synchronize()  # nothing was launched, the device is still healthy


def test_noalloc_conv1_wgrad_rejects_null_pointers():
    lib = _lib()
    stream = torch.cuda.current_stream().cuda_stream
    d = torch.zeros(64, device="cuda")
    ok = [d.data_ptr()] * 6  # images, idx, ctrl, da1, dw, db
    for missing in (0, 2, 3, 4, 5):  # idx (position 1) may be NULL: sequential addressing
        args = list(ok)
        args[missing] = None
        rc = lib.b200_conv1_wgrad_noalloc(*args, 8, stream)
        assert rc != 0, f"NULL argument {missing} was accepted"
        assert "non-NULL" in lib.b200_last_error().decode()
    torch.cuda.synchronize()  # nothing was launched
