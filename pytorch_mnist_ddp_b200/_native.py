"""Build and load libb200mnist.so (the sm_100a kernels and the C++ engine) through ctypes.

The library is plain CUDA C++ with a C ABI (include/b200mnist.h); it does not include torch
headers, so a full rebuild takes seconds with nvcc alone and cross-compiles on a machine
without a GPU.  The shared object is built IN-TREE (pytorch_mnist_ddp_b200/lib/) so that it
travels with the repository snapshot to the GPU box.

On a machine with a CUDA device a missing/unbuildable library is a hard error: there is no
silent fallback to PyTorch ops on the GPU path.
"""
from __future__ import annotations

import ctypes
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import List, Optional

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "csrc")
INCLUDE = os.path.join(ROOT, "include")
BUILD = os.path.join(ROOT, "build")
LIBDIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")
LIBPATH = os.path.join(LIBDIR, "libb200mnist.so")
STAMP = os.path.join(LIBDIR, "libb200mnist.stamp")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-I", INCLUDE,  # no --use_fast_math: fp32 parity with the reference
] + os.environ.get("B200_EXTRA_NVCC_FLAGS", "").split()  # e.g. -DB200_STAMPS (in-kernel timelines, debug only)


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isfile(cand) or cand == "nvcc"):
            return cand
    return "nvcc"


def sources() -> List[str]:
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _fingerprint() -> str:
    h = hashlib.sha256()
    for d in (CSRC, INCLUDE):
        for f in sorted(os.listdir(d)):
            if f.endswith((".cu", ".cuh", ".h")):
                h.update(f.encode())
                with open(os.path.join(d, f), "rb") as fh:
                    h.update(fh.read())
    # the flags without the checkout's absolute include path: the library built here must still count as fresh
    # when the tree is copied to another directory (the GPU box)
    h.update(" ".join(f for f in NVCC_FLAGS if f != INCLUDE).encode())
    return h.hexdigest()


def is_fresh() -> bool:
    if not (os.path.isfile(LIBPATH) and os.path.isfile(STAMP)):
        return False
    with open(STAMP) as f:
        return f.read().strip() == _fingerprint()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every csrc/*.cu for sm_100a and link the shared library.  Returns its path.

    Safe under torchrun: ranks serialise on a lock file, whoever gets it first builds (objects in a private
    directory, the library written under a temporary name and renamed into place), the others find it fresh."""
    if not force and is_fresh():
        return LIBPATH
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    import fcntl
    with open(os.path.join(LIBDIR, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and is_fresh():  # another process built it while this one waited
                return LIBPATH
            return _build_locked(verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(verbose: bool) -> str:
    nvcc = _nvcc()
    objdir = os.path.join(BUILD, f"obj.{os.getpid()}")
    os.makedirs(objdir, exist_ok=True)

    def compile_one(src: str) -> str:
        obj = os.path.join(objdir, src[:-3] + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{res.stdout}\n{res.stderr}")
        if verbose:
            sys.stderr.write(res.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        objs = list(ex.map(compile_one, sources()))
    tmp = LIBPATH + f".tmp.{os.getpid()}"
    link = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp, *objs]
    res = subprocess.run(link, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    os.replace(tmp, LIBPATH)  # atomic: nobody ever maps a half-written library
    with open(STAMP + ".tmp", "w") as f:
        f.write(_fingerprint())
    os.replace(STAMP + ".tmp", STAMP)
    import shutil
    shutil.rmtree(objdir, ignore_errors=True)
    return LIBPATH


_lib: Optional[ctypes.CDLL] = None

_vp, _i, _l, _f, _ull = ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_float, ctypes.c_ulonglong


class EngineConfig(ctypes.Structure):
    _fields_ = [
        ("device", _i), ("max_batch", _i), ("world", _i), ("rank", _i), ("seed", _ull), ("use_tc", _i),
        ("train_n", _i), ("test_n", _i), ("param_floats", _l), ("offsets", _l * 8),
        ("bucket_begin", _l * 2), ("bucket_end", _l * 2),
        ("params", _vp), ("grads", _vp * 2), ("sq_avg", _vp), ("acc_delta", _vp),
    ]


_SIGNATURES = {
    "b200_ctrl_set": [_vp, _ull, _ull, _i, _i, _f, _f, _vp],
    "b200_conv1_fwd": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "b200_conv1_wgrad": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "b200_reduce_partials": [_vp, _vp, _i, _i, _i, _vp],
    "b200_conv1_wgrad_noalloc": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "b200_conv1_wgrad_noalloc_max_batch": [],
    "b200_conv2_fwd_simt": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "b200_conv2_dgrad_simt": [_vp, _vp, _vp, _vp, _vp, _i, _vp],
    "b200_conv2_wgrad_simt": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "b200_sgemm": [_vp, _l, _l, _vp, _l, _l, _vp, _i, _i, _i, _i, _i, _vp, _f, _vp],
    "b200_tc_gemm": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _f, _vp],
    "b200_tc_gemm_ex": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _f, _i, _vp],
    "b200_conv2_fwd_tc": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "b200_conv12_fwd_tc": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "b200_conv2_dgrad_tc": [_vp, _vp, _vp, _vp, _vp, _i, _vp],
    "b200_conv2_wgrad_tc_parts": [_i],
    "b200_conv2_wgrad_tc": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "b200_head_fused": [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "b200_head_wgrad": [_vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp],
    "b200_engine_host_loss_ring": [_vp, ctypes.POINTER(ctypes.POINTER(ctypes.c_float))],
    "b200_eval_accumulate": [_vp, _vp, _i, _vp, _vp],
    "b200_adadelta": [_vp, _vp, _vp, _vp, _l, _vp, _f, _f, _i, _vp, _vp],
    "b200_symm_create": [_i, _i, ctypes.c_size_t, ctypes.POINTER(_vp)],
    "b200_symm_export": [_vp, _vp],
    "b200_symm_import": [_vp, _vp],
    "b200_symm_wrap": [_i, _i, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint64), ctypes.c_uint64, ctypes.POINTER(_vp)],
    "b200_symm_has_multicast": [_vp],
    "b200_symm_create_emulated": [_i, ctypes.c_size_t, ctypes.POINTER(_vp)],
    "b200_symm_peer": [_vp, _i, ctypes.POINTER(_vp)],
    "b200_symm_pull": [_vp, _i, ctypes.c_size_t, ctypes.c_size_t, _vp],
    "b200_symm_destroy": [_vp],
    "b200_comm_plan_create": [_vp, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t,
                              ctypes.c_size_t, _i, _vp, ctypes.POINTER(_vp)],
    "b200_comm_plan_create_emulated": [ctypes.POINTER(_vp), _i, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t,
                                       ctypes.c_size_t, ctypes.c_size_t, _i, ctypes.POINTER(_vp),
                                       ctypes.POINTER(_vp)],
    "b200_comm_plan_destroy": [_vp],
    "b200_comm_conv_plan_create_emulated": [ctypes.POINTER(_vp), _i, ctypes.c_size_t, ctypes.POINTER(_vp),
                                            ctypes.POINTER(_vp)],
    "b200_comm_conv_bucket_emulated": [_vp, _i, ctypes.POINTER(_vp), _i, ctypes.POINTER(_vp), _i, _i, _vp, _vp],
    "b200_comm_run": [_vp, _i, _i, ctypes.c_uint, _i, _vp],
    "b200_ddp_attach": [_vp, _vp, _i, _i],
    "b200_engine_create": [ctypes.POINTER(EngineConfig), ctypes.POINTER(_vp)],
    "b200_engine_destroy": [_vp],
    "b200_engine_buffer": [_vp, _i, ctypes.POINTER(_vp), ctypes.POINTER(ctypes.c_size_t)],
    "b200_engine_load_data": [_vp, _i, _vp, _vp, _i],
    "b200_engine_begin_epoch": [_vp, _vp, _i, _vp],
    "b200_engine_set_lr": [_vp, _f, _vp],
    "b200_engine_set_dropout": [_vp, _i, _vp],
    "b200_engine_set_defer_update": [_vp, _i],
    "b200_engine_apply_update": [_vp, _i, _vp],
    "b200_engine_train_step": [_vp, _i, _i, _vp],
    "b200_engine_train_step_host": [_vp, _vp, _vp, _i, _i, _vp],
    "b200_engine_train_step_gather": [_vp, _vp, _vp, _i, _vp, _i, _i, _vp],
    "b200_engine_eval_batch": [_vp, _i, _i, _i, _vp],
    "b200_engine_eval_reset": [_vp, _vp],
    "b200_engine_eval_result": [_vp, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_longlong), _vp],
    "b200_engine_read_loss": [_vp, _ull, ctypes.POINTER(_f), _vp],
    "b200_engine_step_count": [_vp, ctypes.POINTER(_ull)],
    "b200_engine_launches_per_step": [_vp, _i, ctypes.POINTER(_i)],
}


class NativeError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    """The loaded library (building it first if the in-tree copy is stale or absent)."""
    global _lib
    if _lib is not None:
        return _lib
    path = LIBPATH
    if not is_fresh():
        try:
            path = build()
        except Exception as e:
            # A library that does not match csrc/ would silently run other kernels than the ones in the tree: refuse,
            # unless the caller says that is what they want (a deployment box without nvcc).
            if os.path.isfile(LIBPATH) and os.environ.get("B200_ALLOW_STALE_LIB") == "1":
                sys.stderr.write(f"[b200] WARNING: using a libb200mnist.so that does not match the sources ({e})\n")
            else:
                raise NativeError("libb200mnist.so is missing or older than csrc/ and cannot be rebuilt here "
                                  f"(set B200_ALLOW_STALE_LIB=1 to load a stale library anyway): {e}") from e
    L = ctypes.CDLL(path)
    L.b200_last_error.restype = ctypes.c_char_p
    L.b200_comm_fault_message.restype = ctypes.c_char_p
    L.b200_launch_count.restype = ctypes.c_longlong
    for name, sig in _SIGNATURES.items():
        fn = getattr(L, name, None)
        if fn is None:
            continue
        fn.argtypes = sig
        fn.restype = _i
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        raise NativeError(f"libb200mnist error {rc}: {lib().b200_last_error().decode()}")


def launch_count() -> int:
    return int(lib().b200_launch_count())
