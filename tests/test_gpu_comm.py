"""Fused all-reduce (+ 1/W + Adadelta) kernels.

On a single GPU the W ranks are emulated inside ONE cooperative launch (blockIdx.y = rank) over
W cross-linked arenas, which exercises the same kernel code, flag protocol and fixed-order
reduction as the real multi-GPU path.  The real NVLink path is covered by the multi-process
tests at the bottom, which need >= 2 GPUs and are skipped otherwise.
"""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# (floats) the two real DDP bucket sizes of this model rounded to the flat layout, plus
# odd sizes: smaller than one CTA, not a multiple of the world size
SIZES = [4, 1024, 18816, 19200, 1181184, 1048580]


def _setup(world, n, seed=0):
    from pytorch_mnist_ddp_b200.parallel.fused_ddp import SymmetricArena
    arenas = SymmetricArena.create_emulated(world, 4 * n)
    g = torch.Generator(device="cuda").manual_seed(seed)
    for a in arenas:
        v = a.view()
        v[:n] = torch.randn(n, device="cuda", generator=g)          # params / output
        v[n:2 * n] = torch.randn(n, device="cuda", generator=g)     # grads
        v[2 * n:] = 0                                               # Adadelta state
    torch.cuda.synchronize()
    return arenas


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("n", SIZES)
def test_allreduce_average_emulated(world, mode, n):
    from pytorch_mnist_ddp_b200.parallel.fused_ddp import CommPlan
    arenas = _setup(world, n)
    grads = [a.view()[n:2 * n].clone() for a in arenas]
    plan = CommPlan.create_emulated(arenas, n, 0, 2 * n, 3 * n, n, channel=2)
    for rep in range(3):  # repeated calls reuse the flag words with growing tokens
        plan.run(mode, opt=False)
    torch.cuda.synchronize()
    acc = torch.zeros(n, device="cuda")
    for gq in grads:  # same fixed rank order, fp32: must match bit for bit
        acc += gq
    ref = acc * (1.0 / world)
    outs = [a.view()[:n] for a in arenas]
    for r in range(world):
        assert torch.equal(outs[r], outs[0]), f"rank {r} differs from rank 0"
    assert torch.equal(outs[0], ref)
    ref64 = sum(gq.double() for gq in grads) / world
    torch.testing.assert_close(outs[0].double(), ref64, rtol=1e-5, atol=1e-6)
    plan.destroy()
    for a in arenas:
        a.destroy()


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("mode", [0, 1])
def test_fused_allreduce_adadelta_emulated(world, mode):
    from pytorch_mnist_ddp_b200.ops import native_ops as ops
    from pytorch_mnist_ddp_b200.parallel.fused_ddp import CommPlan
    n = 19200
    arenas = _setup(world, n, seed=3)
    p0 = arenas[0].view()[:n].clone()
    for a in arenas:  # replicas start identical, as after the initial broadcast
        a.view()[:n] = p0
    ctrls = [ops.make_ctrl(step=0, lr=1.0) for _ in range(world)]
    plan = CommPlan.create_emulated(arenas, n, 0, 2 * n, 3 * n, n, channel=0, ctrls=ctrls)
    ref_p = p0.clone().requires_grad_(True)
    opt = torch.optim.Adadelta([ref_p], lr=1.0)
    g = torch.Generator(device="cuda").manual_seed(5)
    for step in range(5):
        grads = [torch.randn(n, device="cuda", generator=g) * 0.01 for _ in range(world)]
        for a, gq in zip(arenas, grads):
            a.view()[n:2 * n] = gq
        for c in ctrls:
            c.copy_(ops.make_ctrl(step=step, lr=1.0))
        torch.cuda.synchronize()
        plan.run(mode, opt=True, token=0)  # token from ctrl->step, as in the engine
        torch.cuda.synchronize()
        acc = torch.zeros(n, device="cuda")
        for gq in grads:
            acc += gq
        ref_p.grad = acc / world
        opt.step()
        for r in range(world):
            assert torch.equal(arenas[r].view()[:n], arenas[0].view()[:n]), f"step {step}: rank {r} drifted"
        torch.testing.assert_close(arenas[0].view()[:n], ref_p.detach(), rtol=1e-4, atol=1e-5)
    plan.destroy()
    for a in arenas:
        a.destroy()


CONV_BUCKET = 19456  # floats: conv2.weight 18432 | conv2.bias 256 | conv1.weight 512 | conv1.bias 256 (ParamLayout)


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("mode", [0, 1])
def test_conv_bucket_kernels_emulated(world, mode):
    """conv2_reduce_kernel (partial rows -> gradient -> all-reduce into every rank's reduced buffer) followed by
    conv_finalize_kernel (conv1 partial rows, one-shot exchange, Adadelta on the whole conv bucket, step advance):
    W emulated ranks on one GPU against torch.optim.Adadelta on the averaged gradient."""
    import ctypes
    from pytorch_mnist_ddp_b200 import _native as N
    from pytorch_mnist_ddp_b200.ops import native_ops as ops
    from pytorch_mnist_ddp_b200.parallel.fused_ddp import SymmetricArena
    n, np2, ng = CONV_BUCKET, 37, 13
    arenas = SymmetricArena.create_emulated(world, 5 * n)
    g = torch.Generator(device="cuda").manual_seed(11)
    p0 = torch.randn(n, device="cuda", generator=g)
    real = torch.zeros(n, dtype=torch.bool, device="cuda")  # padding floats carry no gradient and must stay put
    real[:18432 + 64] = True; real[18688:18688 + 288] = True; real[19200:19232] = True
    for a in arenas:
        v = a.view()
        v[:n] = p0
        v[n:2 * n] = float("nan")       # gradients: every float the kernels read must have been written by them
        v[2 * n:4 * n] = 0              # Adadelta state
        v[4 * n:] = float("nan")        # reduced gradients
    ctrls = [ops.make_ctrl(step=0, lr=1.0) for _ in range(world)]
    lib = N.lib()
    hs = (ctypes.c_void_p * world)(*[a.handle for a in arenas])
    cs = (ctypes.c_void_p * world)(*[c.data_ptr() for c in ctrls])
    plan = ctypes.c_void_p()
    N.check(lib.b200_comm_conv_plan_create_emulated(hs, world, n, cs, ctypes.byref(plan)))
    ref_p = p0.clone().requires_grad_(True)
    opt = torch.optim.Adadelta([ref_p], lr=1.0)
    counters = torch.zeros(8, dtype=torch.int32, device="cuda")
    for step in range(3):
        c2 = [torch.randn(np2, 18496, device="cuda", generator=g) * 0.01 for _ in range(world)]
        c1 = [torch.randn(ng, 320, device="cuda", generator=g) * 0.01 for _ in range(world)]
        torch.cuda.synchronize()
        p2 = (ctypes.c_void_p * world)(*[t.data_ptr() for t in c2])
        p1 = (ctypes.c_void_p * world)(*[t.data_ptr() for t in c1])
        N.check(lib.b200_comm_conv_bucket_emulated(plan, mode, p2, np2, p1, ng, 200, counters.data_ptr(),
                                                   torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        grad = torch.zeros(n, device="cuda", dtype=torch.float64)
        for r in range(world):
            grad[:18496] += c2[r].double().sum(0)
            grad[18688:18688 + 288] += c1[r][:, :288].double().sum(0)
            grad[19200:19232] += c1[r][:, 288:].double().sum(0)
        ref_p.grad = (grad / world).float()
        opt.step()
        for r in range(world):
            assert torch.equal(arenas[r].view()[:n], arenas[0].view()[:n]), f"step {step}: rank {r} drifted"
            c = ops.read_ctrl(ctrls[r])
            assert c["step"] == step + 1 and c["pos"] == 200 * (step + 1), c
        got = arenas[0].view()[:n]
        torch.testing.assert_close(got[real], ref_p.detach()[real], rtol=1e-4, atol=1e-5)
        assert torch.equal(got[~real], p0[~real])
    N.check(lib.b200_comm_plan_destroy(plan))
    for a in arenas:
        a.destroy()


def _torchrun(nproc, script_args, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", "29533"] + script_args
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("comm", ["oneshot", "twoshot", "nvls", "auto"])
def test_multi_gpu_ddp_equivalence(comm, world):
    """W ranks x batch b without dropout == 1 rank x batch W*b ('DDP equivalence'), replicas
    bit-identical, CUDA-graph replay with the communication kernels captured inside."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs >= {world} GPUs")
    res = _torchrun(world, ["tests/ddp_worker.py", "--comm", comm])
    try:  # keep every rank's output where a GPU job's artefacts are collected
        os.makedirs(os.path.join(ROOT, "gpurun_out", "ddp_logs"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "ddp_logs", f"ddp_worker_{comm}_{world}.log"), "w") as f:
            f.write(res.stdout + "\n---- stderr ----\n" + res.stderr)
    except OSError:
        pass
    print("\n".join(l for l in res.stdout.splitlines() if "DDP-vs-single" in l or "loss" in l))
    if comm == "nvls" and "NVLS unavailable" in res.stdout + res.stderr:
        pytest.skip("no multicast mapping on this host")
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "DDP_EQUIVALENCE_OK" in res.stdout, res.stdout[-3000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_hung_peer_is_reported_with_rank_and_peer():
    """a peer that stops arriving must end the job with a message naming who waited for whom."""
    env = dict(os.environ, B200_COMM_TIMEOUT_MS="2500")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29534", "tests/hung_peer_worker.py"]
    res = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=300, env=env)
    out = res.stdout + res.stderr
    line = [l for l in out.splitlines() if "HUNG_PEER_DIAGNOSTIC" in l]
    assert line, out[-3000:]
    print(line[0])
    assert "rank 0 of 2 waited" in line[0] and "for rank 1" in line[0] and "B200_COMM_TIMEOUT_MS" in line[0], line[0]
