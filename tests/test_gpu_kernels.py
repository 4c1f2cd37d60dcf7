"""Numerics of every hand-written kernel against a plain PyTorch fp32 reference of the same op.

Oracle = torch ops on the GPU with TF32 disabled (the reference trains convs in TF32 and fc in
fp32; the CUDA-core kernels here are full fp32, the tcgen05 ones TF32 — tolerances
below reflect that).  Layout conversions are done here, never in the kernels under test.
"""

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

BATCHES = [1, 7, 64, 100, 200, 1000]          # SURVEY section 4; the two large ragged/full sizes of the
BIG_BATCHES = [3404, 4096]                     # batch-4096 configuration run on the tensor-core kernels only


@pytest.fixture(scope="module", autouse=True)
def _fp32_oracle():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield


@pytest.fixture(scope="module")
def ops():
    from pytorch_mnist_ddp_b200.ops import native_ops
    return native_ops


def _need_tc(tc):
    from pytorch_mnist_ddp_b200 import _native as N
    if tc and not N.lib().b200_has_tc():
        pytest.skip("tensor-core kernels not built")


def _net(seed=1):
    from pytorch_mnist_ddp_b200.models.net import Net
    torch.manual_seed(seed)
    return Net().cuda()


def _internal(net):
    from pytorch_mnist_ddp_b200.models.net import ParamLayout
    sd = net.state_dict()
    return {k: ParamLayout.to_internal(k, v.cpu()).cuda().contiguous() for k, v in sd.items()}


def _images(n, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, (n, 28, 28), dtype=torch.uint8, generator=g).cuda()


def _normalize(u8):
    return ((u8.float() / 255) - 0.1307) / 0.3081


@pytest.mark.parametrize("batch", BATCHES)
def test_conv1_fwd_gather_normalize(ops, batch):
    net, imgs = _net(), _images(batch + 100)
    w = _internal(net)
    idx = torch.randperm(batch + 100)[: batch + 5].to(torch.int32).cuda()
    ctrl = ops.make_ctrl(pos=3)
    a1 = ops.conv1_fwd(imgs, idx, ctrl, w["conv1.weight"], w["conv1.bias"], batch)
    x = _normalize(imgs[idx[3: 3 + batch].long()]).unsqueeze(1)
    ref = F.relu(net.conv1(x)).permute(0, 2, 3, 1)
    torch.testing.assert_close(a1, ref, rtol=1e-5, atol=1e-5)
    # sequential addressing (idx = NULL)
    a1s = ops.conv1_fwd(imgs, None, ctrl, w["conv1.weight"], w["conv1.bias"], batch)
    refs = F.relu(net.conv1(_normalize(imgs[3: 3 + batch]).unsqueeze(1))).permute(0, 2, 3, 1)
    torch.testing.assert_close(a1s, refs, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("batch,tc", [(b, t) for b in BATCHES for t in (False, True)] + [(b, True) for b in BIG_BATCHES])
def test_conv2_fwd_eval(ops, batch, tc):
    _need_tc(tc)
    net = _net()
    w = _internal(net)
    a1 = torch.rand(batch, 26, 26, 32, device="cuda") * 2 - 0.5
    ctrl = ops.make_ctrl(train=False)
    pooled, amax = ops.conv2_fwd(a1, w["conv2.weight"], w["conv2.bias"], ctrl, tc=tc)
    y = F.relu(net.conv2(a1.permute(0, 3, 1, 2)))
    ref = F.max_pool2d(y, 2).permute(0, 2, 3, 1)
    tol = dict(rtol=2e-2, atol=2e-2) if tc else dict(rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(pooled, ref, **tol)
    # the stored argmax must point at a maximal element of its window
    win = y.reshape(batch, 64, 12, 2, 12, 2).permute(0, 2, 4, 1, 3, 5).reshape(batch, 12, 12, 64, 4)
    picked = win.gather(-1, amax.long().unsqueeze(-1)).squeeze(-1)
    live = ref > 0
    torch.testing.assert_close(picked[live], ref[live], **tol)


@pytest.mark.parametrize("tc", [False, True])
def test_conv2_fwd_dropout(ops, tc):
    _need_tc(tc)
    net = _net()
    w = _internal(net)
    a1 = torch.rand(64, 26, 26, 32, device="cuda")
    ev, _ = ops.conv2_fwd(a1, w["conv2.weight"], w["conv2.bias"], ops.make_ctrl(train=False), tc=tc)
    tr, _ = ops.conv2_fwd(a1, w["conv2.weight"], w["conv2.bias"], ops.make_ctrl(train=True, step=5, seed=1), tc=tc)
    tr2, _ = ops.conv2_fwd(a1, w["conv2.weight"], w["conv2.bias"], ops.make_ctrl(train=True, step=6, seed=1), tc=tc)
    live = ev > 0
    kept = (tr != 0) & live
    frac = kept.sum().item() / live.sum().item()
    assert abs(frac - 0.75) < 0.01, frac
    torch.testing.assert_close(tr[kept], ev[kept] / 0.75, rtol=1e-5, atol=1e-6)
    assert (tr[~kept] == 0).all()
    assert ((tr2 != 0) & live).ne(kept).any(), "mask must change with the step counter"


def test_dropout_masks_identical_simt_vs_tc(ops):
    _need_tc(True)
    net = _net()
    w = _internal(net)
    a1 = torch.rand(16, 26, 26, 32, device="cuda") + 0.1
    c = ops.make_ctrl(train=True, step=11, seed=7)
    s, _ = ops.conv2_fwd(a1, w["conv2.weight"], w["conv2.bias"], c, tc=False)
    t, _ = ops.conv2_fwd(a1, w["conv2.weight"], w["conv2.bias"], c, tc=True)
    assert ((s != 0) == (t != 0)).float().mean().item() > 0.9999


LAYOUTS = [(False, False), (False, True), (True, True), (True, False)]


@pytest.mark.parametrize("a_mn,b_mn", LAYOUTS)
@pytest.mark.parametrize("shape", [(200, 128, 9216, 36), (128, 9216, 200, 1), (200, 9216, 128, 1), (7, 128, 9216, 4),
                                   (100, 9216, 128, 1), (128, 9216, 100, 1), (1000, 128, 9216, 2),
                                   # the batch-4096 configuration: full and ragged last batches (M-tail predication)
                                   (4096, 128, 9216, 4), (3404, 128, 9216, 4), (2656, 9216, 128, 1), (128, 9216, 3404, 1)])
@pytest.mark.parametrize("tc", [False, True, "fp32"])
def test_gemm(ops, a_mn, b_mn, shape, tc):
    _need_tc(tc)
    M, N, K, splits = shape
    if tc and ((a_mn and M % 32) or (b_mn and N % 32)):
        pytest.skip("MN-major tensor-core operands need the MN extent to be a multiple of 32")
    A = torch.randn(M, K, device="cuda")
    B = torch.randn(N, K, device="cuda")
    Aarg = A.t().contiguous() if a_mn else A
    Barg = B.t().contiguous() if b_mn else B
    D = ops.gemm(Aarg, Barg, a_mn, b_mn, M, N, K, splits=splits, tc=bool(tc), fp32=tc == "fp32").sum(0)
    ref = A @ B.t()
    if tc == "fp32":
        # 3xTF32 split-operand product: fp32-class.  Measured against an fp64 product, relative to the size of
        # the terms (|A|.|B|^T), like an fp32 SGEMM's own rounding error
        ref64 = (A.double() @ B.double().t())
        scale = (A.abs().double() @ B.abs().double().t())
        err = ((D.double() - ref64).abs() / scale).max().item()
        ref_err = ((ref.double() - ref64).abs() / scale).max().item()
        nrm = (D.double() - ref64).norm().item() / ref64.norm().item()
        print(f"3xTF32 {shape} a_mn={a_mn} b_mn={b_mn}: max err/(|A||B|) {err:.2e} (cuBLAS fp32: {ref_err:.2e}), norm-wise {nrm:.2e}")
        assert err < 1e-5, (err, ref_err)
        # Norm-wise against fp64.  The operand split itself is good to ~1e-6; what is left on long K chains is the
        # tensor core's accumulator, which truncates instead of rounding (measured: the error grows with the number
        # of accumulating MMAs of one output, ~K/8*3 of them).  The engine's GEMMs keep their chains short (fc1
        # forward: split-K 36, 256 deep; dgrad: K = 128; wgrad: K = batch).
        chain = K / splits
        assert nrm < (1e-5 if chain <= 512 else 2e-4), (nrm, chain)
    elif tc:
        torch.testing.assert_close(D, ref, rtol=2e-2, atol=2e-2 * (K ** 0.5))
    else:
        torch.testing.assert_close(D, ref, rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("M", [200, 3404])
@pytest.mark.parametrize("tc", [False, True, "fp32"])
def test_gemm_mask_epilogue(ops, tc, M):
    _need_tc(tc)
    N, K = 9216, 128
    A, B = torch.randn(M, K, device="cuda"), torch.randn(K, N, device="cuda")
    aux = torch.randn(M, N, device="cuda")
    D = ops.gemm(A, B, False, True, M, N, K, tc=bool(tc), aux=aux, scale=1 / 0.75, fp32=tc == "fp32")[0]
    ref = (A @ B) * (aux > 0) / 0.75
    tol = dict(rtol=1e-5, atol=3e-4) if tc == "fp32" else dict(rtol=2e-2, atol=0.3) if tc else dict(rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(D, ref, **tol)


def test_tf32_operand_rounding_probe(ops):
    """Documents how tcgen05 kind::tf32 reads an fp32 operand: 1 + 3*2^-12 lies between two TF32 values; truncation
    gives 1.0, round-to-nearest 1 + 2^-10.  The 3xTF32 path does not depend on the answer (it clears the low
    mantissa bits itself), this only records the hardware's behaviour in the test log."""
    _need_tc(True)
    M, N, K = 128, 128, 32
    A = torch.full((M, K), 1.0 + 3 * 2.0 ** -12, device="cuda")
    B = torch.ones(N, K, device="cuda")
    d = ops.gemm(A, B, False, False, M, N, K, tc=True)[0][0, 0].item() / K
    print(f"tcgen05 tf32 operand read of 1+3*2^-12 -> {d!r} ({'truncation' if d == 1.0 else 'rounding'})")
    assert d in (1.0, 1.0 + 2.0 ** -10)


@pytest.mark.parametrize("batch", BATCHES)
def test_head_eval_and_train(ops, batch):
    net = _net()
    w = _internal(net)
    splits = 3
    part = torch.randn(splits, batch, 128, device="cuda")
    labels = torch.randint(0, 10, (batch + 4,), dtype=torch.uint8, device="cuda")
    # eval
    ctrl = ops.make_ctrl(train=False, pos=4)
    h, _, _, loss_i, correct, logp = ops.head_fused(part, w["fc1.bias"], w["fc2.weight"], w["fc2.bias"], labels, None,
                                                   ctrl, train=False)
    z = F.relu(part.sum(0) + net.fc1.bias)
    ref_logp = F.log_softmax(net.fc2(z), dim=1)
    y = labels[4: 4 + batch].long()
    torch.testing.assert_close(h, z, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(logp, ref_logp, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(loss_i, F.nll_loss(ref_logp, y, reduction="none"), rtol=1e-4, atol=1e-4)
    assert (correct.bool() == ref_logp.argmax(1).eq(y)).all()
    # train with dropout: recover the mask from h, then check the backward seed
    ctrl = ops.make_ctrl(train=True, pos=4, step=3)
    h, dlogits, dh, loss_i, _, _ = ops.head_fused(part, w["fc1.bias"], w["fc2.weight"], w["fc2.bias"], labels, None,
                                                  ctrl, train=True)
    keep = (h != 0) | (z == 0)
    if batch >= 64:
        assert abs(keep[z > 0].float().mean().item() - 0.5) < 0.05
    hz = (z * keep * 2).detach().requires_grad_(True)
    loss = F.nll_loss(F.log_softmax(net.fc2(hz), dim=1), y)
    loss.backward()
    torch.testing.assert_close(h, hz.detach(), rtol=1e-5, atol=1e-5)
    ref_dh = hz.grad * (hz > 0) * 2
    torch.testing.assert_close(dh, ref_dh, rtol=1e-4, atol=1e-6)
    dw2, db2, db1, mean_loss = ops.head_wgrad(h, dlogits, dh, loss_i, ctrl)
    torch.testing.assert_close(mean_loss, loss.detach(), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(dw2, net.fc2.weight.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(db2, net.fc2.bias.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(db1, ref_dh.sum(0), rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("batch,tc", [(b, t) for b in (1, 7, 64, 200, 1000) for t in (False, True)] + [(4096, True)])
def test_conv2_backward(ops, batch, tc):
    _need_tc(tc)
    net = _net()
    w = _internal(net)
    a1 = (torch.rand(batch, 26, 26, 32, device="cuda") - 0.3).clamp_min(0)  # post-ReLU input with zeros
    pooled, amax = ops.conv2_fwd(a1, w["conv2.weight"], w["conv2.bias"], ops.make_ctrl(train=False))
    dp = torch.randn(batch, 12, 12, 64, device="cuda") * (pooled > 0)  # masked as the fc1 dgrad epilogue does
    x = a1.permute(0, 3, 1, 2).detach().requires_grad_(True)
    net.zero_grad()
    out = F.max_pool2d(F.relu(net.conv2(x)), 2)
    out.backward(dp.permute(0, 3, 1, 2))
    from pytorch_mnist_ddp_b200.models.net import ParamLayout
    ref_dw = ParamLayout.to_internal("conv2.weight", net.conv2.weight.grad.cpu()).cuda().view(9, 64, 32)
    ref_da1 = x.grad.permute(0, 2, 3, 1) * (a1 > 0)
    dw, db = ops.conv2_wgrad(dp, amax, a1, tc=tc)
    da1 = ops.conv2_dgrad(dp, amax, w["conv2.weight"], a1, tc=tc)
    # The oracle recomputes the forward with cuDNN, so a near-tie inside a 2x2 window can pick a
    # different argmax than the kernel under test did: compare norm-wise, not element-wise.
    def rel(a, b):
        return (a - b).norm().item() / max(b.norm().item(), 1e-12)
    lim = 3e-2 if tc else 2e-3
    assert rel(dw, ref_dw) < lim, rel(dw, ref_dw)
    assert rel(db, net.conv2.bias.grad) < 2e-3, rel(db, net.conv2.bias.grad)
    assert rel(da1, ref_da1) < lim, rel(da1, ref_da1)
    assert (da1[a1 <= 0] == 0).all(), "ReLU mask of conv1 must be applied"


@pytest.mark.parametrize("batch", [1, 7, 64, 200, 1000])
def test_conv2_backward_tc_close_to_fp32_kernels(ops, batch):
    """Same (dp, argmax) into the CUDA-core and the tensor-core backward: no max-pool tie ambiguity here, so the two
    must agree element-wise to TF32 accuracy, and the bias gradient (plain fp32 sums of dp) almost exactly."""
    _need_tc(True)
    net = _net()
    w = _internal(net)
    a1 = (torch.rand(batch, 26, 26, 32, device="cuda") - 0.3).clamp_min(0)
    pooled, amax = ops.conv2_fwd(a1, w["conv2.weight"], w["conv2.bias"], ops.make_ctrl(train=False), tc=True)
    dp = torch.randn(batch, 12, 12, 64, device="cuda") * (pooled > 0)
    dw0, db0 = ops.conv2_wgrad(dp, amax, a1, tc=False)
    dw1, db1 = ops.conv2_wgrad(dp, amax, a1, tc=True)
    da0 = ops.conv2_dgrad(dp, amax, w["conv2.weight"], a1, tc=False)
    da1 = ops.conv2_dgrad(dp, amax, w["conv2.weight"], a1, tc=True)
    rel = lambda a, b: (a - b).norm().item() / max(b.norm().item(), 1e-12)
    print(f"batch {batch}: TC vs fp32 kernels, norm-wise: dw {rel(dw1, dw0):.2e} db {rel(db1, db0):.2e} da1 {rel(da1, da0):.2e}")
    torch.testing.assert_close(db1, db0, rtol=1e-4, atol=1e-4 * db0.abs().max().item())
    torch.testing.assert_close(dw1, dw0, rtol=2e-2, atol=2e-3 * dw0.abs().max().item())
    torch.testing.assert_close(da1, da0, rtol=2e-2, atol=2e-3 * da0.abs().max().item())
    assert rel(dw1, dw0) < 3e-3 and rel(da1, da0) < 3e-3


@pytest.mark.parametrize("train", [False, True])
@pytest.mark.parametrize("batch", BATCHES + BIG_BATCHES)
def test_conv12_fwd_equals_conv1_then_conv2(ops, batch, train):
    """The fused conv1+conv2 forward must reproduce conv1_fwd followed by the tensor-core conv2 forward bit for bit
    (same FMA order in conv1, same MMAs and epilogue in conv2, same Philox offsets)."""
    _need_tc(True)
    net, imgs = _net(), _images(batch + 50)
    w = _internal(net)
    idx = torch.randperm(batch + 50)[: batch + 5].to(torch.int32).cuda()
    ctrl = ops.make_ctrl(pos=3, step=5, seed=11, train=train)
    a1_ref = ops.conv1_fwd(imgs, idx, ctrl, w["conv1.weight"], w["conv1.bias"], batch)
    pooled_ref, amax_ref = ops.conv2_fwd(a1_ref, w["conv2.weight"], w["conv2.bias"], ctrl, tc=True)
    a1, pooled, amax = ops.conv12_fwd(imgs, idx, ctrl, w["conv1.weight"], w["conv1.bias"], w["conv2.weight"],
                                      w["conv2.bias"], batch)
    torch.cuda.synchronize()
    assert torch.equal(a1, a1_ref), (a1 - a1_ref).abs().max().item()
    assert torch.equal(pooled, pooled_ref), (pooled - pooled_ref).abs().max().item()
    assert torch.equal(amax, amax_ref)
    # sequential addressing (idx = NULL), as the evaluation pass uses it
    a1s, pooleds, _ = ops.conv12_fwd(imgs, None, ctrl, w["conv1.weight"], w["conv1.bias"], w["conv2.weight"],
                                     w["conv2.bias"], batch)
    a1s_ref = ops.conv1_fwd(imgs, None, ctrl, w["conv1.weight"], w["conv1.bias"], batch)
    assert torch.equal(a1s, a1s_ref)


@pytest.mark.parametrize("batch", [1, 7, 64, 200, 1000, 4096])
def test_conv1_wgrad(ops, batch):
    net, imgs = _net(), _images(batch + 36)
    idx = torch.randperm(batch + 36)[:batch].to(torch.int32).cuda()
    ctrl = ops.make_ctrl()
    da1 = torch.randn(batch, 26, 26, 32, device="cuda")
    dw, db = ops.conv1_wgrad(imgs, idx, ctrl, da1)
    x = _normalize(imgs[idx.long()]).unsqueeze(1)
    net.zero_grad()
    net.conv1(x).backward(da1.permute(0, 3, 1, 2))
    ref = net.conv1.weight.grad.reshape(32, 9).t()
    s = ref.abs().max().item()
    torch.testing.assert_close(dw, ref, rtol=1e-3, atol=1e-4 * s)
    torch.testing.assert_close(db, net.conv1.bias.grad, rtol=1e-3, atol=1e-4 * s)


def test_adadelta_matches_torch(ops):
    n = 1024 * 37
    p = torch.randn(n, device="cuda")
    ref_p = p.clone().requires_grad_(True)
    opt = torch.optim.Adadelta([ref_p], lr=1.0)
    v, u = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    lr = 1.0
    for step in range(100):
        g = torch.randn(n, device="cuda") * (0.1 if step % 3 else 1e-4)
        ctrl = ops.make_ctrl(lr=lr)
        ops.adadelta(p, g, v, u, ctrl)
        ref_p.grad = g.clone()
        for grp in opt.param_groups:
            grp["lr"] = lr
        opt.step()
        if step % 20 == 19:
            lr *= 0.7
    torch.testing.assert_close(p, ref_p.detach(), rtol=1e-4, atol=1e-5)
    st = opt.state[ref_p]
    torch.testing.assert_close(v, st["square_avg"], rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(u, st["acc_delta"], rtol=1e-4, atol=1e-7)


def test_adadelta_advances_ctrl(ops):
    n = 4096
    p, g = torch.zeros(n, device="cuda"), torch.ones(n, device="cuda")
    v, u = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    ctrl = ops.make_ctrl(step=5, pos=100)
    ops.adadelta(p, g, v, u, ctrl, advance=200)
    c = ops.read_ctrl(ctrl)
    assert c["step"] == 6 and c["pos"] == 300
