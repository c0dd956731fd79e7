"""Numerics of every sm_100a kernel against a plain PyTorch fp32 restatement of the same op
(SURVEY.md §4.2 "Unit, GPU").  torch.nn.functional is the oracle only — it is never on the
product path.  Tolerances are per-op: the tensor-core convolutions take bf16 operands with fp32
accumulation (the oracle is fed the same bf16-rounded operands, so the remaining difference is
accumulation order and the bf16 rounding of the stored result); everything else is fp32 math.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

H = W = 28


@pytest.fixture(scope="module")
def nat():
    from ddp_b200.ops import native
    native.lib()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return native


def _lay():
    from ddp_b200.ops import layout
    return layout


def _bf(x):
    return x.to(torch.bfloat16).float()


def _report(name, got, want, atol, rtol):
    err = (got - want).abs()
    tol = atol + rtol * want.abs()
    bad = err > tol
    if bad.any():
        idx = bad.nonzero()[:8].tolist()
        msg = (f"{name}: {int(bad.sum())}/{bad.numel()} mismatches, max abs err {err.max().item():.4g}, "
               f"max |want| {want.abs().max().item():.4g}; first bad idx {idx}; "
               f"got {[got[tuple(i)].item() for i in idx[:4]]} want {[want[tuple(i)].item() for i in idx[:4]]}")
        pytest.fail(msg)


@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (5, 16)])
def test_conv1_fwd(nat, batch, cap):
    torch.manual_seed(0)
    x = torch.rand(batch, 1, H, W, device="cuda")
    w = torch.randn(32, 1, 3, 3, device="cuda") * 0.3
    b = torch.randn(32, device="cuda") * 0.1
    out = nat.alloc_planes(32, cap, H, W)
    nat.conv1_fwd(x, w, b, out, batch, cap, H, W)
    torch.cuda.synchronize()
    got = _lay().planes_to_nchw(out, batch, 32, H, W)
    want = F.relu(F.conv2d(x, w, b, padding=1))
    _report("conv1_fwd", got, want, 1e-2, 1e-2)
    assert _lay().border_rows_are_zero(out, batch, H, W)


# 2500 images = 17.6K tiles = ~60 tiles per CTA: long enough for the two producer groups to drift apart
@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (5, 16), (300, 300), (2500, 2500)])
def test_conv1_fwd_tcgen05_tf32(nat, batch, cap):
    torch.manual_seed(10)
    x = torch.rand(batch, 1, H, W, device="cuda")
    w = torch.randn(32, 1, 3, 3, device="cuda") * 0.3
    b = torch.randn(32, device="cuda") * 0.1
    out = nat.alloc_planes(32, cap, H, W)
    nat.conv1_fwd_tc(x, w, b, out, batch, cap, H, W)
    torch.cuda.synchronize()
    got = _lay().planes_to_nchw(out, batch, 32, H, W)
    want = F.relu(F.conv2d(x, w, b, padding=1))
    _report("conv1_fwd_tc", got, want, 1e-2, 1e-2)   # tf32 operands + bf16 stored result
    assert _lay().border_rows_are_zero(out, batch, H, W)


@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (37, 64), (300, 300), (2500, 2500)])
def test_conv1_wgrad_tcgen05(nat, batch, cap):
    torch.manual_seed(11)
    x = torch.rand(batch, 1, H, W, device="cuda")
    dact = torch.randn(batch, 32, H, W, device="cuda") * 0.1
    dact_p = _lay().nchw_to_planes(dact, cap)
    out = torch.zeros(32 * 9 + 32, device="cuda")
    scr = nat.scratch(nat.conv1_wgrad_tc_scratch_bytes(1, 32))
    nat.conv1_wgrad_tc(dact_p, x, out, scr, 1, 32, batch, cap, H, W)
    torch.cuda.synchronize()
    wref = torch.zeros(32, 1, 3, 3, device="cuda", requires_grad=True)
    bref = torch.zeros(32, device="cuda", requires_grad=True)
    F.conv2d(x, wref, bref, padding=1).backward(_bf(dact))
    _report("conv1_wgrad_tc dW", out[:288].view(32, 1, 3, 3), wref.grad, 1e-3 * wref.grad.abs().max().item(), 1e-4)
    _report("conv1_wgrad_tc db", out[288:], bref.grad, 1e-3 * bref.grad.abs().max().item(), 1e-4)


@pytest.mark.parametrize("batch,cap", [(2, 2), (64, 64), (37, 64), (300, 300)])
def test_conv3x3_fwd_tcgen05(nat, batch, cap):
    torch.manual_seed(1)
    a = torch.relu(torch.randn(batch, 32, H, W, device="cuda"))
    w = torch.randn(64, 32, 3, 3, device="cuda") * 0.06
    b = torch.randn(64, device="cuda") * 0.1
    inp = _lay().nchw_to_planes(a, cap)
    out = nat.alloc_planes(64, cap, H, W)
    nat.conv3x3_fwd(inp, w, b, out, batch, cap, H, W)
    torch.cuda.synchronize()
    got = _lay().planes_to_nchw(out, batch, 64, H, W)
    want = F.relu(F.conv2d(_bf(a), _bf(w), b, padding=1))
    _report("conv3x3_fwd", got, want, 2e-2, 1e-2)
    assert _lay().border_rows_are_zero(out, batch, H, W)


@pytest.mark.parametrize("batch,cap", [(2, 2), (64, 64), (37, 64)])
def test_conv3x3_dgrad_tcgen05(nat, batch, cap):
    torch.manual_seed(2)
    act1 = torch.relu(torch.randn(batch, 32, H, W, device="cuda"))
    dout = torch.randn(batch, 64, H, W, device="cuda") * 0.1
    w = torch.randn(64, 32, 3, 3, device="cuda") * 0.06
    dout_p = _lay().nchw_to_planes(dout, cap)
    mask_p = _lay().nchw_to_planes(act1, cap)
    din = nat.alloc_planes(32, cap, H, W)
    nat.conv3x3_dgrad(dout_p, w, mask_p, din, batch, cap, H, W)
    torch.cuda.synchronize()
    got = _lay().planes_to_nchw(din, batch, 32, H, W)
    want = F.conv_transpose2d(_bf(dout), _bf(w), padding=1) * (_bf(act1) > 0)
    _report("conv3x3_dgrad", got, want, 2e-2, 1e-2)
    assert _lay().border_rows_are_zero(din, batch, H, W)


@pytest.mark.parametrize("batch,cap", [(2, 2), (64, 64), (37, 64), (300, 300)])
def test_conv3x3_wgrad_tcgen05(nat, batch, cap):
    torch.manual_seed(3)
    act1 = torch.relu(torch.randn(batch, 32, H, W, device="cuda"))
    dout = torch.randn(batch, 64, H, W, device="cuda") * 0.1
    dout_p = _lay().nchw_to_planes(dout, cap)
    in_p = _lay().nchw_to_planes(act1, cap)
    dw = torch.zeros(64, 32, 3, 3, device="cuda")
    db = torch.zeros(64, device="cuda")
    scr = nat.scratch(nat.conv3x3_wgrad_scratch_bytes(32, 64))
    nat.conv3x3_wgrad(dout_p, in_p, dw, db, scr, 32, 64, batch, cap, H, W)
    torch.cuda.synchronize()
    a = _bf(act1).requires_grad_(False)
    wref = torch.zeros(64, 32, 3, 3, device="cuda", requires_grad=True)
    bref = torch.zeros(64, device="cuda", requires_grad=True)
    y = F.conv2d(a, wref, bref, padding=1)
    y.backward(_bf(dout))
    scale = wref.grad.abs().max().item()
    _report("conv3x3_wgrad dW", dw, wref.grad, 2e-3 * scale, 2e-3)
    _report("conv3x3_wgrad db", db, bref.grad, 2e-3 * bref.grad.abs().max().item(), 2e-3)


@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (37, 64)])
def test_conv1_wgrad(nat, batch, cap):
    torch.manual_seed(4)
    x = torch.rand(batch, 1, H, W, device="cuda")
    dact = torch.randn(batch, 32, H, W, device="cuda") * 0.1
    dact_p = _lay().nchw_to_planes(dact, cap)
    out = torch.zeros(32 * 9 + 32, device="cuda")
    scr = nat.scratch(nat.conv1_wgrad_scratch_bytes(1, 32, cap, H, W))
    nat.conv1_wgrad(dact_p, x, out, scr, 1, 32, batch, cap, H, W)
    torch.cuda.synchronize()
    wref = torch.zeros(32, 1, 3, 3, device="cuda", requires_grad=True)
    bref = torch.zeros(32, device="cuda", requires_grad=True)
    F.conv2d(x, wref, bref, padding=1).backward(_bf(dact))
    _report("conv1_wgrad dW", out[:288].view(32, 1, 3, 3), wref.grad, 1e-3 * wref.grad.abs().max().item(), 1e-4)
    _report("conv1_wgrad db", out[288:], bref.grad, 1e-3 * bref.grad.abs().max().item(), 1e-4)


@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (37, 64), (130, 130)])
def test_linear_fwd_loss_bwd(nat, batch, cap):
    torch.manual_seed(5)
    C = 64
    K = C * H * W
    act = torch.relu(torch.randn(batch, C, H, W, device="cuda"))
    Wfl = torch.randn(10, K, device="cuda") * 0.01
    bfl = torch.randn(10, device="cuda") * 0.1
    y = torch.randint(0, 10, (batch,), device="cuda")
    act_p = _lay().nchw_to_planes(act, cap)

    part = nat.scratch(nat.linear_fwd_scratch_bytes(10, C, cap, H, W))
    S = nat.linear_fwd(act_p, Wfl, part, 10, C, batch, cap, H, W)
    logits = torch.zeros(batch, 10, device="cuda")
    dlogits = torch.zeros(batch, 10, device="cuda")
    dbias = torch.zeros(10, device="cuda")
    loss = torch.zeros(1, device="cuda")
    nat.softmax_xent(part, S, bfl, y, logits, dlogits, dbias, loss, 10, batch)

    dact = nat.alloc_planes(C, cap, H, W)
    lay = _lay()
    first = lay.GUARD + batch * 900
    last = lay.GUARD + lay.tiles_for(batch, H, W) * 128
    dact[:, first:last] = 7.0  # stale rows of an earlier, larger batch: the kernel must zero them
    dW = torch.zeros(10, K, device="cuda")
    scr = nat.scratch(nat.linear_bwd_scratch_bytes(10, C, cap, H, W))
    nat.linear_bwd(act_p, dlogits, Wfl, dact, dW, scr, 10, C, batch, cap, H, W)
    torch.cuda.synchronize()

    a = _bf(act).reshape(batch, K).requires_grad_(True)
    wr = Wfl.clone().requires_grad_(True)
    br = bfl.clone().requires_grad_(True)
    z = a @ wr.t() + br
    lref = F.cross_entropy(z, y)
    lref.backward()
    _report("logits", logits, z.detach(), 1e-3, 1e-4)
    assert abs(loss.item() - lref.item()) < 1e-4 * max(1.0, abs(lref.item()))
    dz = (torch.softmax(z.detach(), 1) - F.one_hot(y, 10).float()) / batch
    _report("dlogits", dlogits, dz, 1e-6, 1e-4)
    _report("dbias", dbias, br.grad, 1e-6, 1e-4)
    _report("dW_fl", dW, wr.grad, 1e-4 * wr.grad.abs().max().item() + 1e-7, 1e-3)
    got_dact = _lay().planes_to_nchw(dact, batch, C, H, W)
    want_dact = (a.grad.reshape(batch, C, H, W)) * (_bf(act) > 0)
    _report("dact2", got_dact, want_dact, 1e-2 * want_dact.abs().max().item(), 1e-2)
    # rows between the end of the batch and the end of its last tile must be zero
    assert (dact[:, first:last].float() == 0).all()


@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (37, 64), (130, 130), (300, 300)])
def test_linear_fwd_tcgen05(nat, batch, cap):
    torch.manual_seed(12)
    C = 64
    K = C * H * W
    act = torch.relu(torch.randn(batch, C, H, W, device="cuda"))
    Wfl = torch.randn(10, K, device="cuda") * 0.01
    bfl = torch.randn(10, device="cuda") * 0.1
    y = torch.randint(0, 10, (batch,), device="cuda")
    act_p = _lay().nchw_to_planes(act, cap)
    packed = torch.empty(nat.linear_pack_bytes(10, C, H, W), dtype=torch.uint8, device="cuda")
    nat.linear_pack(Wfl, packed, 10, C, H, W)
    part = nat.scratch(nat.linear_fwd_tc_scratch_bytes(10, C, cap, H, W))
    S = nat.linear_fwd_tc(act_p, packed, part, 10, C, batch, cap, H, W)
    logits = torch.zeros(batch, 10, device="cuda")
    dlogits = torch.zeros(batch, 10, device="cuda")
    dbias = torch.zeros(10, device="cuda")
    loss = torch.zeros(1, device="cuda")
    nat.softmax_xent(part, S, bfl, y, logits, dlogits, dbias, loss, 10, batch)
    torch.cuda.synchronize()
    z = _bf(act).reshape(batch, K) @ Wfl.t() + bfl
    _report("logits (tensor cores, hi+lo weights)", logits, z, 2e-3, 1e-3)
    lref = F.cross_entropy(z, y)
    assert abs(loss.item() - lref.item()) < 1e-3 * max(1.0, abs(lref.item()))
    dz = (torch.softmax(z, 1) - F.one_hot(y, 10).float()) / batch
    _report("dlogits", dlogits, dz, 1e-5, 1e-3)
    _report("dbias", dbias, dz.sum(0), 1e-5, 1e-3)


@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (37, 64), (130, 130), (300, 300), (1100, 1100)])
def test_linear_bwd_tcgen05(nat, batch, cap):
    torch.manual_seed(13)
    C = 64
    K = C * H * W
    act = torch.relu(torch.randn(batch, C, H, W, device="cuda"))
    Wfl = torch.randn(10, K, device="cuda") * 0.01
    bfl = torch.randn(10, device="cuda") * 0.1
    y = torch.randint(0, 10, (batch,), device="cuda")
    act_p = _lay().nchw_to_planes(act, cap)
    packed = torch.empty(nat.linear_pack_bytes(10, C, H, W), dtype=torch.uint8, device="cuda")
    nat.linear_pack(Wfl, packed, 10, C, H, W)
    # dlogits (fp32 and the packed bf16 hi/lo tiles) come from the loss kernel
    part = nat.scratch(nat.linear_fwd_tc_scratch_bytes(10, C, cap, H, W))
    S = nat.linear_fwd_tc(act_p, packed, part, 10, C, batch, cap, H, W)
    logits = torch.zeros(batch, 10, device="cuda")
    dl = torch.zeros(batch, 10, device="cuda")
    dbias = torch.zeros(10, device="cuda")
    loss = torch.zeros(1, device="cuda")
    dlp = torch.full((nat.dl_packed_bytes(cap),), 0x7f, dtype=torch.uint8, device="cuda")  # garbage
    nat.softmax_xent(part, S, bfl, y, logits, dl, dbias, loss, 10, batch, dl_packed=dlp)
    dact = nat.alloc_planes(C, cap, H, W)
    lay = _lay()
    first = lay.GUARD + batch * 900
    last = lay.GUARD + lay.tiles_for(batch, H, W) * 128
    dact[:, first:last] = 7.0
    dW = torch.full((10, K), float("nan"), device="cuda")
    nat.linear_bwd_tc(act_p, dlp, packed, dact, dW, 10, C, batch, cap, H, W)
    torch.cuda.synchronize()
    a = _bf(act).reshape(batch, K)
    want_dW = dl.t() @ a
    _report("dW_fl (tensor cores)", dW, want_dW, 2e-3 * want_dW.abs().max().item() + 1e-8, 2e-3)
    want_dact = ((dl @ Wfl).reshape(batch, C, H, W)) * (_bf(act) > 0)
    got_dact = lay.planes_to_nchw(dact, batch, C, H, W)
    _report("dact2 (tensor cores)", got_dact, want_dact, 1e-2 * want_dact.abs().max().item(), 1e-2)
    assert (dact[:, first:last].float() == 0).all()
    assert lay.border_rows_are_zero(dact, batch, H, W)


def test_sgd(nat):
    torch.manual_seed(6)
    p = torch.randn(520586, device="cuda")
    g = torch.randn(520586, device="cuda")
    lr = torch.tensor([0.01], device="cuda")
    want = p - 0.01 * g
    nat.sgd(p, g, lr)
    torch.cuda.synchronize()
    assert torch.allclose(p, want, atol=1e-6, rtol=1e-6)


def test_gather_u8(nat):
    torch.manual_seed(7)
    imgs = torch.randint(0, 256, (1000, 28, 28), dtype=torch.uint8, device="cuda")
    labels = torch.randint(0, 10, (1000,), device="cuda")
    idx = torch.randperm(1000, device="cuda")[:64]
    x = torch.empty(64, 1, 28, 28, device="cuda")
    y = torch.empty(64, dtype=torch.int64, device="cuda")
    nat.gather_u8(imgs, labels, idx, x, y)
    torch.cuda.synchronize()
    # ToTensor semantics = true division on the CPU (CUDA's scalar div multiplies by 1/255)
    assert torch.equal(x.cpu(), imgs[idx].cpu().float().div(255).unsqueeze(1))
    assert torch.equal(y, labels[idx])


@pytest.mark.parametrize("algo_name", ["push", "pull"])
@pytest.mark.parametrize("nparts,n", [(1, 320), (37, 320), (148, 320), (5, 4096)])
def test_reduce_allreduce_sgd_single_rank(nat, algo_name, nparts, n):
    """Fused split-K reduce + all-reduce + SGD kernel at world size 1 (the exchange degenerates to
    this rank's own staging lines / flags); the multi-rank case is tests/test_multi_gpu.py."""
    from ddp_b200.parallel.comm import ALGO_ONESHOT, ALGO_PUSH, SymmetricComm
    algo = ALGO_PUSH if algo_name == "push" else ALGO_ONESHOT
    torch.manual_seed(8)
    region = (n * 4 + 255) // 256 * 256
    comm = SymmetricComm(0, 1, 2 * region)
    try:
        p, g = comm.tensor(0, n), comm.tensor(region, n)
        lr = torch.tensor([0.05], device="cuda")
        p.copy_(torch.randn(n, device="cuda"))
        for it in range(3):  # consecutive epochs reuse the staging lines
            parts = torch.randn(nparts, n, device="cuda")
            want = p - 0.05 * parts.sum(0)
            comm.reduce_allreduce_sgd(parts, nparts, n, region, 0, lr, algo=algo, slot=2)
            torch.cuda.synchronize()
            assert torch.allclose(g, parts.sum(0), atol=1e-4, rtol=1e-5)
            assert torch.allclose(p, want, atol=1e-5, rtol=1e-5), (it, (p - want).abs().max().item())
    finally:
        comm.close()


def test_reduce_allreduce_sgd_conv3x3_layout_single_rank(nat):
    """layout = conv3x3 weight-gradient accumulator images: the fused kernel must apply the same
    permutation as wgrad_reduce_kernel (checked against ddpb_conv3x3_wgrad on real data)."""
    from ddp_b200.parallel.comm import (ALGO_PUSH, CONV3X3_WGRAD_IMAGE, LAYOUT_CONV3X3_WGRAD,
                                        SymmetricComm, conv3x3_wgrad_image_dest)
    lay = _lay()
    torch.manual_seed(9)
    batch = 37
    a1 = lay.nchw_to_planes(torch.relu(torch.randn(batch, 32, H, W, device="cuda")), batch)
    d2 = lay.nchw_to_planes(torch.randn(batch, 64, H, W, device="cuda") * 0.1, batch)
    dw = torch.zeros(64, 32, 3, 3, device="cuda")
    db = torch.zeros(64, device="cuda")
    scr = nat.scratch(nat.conv3x3_wgrad_scratch_bytes(32, 64))
    nat.conv3x3_wgrad(d2, a1, dw, db, scr, 32, 64, batch, batch, H, W)
    n_dst = 64 * 32 * 9 + 64
    region = (n_dst * 4 + 255) // 256 * 256
    comm = SymmetricComm(0, 1, 2 * region)
    try:
        p, g = comm.tensor(0, n_dst), comm.tensor(region, n_dst)
        lr = torch.tensor([0.5], device="cuda")
        for it in range(2):
            p.copy_(torch.randn(n_dst, device="cuda"))
            start = p.clone()
            nparts = nat.conv3x3_wgrad_partials(d2, a1, scr, 32, 64, batch, batch, H, W)
            comm.reduce_allreduce_sgd(scr, nparts, CONV3X3_WGRAD_IMAGE, region, 0, lr, algo=ALGO_PUSH,
                                      layout=LAYOUT_CONV3X3_WGRAD, slot=5)
            torch.cuda.synchronize()
            want_g = torch.cat([dw.reshape(-1), db])
            assert torch.allclose(g, want_g, atol=1e-5, rtol=1e-5), (g - want_g).abs().max().item()
            assert torch.allclose(p, start - 0.5 * want_g, atol=1e-5, rtol=1e-5)
        # the Python restatement of the index map agrees with the kernel
        dest = conv3x3_wgrad_image_dest().cuda()
        img = scr.view(torch.float32)[: nparts * CONV3X3_WGRAD_IMAGE].view(nparts, -1).sum(0)
        ref = torch.zeros(n_dst, device="cuda")
        ref[dest[dest >= 0]] = img[dest >= 0]
        assert torch.allclose(ref, g, atol=1e-4, rtol=1e-4)
    finally:
        comm.close()
