"""Numerics of the reference-precision ("bf16x3") kernels — split planes x = hi + lo with three MMAs
per K step, fp32 linear layer — and of the tf32 forward / data-gradient building blocks, against an
fp64 PyTorch restatement of the same op.  The oracle is fed the operand values the kernel documents it
uses (hi + lo of the inputs for split operands, tf32-rounded inputs for the tf32 blocks, exact fp32
for the linear layer), so what is left is the dropped lo.lo term (< 2^-16 relative per product) and
fp32 accumulation order.  How far the kernels are from an exact computation on UNrounded inputs, next
to the reference's own cuDNN/cuBLAS path, is measured in tests/test_precision_gpu.py.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

H = W = 28
F32 = torch.float32


@pytest.fixture(scope="module")
def nat():
    from ddp_b200.ops import native
    native.lib()
    return native


def _lay():
    from ddp_b200.ops import layout
    return layout


def _t(x):
    return _lay().round_tf32(x)


def _sv(x):
    return _lay().split_value(x)


def _check(name, got, want64, rel):
    want = want64.to(torch.float64)
    err = (got.to(torch.float64) - want).abs().max().item()
    scale = want.abs().max().item()
    print(f"[x3-kernels] {name}: max abs err {err:.3e}, scale {scale:.3e}, relative {err / max(scale, 1e-30):.2e}")
    assert err <= rel * scale + 1e-30, f"{name}: max abs err {err:.3e} vs scale {scale:.3e} (allowed {rel:.1e} relative)"


def test_host_restatements():
    torch.manual_seed(0)
    x = torch.randn(4096)
    r = _t(x)
    assert ((r.view(torch.int32) & 0x1FFF) == 0).all()
    assert ((r - x).abs() <= x.abs() * 2.0 ** -11).all()
    s = _sv(x)
    assert ((s - x).abs() <= x.abs() * 2.0 ** -16).all()


def _packed_split(nat, w):
    packed = torch.zeros(nat.conv3x3_pack_split_bytes(w.shape[1], w.shape[0]) // 2, dtype=torch.bfloat16,
                         device="cuda")
    nat.conv3x3_pack_split(w, packed)
    return packed


# ------------------------------------------------------------------------------- split planes
@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (5, 16), (300, 300)])
def test_conv1_fwd_split(nat, batch, cap):
    torch.manual_seed(20)
    x = torch.rand(batch, 1, H, W, device="cuda")
    w = torch.randn(32, 1, 3, 3, device="cuda") * 0.3
    b = torch.randn(32, device="cuda") * 0.1
    out = nat.alloc_planes(32, cap, H, W, fmt="split")
    nat.conv1_fwd_tc_split(x, w, b, out, batch, cap, H, W)
    torch.cuda.synchronize()
    got = _lay().planes_to_nchw(out, batch, 32, H, W)
    want = F.relu(F.conv2d(x.double(), w.double(), b.double(), padding=1))  # 3xTF32: fp32-level GEMM
    _check("conv1_fwd_split", got, want, 2e-5)      # stored as hi + lo: 2^-16 relative
    assert _lay().border_rows_are_zero(out, batch, H, W)


@pytest.mark.parametrize("batch,cap", [(2, 2), (64, 64), (37, 64), (300, 300)])
def test_conv3x3_fwd_split(nat, batch, cap):
    torch.manual_seed(21)
    a = torch.relu(torch.randn(batch, 32, H, W, device="cuda"))
    w = torch.randn(64, 32, 3, 3, device="cuda") * 0.06
    b = torch.randn(64, device="cuda") * 0.1
    inp = _lay().nchw_to_split_planes(a, cap)
    out = nat.alloc_planes(64, cap, H, W, fmt="f32")
    nat.conv3x3_fwd_split(inp, _packed_split(nat, w), b, out, 32, 64, batch, cap, H, W, out_f32=True)
    torch.cuda.synchronize()
    got = _lay().planes_to_nchw(out, batch, 64, H, W)
    want = F.relu(F.conv2d(_sv(a).double(), _sv(w).double(), b.double(), padding=1))
    _check("conv3x3_fwd_split (fp32 out)", got, want, 1e-5)
    assert _lay().border_rows_are_zero(out, batch, H, W)
    out2 = nat.alloc_planes(64, cap, H, W, fmt="split")
    nat.conv3x3_fwd_split(inp, _packed_split(nat, w), b, out2, 32, 64, batch, cap, H, W, out_f32=False)
    torch.cuda.synchronize()
    _check("conv3x3_fwd_split (split out)", _lay().planes_to_nchw(out2, batch, 64, H, W), want, 2e-5)


@pytest.mark.parametrize("batch,cap", [(2, 2), (64, 64), (37, 64)])
def test_conv3x3_dgrad_split(nat, batch, cap):
    torch.manual_seed(22)
    act1 = torch.relu(torch.randn(batch, 32, H, W, device="cuda"))
    dout = torch.randn(batch, 64, H, W, device="cuda") * 0.1
    w = torch.randn(64, 32, 3, 3, device="cuda") * 0.06
    dout_p = _lay().nchw_to_split_planes(dout, cap)
    mask_p = _lay().nchw_to_split_planes(act1, cap)
    din = nat.alloc_planes(32, cap, H, W, fmt="split")
    nat.conv3x3_dgrad_split(dout_p, _packed_split(nat, w), mask_p, din, 32, 64, batch, cap, H, W)
    torch.cuda.synchronize()
    got = _lay().planes_to_nchw(din, batch, 32, H, W)
    want = F.conv_transpose2d(_sv(dout).double(), _sv(w).double(), padding=1) * (act1 > 0)
    _check("conv3x3_dgrad_split", got, want, 2e-5)
    assert _lay().border_rows_are_zero(din, batch, H, W)
    # fp32 output planes (what the exact-fp32 conv1 weight gradient reads): same values, unrounded
    din32 = nat.alloc_planes(32, cap, H, W, fmt="f32")
    nat.conv3x3_dgrad_split(dout_p, _packed_split(nat, w), mask_p, din32, 32, 64, batch, cap, H, W, out_f32=True)
    torch.cuda.synchronize()
    _check("conv3x3_dgrad_split (fp32 out)", _lay().planes_to_nchw(din32, batch, 32, H, W), want, 1e-5)
    assert _lay().border_rows_are_zero(din32, batch, H, W)


@pytest.mark.parametrize("batch,cap", [(2, 2), (64, 64), (37, 64), (300, 300)])
def test_conv3x3_wgrad_split(nat, batch, cap):
    torch.manual_seed(23)
    act1 = torch.relu(torch.randn(batch, 32, H, W, device="cuda"))
    dout = torch.randn(batch, 64, H, W, device="cuda") * 0.1
    dout_p = _lay().nchw_to_split_planes(dout, cap)
    in_p = _lay().nchw_to_split_planes(act1, cap)
    dw = torch.zeros(64, 32, 3, 3, device="cuda")
    db = torch.zeros(64, device="cuda")
    scr = nat.scratch(nat.conv3x3_wgrad_scratch_bytes(32, 64))
    nat.conv3x3_wgrad_split(dout_p, in_p, dw, db, scr, 32, 64, batch, cap, H, W)
    torch.cuda.synchronize()
    wref = torch.zeros(64, 32, 3, 3, device="cuda", dtype=torch.float64, requires_grad=True)
    bref = torch.zeros(64, device="cuda", dtype=torch.float64, requires_grad=True)
    F.conv2d(_sv(act1).double(), wref, bref, padding=1).backward(_sv(dout).double())
    _check("conv3x3_wgrad_split dW", dw, wref.grad, 2e-5)
    _check("conv3x3_wgrad_split db", db, bref.grad, 2e-5)


@pytest.mark.parametrize("batch,cap", [(1, 1), (3, 3), (64, 64), (37, 64), (300, 300)])
def test_conv1_wgrad_exact_fp32(nat, batch, cap):
    torch.manual_seed(27)
    x = torch.rand(batch, 1, H, W, device="cuda")
    dact = torch.randn(batch, 32, H, W, device="cuda") * 0.1
    dact_p = _lay().nchw_to_planes(dact, cap, dtype=F32)
    out = torch.full((32 * 9 + 32,), float("nan"), device="cuda")
    scr = nat.scratch(nat.conv1_wgrad_f32_scratch_bytes(1, 32))
    nat.conv1_wgrad_f32(dact_p, x, out, scr, 1, 32, batch, cap, H, W)
    torch.cuda.synchronize()
    wref = torch.zeros(32, 1, 3, 3, device="cuda", dtype=torch.float64, requires_grad=True)
    bref = torch.zeros(32, device="cuda", dtype=torch.float64, requires_grad=True)
    F.conv2d(x.double(), wref, bref, padding=1).backward(dact.double())
    _check("conv1_wgrad_f32 dW", out[:288].view(32, 1, 3, 3), wref.grad, 2e-6)
    _check("conv1_wgrad_f32 db", out[288:], bref.grad, 2e-6)


@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (37, 64), (300, 300)])
def test_conv1_wgrad_split(nat, batch, cap):
    torch.manual_seed(24)
    x = torch.rand(batch, 1, H, W, device="cuda")
    dact = torch.randn(batch, 32, H, W, device="cuda") * 0.1
    dact_p = _lay().nchw_to_split_planes(dact, cap)
    out = torch.zeros(32 * 9 + 32, device="cuda")
    scr = nat.scratch(nat.conv1_wgrad_tc_scratch_bytes(1, 32))
    nat.conv1_wgrad_tc_split(dact_p, x, out, scr, 1, 32, batch, cap, H, W)
    torch.cuda.synchronize()
    wref = torch.zeros(32, 1, 3, 3, device="cuda", dtype=torch.float64, requires_grad=True)
    bref = torch.zeros(32, device="cuda", dtype=torch.float64, requires_grad=True)
    F.conv2d(_sv(x).double(), wref, bref, padding=1).backward(_sv(dact).double())
    _check("conv1_wgrad_split dW", out[:288].view(32, 1, 3, 3), wref.grad, 2e-5)
    _check("conv1_wgrad_split db", out[288:], bref.grad, 2e-5)


# 600: the large-batch forward variant (4 images per thread, two CTAs per SM)
@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (37, 64), (130, 130), (600, 600)])
def test_linear_f32_fwd_loss_bwd(nat, batch, cap):
    torch.manual_seed(25)
    C = 64
    K = C * H * W
    act = torch.relu(torch.randn(batch, C, H, W, device="cuda"))
    Wfl = torch.randn(10, K, device="cuda") * 0.01
    bfl = torch.randn(10, device="cuda") * 0.1
    y = torch.randint(0, 10, (batch,), device="cuda")
    act_p = _lay().nchw_to_planes(act, cap, dtype=F32)
    part = nat.scratch(nat.linear_fwd_f32_scratch_bytes(10, C, cap, H, W))
    S = nat.linear_fwd_f32(act_p, Wfl, part, 10, C, batch, cap, H, W)
    logits = torch.zeros(batch, 10, device="cuda")
    dlogits = torch.zeros(batch, 10, device="cuda")
    dbias = torch.zeros(10, device="cuda")
    loss = torch.zeros(1, device="cuda")
    nat.softmax_xent(part, S, bfl, y, logits, dlogits, dbias, loss, 10, batch)
    lay = _lay()
    first = lay.GUARD + batch * 900
    last = lay.GUARD + lay.tiles_for(batch, H, W) * 128
    scr = nat.scratch(nat.linear_bwd_scratch_bytes(10, C, cap, H, W))
    a64 = act.double().reshape(batch, K)
    z = a64 @ Wfl.double().t() + bfl.double()
    _check("logits (fp32 FMA)", logits, z, 2e-6)
    lref = F.cross_entropy(z, y)
    assert abs(loss.item() - lref.item()) < 2e-6 * max(1.0, abs(lref.item()))
    dz = (torch.softmax(z, 1) - F.one_hot(y, 10).double()) / batch
    _check("dlogits", dlogits, dz, 1e-5)
    _check("dbias", dbias, dz.sum(0), 1e-5)
    dl64 = dlogits.double()
    want_dact = ((dl64 @ Wfl.double()).reshape(batch, C, H, W)) * (act > 0)
    for split_out, fmt, tol in ((True, "split", 2e-5), (False, "f32", 5e-4)):
        dact = nat.alloc_planes(C, cap, H, W, fmt=fmt)
        dact[:, first:last] = 7.0  # stale rows of an earlier, larger batch: the kernel must zero them
        dW = torch.full((10, K), float("nan"), device="cuda")
        bpart = nat.scratch(nat.linear_bwd_f32_bias_part_bytes(C, cap, H, W))
        bpart.fill_(float("nan"))
        ns, nb = nat.linear_bwd_f32(act_p, dlogits, Wfl, dact, dW, scr, 10, C, batch, cap, H, W,
                                    split_out=split_out, bias_part=bpart)
        torch.cuda.synchronize()
        # the exact-fp32 bias-gradient parts: host restatement of sum_bias_parts (csrc/layout.cuh)
        bp = bpart[:ns * nb * 8].view(ns, nb, 2, 4).double().cpu()
        db = torch.zeros(C, dtype=torch.float64)
        for ch in range(C):
            c4, e = ch // 4, ch % 4
            for bx in range((c4 * H * W) // 256, min(((c4 + 1) * H * W - 1) // 256, nb - 1) + 1):
                slot = c4 - (bx * 256) // (H * W)
                if slot in (0, 1):
                    db[ch] += bp[:, bx, slot, e].sum()
        _check(f"conv bias gradient from the linear backward ({fmt})", db.cuda(), want_dact.sum((0, 2, 3)), 2e-6)
        _check(f"dW_fl (fp32 FMA, {fmt} dact)", dW, dl64.t() @ a64, 2e-6)
        _check(f"dact2 ({fmt})", lay.planes_to_nchw(dact, batch, C, H, W), want_dact, tol)
        assert (dact[:, first:last] == 0).all()
        assert lay.border_rows_are_zero(dact, batch, H, W)


def test_fused_net2_update_rewrites_split_tiles(nat):
    """The fused split-K reduce + all-reduce + SGD kernel (layout = conv3x3 accumulator image) must leave
    the split tile images equal to a fresh pack of the updated weights."""
    from ddp_b200.parallel.comm import (ALGO_PUSH, CONV3X3_WGRAD_IMAGE, LAYOUT_CONV3X3_WGRAD, SymmetricComm)
    lay = _lay()
    torch.manual_seed(26)
    batch = 37
    a1 = lay.nchw_to_split_planes(torch.relu(torch.randn(batch, 32, H, W, device="cuda")), batch)
    d2 = lay.nchw_to_split_planes(torch.randn(batch, 64, H, W, device="cuda") * 0.1, batch)
    scr = nat.scratch(nat.conv3x3_wgrad_scratch_bytes(32, 64))
    n_dst = 64 * 32 * 9 + 64
    region = (n_dst * 4 + 255) // 256 * 256
    comm = SymmetricComm(0, 1, 2 * region)
    try:
        p = comm.tensor(0, n_dst)
        lr = torch.tensor([0.5], device="cuda")
        p.copy_(torch.randn(n_dst, device="cuda") * 0.05)
        packed = torch.full((nat.conv3x3_pack_split_bytes(32, 64) // 2,), float("nan"), dtype=torch.bfloat16,
                            device="cuda")
        nparts = nat.conv3x3_wgrad_split_partials(d2, a1, scr, 32, 64, batch, batch, H, W)
        comm.reduce_allreduce_sgd(scr, nparts, CONV3X3_WGRAD_IMAGE, region, 0, lr, algo=ALGO_PUSH,
                                  layout=LAYOUT_CONV3X3_WGRAD, slot=5, pack=packed)
        torch.cuda.synchronize()
        fresh = _packed_split(nat, p[:64 * 32 * 9].view(64, 32, 3, 3))
        torch.cuda.synchronize()
        assert torch.equal(packed.view(torch.int16), fresh.view(torch.int16))
    finally:
        comm.close()


# ------------------------------------------------------------------------------- tf32 building blocks
@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (300, 300)])
def test_conv1_fwd_f32_planes(nat, batch, cap):
    torch.manual_seed(30)
    x = torch.rand(batch, 1, H, W, device="cuda")
    w = torch.randn(32, 1, 3, 3, device="cuda") * 0.3
    b = torch.randn(32, device="cuda") * 0.1
    out = nat.alloc_planes(32, cap, H, W, fmt="f32")
    nat.conv1_fwd_tc_f32(x, w, b, out, batch, cap, H, W)
    torch.cuda.synchronize()
    got = _lay().planes_to_nchw(out, batch, 32, H, W)
    want = F.relu(F.conv2d(_t(x).double(), _t(w).double(), b.double(), padding=1))
    _check("conv1_fwd_f32", got, want, 6e-4)      # stored value is rounded to tf32 (2^-11 relative)
    assert ((out.view(torch.int32) & 0x1FFF) == 0).all(), "conv1 output must be tf32-rounded"
    assert _lay().border_rows_are_zero(out, batch, H, W)


@pytest.mark.parametrize("batch,cap", [(2, 2), (64, 64), (37, 64), (300, 300)])
def test_conv3x3_fwd_dgrad_tf32(nat, batch, cap):
    torch.manual_seed(31)
    a = _t(torch.relu(torch.randn(batch, 32, H, W, device="cuda")))
    dout = _t(torch.randn(batch, 64, H, W, device="cuda") * 0.1)
    w = torch.randn(64, 32, 3, 3, device="cuda") * 0.06
    b = torch.randn(64, device="cuda") * 0.1
    inp = _lay().nchw_to_planes(a, cap, dtype=F32)
    packed = torch.zeros(nat.conv3x3_pack_tf32_bytes(32, 64) // 4, device="cuda")
    nat.conv3x3_pack_tf32(w, packed)
    out = nat.alloc_planes(64, cap, H, W, fmt="f32")
    nat.conv3x3_fwd_tf32(inp, packed, b, out, 32, 64, batch, cap, H, W)
    din = nat.alloc_planes(32, cap, H, W, fmt="f32")
    nat.conv3x3_dgrad_tf32(_lay().nchw_to_planes(dout, cap, dtype=F32), packed, inp, din, 32, 64, batch, cap,
                           H, W, round_out=False)
    torch.cuda.synchronize()
    want = F.relu(F.conv2d(a.double(), _t(w).double(), b.double(), padding=1))
    _check("conv3x3_fwd_tf32", _lay().planes_to_nchw(out, batch, 64, H, W), want, 1e-5)
    want = F.conv_transpose2d(dout.double(), _t(w).double(), padding=1) * (a > 0)
    _check("conv3x3_dgrad_tf32", _lay().planes_to_nchw(din, batch, 32, H, W), want, 1e-5)
    assert _lay().border_rows_are_zero(out, batch, H, W) and _lay().border_rows_are_zero(din, batch, H, W)


@pytest.mark.parametrize("batch,cap", [(1, 1), (3, 3), (64, 64), (37, 64), (128, 128)])
def test_linear_fwd_with_fused_loss_matches_the_two_kernel_route(nat, batch, cap):
    """ddpb_linear_fwd_f32_xent (loss in the tail of the linear forward) against ddpb_linear_fwd_f32 +
    ddpb_softmax_xent and against an fp64 oracle; repeated, because the ticket word must reset itself."""
    torch.manual_seed(40)
    C, K = 64, 64 * H * W
    act = torch.relu(torch.randn(batch, C, H, W, device="cuda"))
    Wfl = torch.randn(10, K, device="cuda") * 0.01
    bfl = torch.randn(10, device="cuda") * 0.1
    y = torch.randint(0, 10, (batch,), device="cuda")
    act_p = _lay().nchw_to_planes(act, cap, dtype=F32)
    part = nat.scratch(nat.linear_fwd_f32_scratch_bytes(10, C, cap, H, W))
    S = nat.linear_fwd_f32(act_p, Wfl, part, 10, C, batch, cap, H, W)
    ref = [torch.zeros(batch, 10, device="cuda"), torch.zeros(batch, 10, device="cuda"),
           torch.zeros(10, device="cuda"), torch.zeros(1, device="cuda")]
    nat.softmax_xent(part, S, bfl, y, *ref, 10, batch)
    z = act.double().reshape(batch, K) @ Wfl.double().t() + bfl.double()
    for rep in range(3):
        got = [torch.full((batch, 10), float("nan"), device="cuda"), torch.full((batch, 10), float("nan"), device="cuda"),
               torch.full((10,), float("nan"), device="cuda"), torch.full((1,), float("nan"), device="cuda")]
        nat.linear_fwd_f32_xent(act_p, Wfl, part, bfl, y, *got, 10, C, batch, cap, H, W)
        torch.cuda.synchronize()
        for name, g, r in zip(("logits", "dlogits", "dbias", "loss"), got, ref):
            assert torch.allclose(g, r, rtol=1e-5, atol=1e-6 * max(1.0, r.abs().max().item())), (rep, name)
        _check("fused logits", got[0], z, 2e-6)
        assert abs(got[3].item() - F.cross_entropy(z, y).item()) < 2e-6 * max(1.0, abs(got[3].item()))
    assert nat.linear_fwd_f32_xent_max_batch() == 128


@pytest.mark.parametrize("batch,cap", [(3, 3), (64, 64), (37, 64), (300, 300)])
def test_linear_bwd_halves_equal_the_whole(nat, batch, cap):
    """mode 1 (dX + bias parts, main stream) + mode 2 (dW, side stream) == mode 0."""
    torch.manual_seed(41)
    C, K = 64, 64 * H * W
    lay = _lay()
    act_p = lay.nchw_to_planes(torch.relu(torch.randn(batch, C, H, W, device="cuda")), cap, dtype=F32)
    Wfl = torch.randn(10, K, device="cuda") * 0.01
    dl = torch.randn(batch, 10, device="cuda") * 0.01
    scr = nat.scratch(nat.linear_bwd_scratch_bytes(10, C, cap, H, W))
    out = {}
    for tag, modes in (("whole", (0,)), ("halves", (1, 2))):
        dact = nat.alloc_planes(C, cap, H, W, fmt="split")
        dW = torch.full((10, K), float("nan"), device="cuda")
        bp = nat.scratch(nat.linear_bwd_f32_bias_part_bytes(C, cap, H, W))
        ext = None
        for m in modes:
            e = nat.linear_bwd_f32(act_p, dl, Wfl, dact, dW, scr, 10, C, batch, cap, H, W, split_out=True,
                                   bias_part=bp, mode=m)
            ext = e if m != 2 else ext
        db = torch.zeros(64, device="cuda")
        dw2 = torch.zeros(64, 32, 3, 3, device="cuda")
        a1 = lay.nchw_to_split_planes(torch.zeros(batch, 32, H, W, device="cuda"), cap)
        nat.conv3x3_wgrad_split(dact, a1, dw2, db, nat.scratch(nat.conv3x3_wgrad_scratch_bytes(32, 64)), 32, 64,
                                batch, cap, H, W, bias_parts=(bp, *ext))
        torch.cuda.synchronize()
        out[tag] = (dact.clone(), dW, db)
    assert torch.equal(out["whole"][0].view(torch.int16), out["halves"][0].view(torch.int16))
    assert torch.equal(out["whole"][1], out["halves"][1])
    assert torch.allclose(out["whole"][2], out["halves"][2], rtol=1e-5, atol=1e-6 * out["whole"][2].abs().max().item())  # different split count
