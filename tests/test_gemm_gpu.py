"""General linear layers on the tensor cores (csrc/gemm_tc.cu, models/linear.py) against fp64 oracles:
the packed operand format, the three operand-major combinations of the GEMM, split-K reduction, the
fused ReLU + max-pool pass, ``LinearB200`` and ``WideCNNB200`` as autograd modules.
What the reference delegates to cuBLAS through nn.Linear (reference model.py:16, train_ddp.py:199)."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from ddp_b200.ops import native
    native.lib()
    return native


def _rel(a, ref):
    return float((a.double() - ref).abs().max() / ref.abs().max())


def test_pack_rows_split_layout_and_padding(nat):
    dev = "cuda"
    torch.manual_seed(0)
    rows, k = 37, 750                       # k is neither a multiple of 8 nor of 4 floats per row start
    x = torch.randn(rows, k, device=dev)
    pm = nat.PackedMatrix(100, k, dev)
    pm.buf.fill_(7.0)                       # poison: everything the pack must define gets overwritten
    pm.pack(x)
    rp, kp = 256, 768
    planes = pm.buf.view(2, kp // 8, rp, 8).float()
    hi = x.bfloat16().float()
    lo = (x - hi).bfloat16().float()
    got_hi = planes[0].permute(1, 0, 2).reshape(rp, kp)
    got_lo = planes[1].permute(1, 0, 2).reshape(rp, kp)
    assert torch.equal(got_hi[:rows, :k], hi) and torch.equal(got_lo[:rows, :k], lo)
    # the tail of the last real chunk and the rows up to the next multiple of 32 are written as zeros
    assert float(got_hi[:rows, k:752].abs().max()) == 0 and float(got_lo[:rows, k:752].abs().max()) == 0
    assert float(got_hi[rows:64, :752].abs().max()) == 0 and float(got_lo[rows:64, :752].abs().max()) == 0
    # the host-side restatement of the format (ops/layout.py) agrees on everything the pack defines
    from ddp_b200.ops.layout import pack_rows_split, unpack_rows_split
    ref = pack_rows_split(x, rows_cap=100)
    got = pm.buf.view(2, kp // 8, rp, 8)
    assert torch.equal(got[:, :94, :64, :], ref[:, :94, :64, :])
    assert torch.equal(unpack_rows_split(got, rows, k), hi + lo)
    # a mask zeroes where it is <= 0
    m = torch.randn(rows, k, device=dev)
    pm.pack(x, m)
    got_hi = pm.buf.view(2, kp // 8, rp, 8)[0].float().permute(1, 0, 2).reshape(rp, kp)
    assert torch.equal(got_hi[:rows, :k], torch.where(m > 0, hi, torch.zeros_like(hi)))


@pytest.mark.parametrize("M,N,K", [(750, 100, 1000), (10, 37, 750), (300, 256, 4096), (128, 300, 64)])
def test_gemm_forward_majors_against_fp64(nat, M, N, K):
    """A = W[M][K] K-major, B = x[N][K] K-major, stored transposed (y[N][M]) with split-K + bias + ReLU."""
    dev = "cuda"
    torch.manual_seed(1)
    w = torch.randn(M, K, device=dev)
    x = torch.randn(N, K, device=dev)
    bias = torch.randn(M, device=dev)
    a = nat.PackedMatrix(M, K, dev).pack(w)
    b = nat.PackedMatrix(N, K, dev).pack(x)
    ref = x.double() @ w.double().t()
    y = torch.full((N, M), float("nan"), device=dev)
    nat.gemm_split(a, False, b, False, M, N, K, y, M, True)
    assert _rel(y, ref) < 2e-5
    for splits in (2, 5):
        part = torch.full((splits, N, M), float("nan"), device=dev)
        nat.gemm_split(a, False, b, False, M, N, K, part, M, True, splits, N * M)
        out = torch.empty(N, M, device=dev)
        nat.splitk_reduce(part, min(splits, (K + 31) // 32), N * M, N, M, M, bias, True, out)
        assert _rel(out, torch.relu(ref + bias.double())) < 2e-5
    # error relative to what torch's own paths make of the same product
    tf32 = torch.backends.cuda.matmul.allow_tf32
    try:
        torch.backends.cuda.matmul.allow_tf32 = True
        e_tf32 = _rel(x @ w.t(), ref)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = tf32
    assert _rel(y, ref) < e_tf32 / 8


@pytest.mark.parametrize("O,B,K", [(750, 100, 1000), (10, 37, 750), (750, 256, 2048)])
def test_gemm_backward_majors_against_fp64(nat, O, B, K):
    """Data gradient (A MN-major, B K-major, transposed store) and weight gradient (both MN-major)."""
    dev = "cuda"
    torch.manual_seed(2)
    w = torch.randn(O, K, device=dev)
    x = torch.randn(B, K, device=dev)
    dy = torch.randn(B, O, device=dev)
    pw = nat.PackedMatrix(O, K, dev).pack(w)
    px = nat.PackedMatrix(B, K, dev).pack(x)
    pdy = nat.PackedMatrix(B, O, dev).pack(dy)
    dx = torch.full((B, K), float("nan"), device=dev)
    nat.gemm_split(pw, True, pdy, False, K, B, O, dx, K, True)
    assert _rel(dx, dy.double() @ w.double()) < 2e-5
    dw = torch.full((O, K), float("nan"), device=dev)
    nat.gemm_split(pdy, True, px, True, O, K, B, dw, K, False)
    assert _rel(dw, dy.double().t() @ x.double()) < 2e-5
    # split-K over the batch rows, summed in fixed order
    part = torch.full((3, O, K), float("nan"), device=dev)
    nat.gemm_split(pdy, True, px, True, O, K, B, part, K, False, 3, O * K)
    out = torch.empty(O, K, device=dev)
    nat.splitk_reduce(part, min(3, (B + 31) // 32), O * K, O, K, K, None, False, out)
    assert _rel(out, dy.double().t() @ x.double()) < 2e-5
    db = torch.empty(O, device=dev)
    nat.colsum(dy, None, db)
    assert _rel(db, dy.double().sum(0)) < 1e-6


def test_gemm_rejects_bad_arguments(nat):
    dev = "cuda"
    a = nat.PackedMatrix(16, 64, dev).pack(torch.randn(16, 64, device=dev))
    out = torch.empty(16, 16, device=dev)
    with pytest.raises(nat.NativeError):
        nat.gemm_split(a, False, a, False, 32, 16, 64, out, 16, False)   # M beyond the packed rows
    with pytest.raises(nat.NativeError):
        nat.gemm_split(a, False, a, True, 16, 16, 64, out, 16, False)    # unsupported major pair


def test_relu_maxpool_matches_torch(nat):
    dev = "cuda"
    torch.manual_seed(3)
    x = torch.randn(5, 7, 32, 32, device=dev, requires_grad=True)
    from ddp_b200.models.linear import ReLUMaxPool2dB200
    y = ReLUMaxPool2dB200(native=True)(x)
    ref = F.max_pool2d(F.relu(x), 2)
    assert torch.equal(y, ref)
    g = torch.randn_like(y)
    (dx,) = torch.autograd.grad(y, x, g)
    (dref,) = torch.autograd.grad(ref, x, g)
    assert torch.equal(dx, dref)
    # with a bias: relu(max(x) + b)
    bias = torch.randn(7, device=dev)
    out = torch.empty(5, 7, 16, 16, device=dev)
    code = torch.empty(5, 7, 16, 16, dtype=torch.uint8, device=dev)
    nat.relu_maxpool2_fwd(x.detach(), bias, out, code)
    assert torch.allclose(out, F.max_pool2d(F.relu(x.detach() + bias.view(1, 7, 1, 1)), 2), atol=1e-6)


@pytest.mark.parametrize("batch,in_f,out_f,relu", [(37, 750, 10, False), (256, 4096, 750, True), (64, 1000, 333, True)])
def test_linear_b200_against_fp64_and_torch(nat, batch, in_f, out_f, relu):
    from ddp_b200.models.linear import LinearB200
    dev = "cuda"
    torch.manual_seed(4)
    lin = LinearB200(in_f, out_f, fused_relu=relu, native=True).to(dev)
    ref = nn.Linear(in_f, out_f).to(dev).double()
    ref.load_state_dict({k: v.double() for k, v in lin.state_dict().items()})
    tl = nn.Linear(in_f, out_f).to(dev)
    tl.load_state_dict(lin.state_dict())
    act = torch.relu if relu else (lambda t: t)
    for b in (batch, max(1, batch // 3)):       # a smaller batch afterwards reuses the packed buffers
        x = torch.randn(b, in_f, device=dev, requires_grad=True)
        g = torch.randn(b, out_f, device=dev)
        y = lin(x)
        dx, dw, db = torch.autograd.grad(y, (x, lin.weight, lin.bias), g)
        xr = x.detach().double().requires_grad_(True)
        yr = act(ref(xr))
        dxr, dwr, dbr = torch.autograd.grad(yr, (xr, ref.weight, ref.bias), g.double())
        xt = x.detach().clone().requires_grad_(True)
        yt = act(tl(xt))
        dxt, dwt, dbt = torch.autograd.grad(yt, (xt, tl.weight, tl.bias), g)
        for name, ours, oracle, theirs in (("y", y, yr, yt), ("dx", dx, dxr, dxt), ("dw", dw, dwr, dwt),
                                           ("db", db, dbr, dbt)):
            e_ours, e_torch = _rel(ours, oracle), _rel(theirs, oracle)
            assert e_ours < 2e-5, (name, b, e_ours, e_torch)


def test_small_output_linear_is_exact_fp32(nat):
    """Layers with at most 16 outputs use the CUDA-core kernels of csrc/wide_ops.cu: fp32-exact results
    (1e-6 relative to fp64, where the tensor-core path's split operands give 1e-5)."""
    dev = "cuda"
    torch.manual_seed(12)
    for batch, k, nout in ((256, 750, 10), (45, 1000, 16), (1, 33, 1)):
        x = torch.randn(batch, k, device=dev)
        w = torch.randn(nout, k, device=dev)
        bias = torch.randn(nout, device=dev)
        dy = torch.randn(batch, nout, device=dev)
        y = torch.full((batch, nout), float("nan"), device=dev)
        nat.small_linear_fwd(x, w, bias, y)
        assert _rel(y, x.double() @ w.double().t() + bias.double()) < 2e-6
        dx = torch.full((batch, k), float("nan"), device=dev)
        dw = torch.full((nout, k), float("nan"), device=dev)
        db = torch.full((nout,), float("nan"), device=dev)
        nat.small_linear_bwd(dy, x, w, dx, dw, db)
        assert _rel(dx, dy.double() @ w.double()) < 2e-6
        assert _rel(dw, dy.double().t() @ x.double()) < 2e-6
        assert _rel(db, dy.double().sum(0)) < 2e-6


def test_linear_b200_weight_grad_sink_and_module_protocol(nat):
    """``weight_grad_out``: the backward writes the weight gradient into a caller's buffer and autograd
    adopts it; a second forward before backward is refused; optimizers work on the parameters."""
    from ddp_b200.models.linear import LinearB200
    dev = "cuda"
    torch.manual_seed(5)
    lin = LinearB200(512, 300, native=True).to(dev)
    sink = torch.full((300, 512), float("nan"), device=dev)
    lin.weight_grad_out = sink
    x = torch.randn(48, 512, device=dev)
    lin(x).square().sum().backward()
    ref = torch.autograd.grad(F.linear(x, lin.weight, lin.bias).square().sum(), lin.weight)[0]
    assert torch.allclose(sink, ref, rtol=1e-4, atol=1e-3)
    assert torch.equal(lin.weight.grad, sink)
    import os
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/grad_sink.txt", "w") as f:   # informative: did autograd adopt the buffer (no copy)?
        f.write(f"adopted={lin.weight.grad.data_ptr() == sink.data_ptr()}\n")
    y1 = lin(x)
    lin(x)
    with pytest.raises(RuntimeError, match="most recent forward"):
        y1.sum().backward()
    opt = torch.optim.SGD(lin.parameters(), lr=0.1)
    w0 = lin.weight.detach().clone()
    opt.zero_grad()
    lin(x).sum().backward()
    opt.step()
    assert not torch.equal(w0, lin.weight)


def test_small_cin_conv_forward_and_weight_gradient_against_fp64(nat):
    """conv1 of the widened CNN: fp32 FMA kernels, NCHW image in, NHWC activation out (csrc/wide_ops.cu)."""
    dev = "cuda"
    torch.manual_seed(7)
    b = 9
    x = torch.randn(b, 3, 32, 32, device=dev)
    conv = nn.Conv2d(3, 64, 3, padding=1).to(dev)
    out = torch.full((b, 32, 32, 64), float("nan"), device=dev)
    nat.conv3x3_small_cin_fwd(x, conv.weight.detach(), conv.bias.detach(), out)
    cd = nn.Conv2d(3, 64, 3, padding=1).to(dev).double()
    cd.load_state_dict({k: v.double() for k, v in conv.state_dict().items()})
    ref = torch.relu(cd(x.double()))
    assert _rel(out.permute(0, 3, 1, 2), ref) < 1e-6
    g = torch.randn(b, 32, 32, 64, device=dev)
    dw, db = nat.conv3x3_small_cin_wgrad(g, out, x, 64)
    dwr, dbr = torch.autograd.grad(ref, (cd.weight, cd.bias), g.permute(0, 3, 1, 2).double())
    assert dw.shape == (64, 3, 3, 3) and _rel(dw, dwr) < 1e-6 and _rel(db, dbr) < 1e-6
    dw2, _ = nat.conv3x3_small_cin_wgrad(g, None, x, 64)       # no ReLU mask
    dwr2 = torch.autograd.grad(cd(x.double()), cd.weight, g.permute(0, 3, 1, 2).double())[0]
    assert _rel(dw2, dwr2) < 1e-6


def test_relu_maxpool_nhwc_and_tall_colsum(nat):
    dev = "cuda"
    torch.manual_seed(8)
    x = torch.randn(6, 32, 32, 128, device=dev)             # NHWC
    bias = torch.randn(128, device=dev)
    out = torch.empty(6, 16, 16, 128, device=dev)
    code = torch.empty(6, 16, 16, 128, dtype=torch.uint8, device=dev)
    nat.relu_maxpool2_nhwc_fwd(x, bias, out, code)
    xn = (x.permute(0, 3, 1, 2) + bias.view(1, -1, 1, 1)).detach().requires_grad_(True)
    ref = F.max_pool2d(F.relu(xn), 2)
    assert torch.allclose(out.permute(0, 3, 1, 2), ref, atol=1e-6)
    g = torch.randn(6, 16, 16, 128, device=dev)
    din = torch.full((6, 32, 32, 128), float("nan"), device=dev)
    nat.relu_maxpool2_nhwc_bwd(g, code, din)
    (dref,) = torch.autograd.grad(ref, xn, g.permute(0, 3, 1, 2))
    assert torch.equal(din.permute(0, 3, 1, 2), dref)
    db = torch.empty(128, device=dev)
    nat.colsum(g.view(-1, 128), out.view(-1, 128), db)      # 1536 rows: the two-pass path
    assert _rel(db, dref.double().sum((0, 2, 3))) < 1e-5


def test_wide_cnn_b200_gradients_against_fp64_and_the_torch_model(nat):
    """Whole-model gradients: the native model is at least as close to an fp64 oracle as torch's own path
    (cuDNN TF32 convolutions, cuBLAS fp32 linear layers) is, up to a factor 2."""
    from ddp_b200.models.wide_cnn import WideCNN, WideCNNB200
    dev = "cuda"
    torch.manual_seed(6)
    a = WideCNNB200(native=True).to(dev)
    b = WideCNN().to(dev)
    b.load_state_dict(a.state_dict())          # (c,h,w) column order out of a, which stores (h,w,c)
    o = WideCNN().to(dev).double()
    o.load_state_dict({k: v.double() for k, v in a.state_dict().items()})
    x = torch.rand(32, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (32,), device=dev)
    la = F.cross_entropy(a(x), y)
    la.backward()
    lb = F.cross_entropy(b(x), y)
    lb.backward()
    lo = F.cross_entropy(o(x.double()), y)
    lo.backward()
    assert abs(float(la) - float(lo)) < 1e-4
    rows = []
    for (n, p), q, r in zip(a.named_parameters(), b.parameters(), o.parameters()):
        ga = a._cols_to_chw(p.grad) if n == "classifier.0.weight" else p.grad
        scale = float(r.grad.abs().max())
        e_ours = float((ga.double() - r.grad).abs().max()) / scale
        e_torch = float((q.grad.double() - r.grad).abs().max()) / scale
        rows.append((n, e_ours, e_torch))
    import os
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/precision_wide.md", "w") as f:   # copied to profiles/ as evidence
        f.write("# Widened CNN (config 4): gradient error against an fp64 oracle, batch 32\n\n"
                "max |g - g_fp64| / max |g_fp64| per parameter; `ours` = WideCNNB200 (fp32 conv1 kernels, cuDNN TF32 "
                "conv2, bf16x3 tcgen05 fc1, fp32 head), `torch` = the plain module (cuDNN TF32 convolutions, cuBLAS "
                "fp32 linear layers).  Written by tests/test_gemm_gpu.py.\n\n"
                "| parameter | ours | torch | ours / torch |\n|---|---:|---:|---:|\n")
        for n, eo, et in rows:
            f.write(f"| `{n}` | {eo:.2e} | {et:.2e} | {eo / max(et, 1e-30):.2f} |\n")
        f.write(f"\nloss: ours {float(la):.7f}, torch {float(lb):.7f}, fp64 {float(lo):.7f}\n")
    for n, e_ours, e_torch in rows:
        assert e_ours < max(2 * e_torch, 1e-5), (n, e_ours, e_torch)


def test_bucketed_engine_graph_replay_matches_eager_steps(nat):
    """BucketedDataParallel(use_graph=True) on one GPU: the captured step (native model, bucket updates on
    the side streams) leaves exactly the parameters the eager engine leaves."""
    from ddp_b200.models.wide_cnn import WideCNNB200
    from ddp_b200.parallel.bucketed import BucketedDataParallel
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(11)
    xs = [torch.rand(16, 3, 32, 32, generator=g).to(dev) for _ in range(3)]
    ys = [torch.randint(0, 10, (16,), generator=g).to(dev) for _ in range(3)]
    finals = []
    for use_graph in (False, True):
        torch.manual_seed(9)
        m = WideCNNB200(hidden=64, native=True).to(dev)
        eng = BucketedDataParallel(m, 0, 1, lr=0.05, bucket_mb=1.0, use_graph=use_graph)
        losses = [float(eng.step(xs[i % 3], ys[i % 3], F.cross_entropy)) for i in range(8)]
        torch.cuda.synchronize()
        finals.append((losses, {k: v.clone() for k, v in m.state_dict().items()}))
        if use_graph:
            assert len(eng._graphs) == 2          # one per step parity
        eng.close()
    (l0, s0), (l1, s1) = finals
    assert l0 == l1, (l0, l1)
    assert all(torch.equal(s0[k], s1[k]) for k in s0)


def test_gemm_two_m_tiles_per_cta_option(nat):
    """DDPB_GEMM_MT=2 (two accumulators per CTA sharing one B tile) is read once per process: run the three
    operand-major pairs in a child process and compare with fp64 there."""
    import os
    import subprocess
    import sys
    code = """
import sys, torch
sys.path.insert(0, %r)
from ddp_b200.ops import native as nat
nat.lib()
dev = 'cuda'
torch.manual_seed(0)
O, B, K = 750, 300, 1000
w, x, dy = torch.randn(O, K, device=dev), torch.randn(B, K, device=dev), torch.randn(B, O, device=dev)
pw, px, pdy = nat.PackedMatrix(O, K, dev).pack(w), nat.PackedMatrix(B, K, dev).pack(x), nat.PackedMatrix(B, O, dev).pack(dy)
rel = lambda a, r: float((a.double() - r).abs().max() / r.abs().max())
y = torch.full((B, O), float('nan'), device=dev)
nat.gemm_split(pw, False, px, False, O, B, K, y, O, True)
dx = torch.full((B, K), float('nan'), device=dev)
nat.gemm_split(pw, True, pdy, False, K, B, O, dx, K, True)
dw = torch.full((O, K), float('nan'), device=dev)
nat.gemm_split(pdy, True, px, True, O, K, B, dw, K, False)
errs = (rel(y, x.double() @ w.double().t()), rel(dx, dy.double() @ w.double()), rel(dw, dy.double().t() @ x.double()))
assert max(errs) < 2e-5, errs
print('mt2 ok', errs)
""" % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, DDPB_GEMM_MT="2")
    res = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and "mt2 ok" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
