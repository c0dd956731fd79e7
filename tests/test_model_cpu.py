"""CPU model path: same function and initialisation as the reference classifier
(reference model.py:4-20), gradients land in the flat arena, buckets are contiguous slices."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from ddp_b200.models.simple_cnn import SimpleCNN


def _torch_modules():
    net = nn.Sequential(nn.Conv2d(1, 32, 3, padding=1), nn.ReLU(), nn.Conv2d(32, 64, 3, padding=1),
                        nn.ReLU(), nn.Flatten())
    fl = nn.Linear(50176, 10)
    return net, fl


def test_param_count_and_layout():
    m = SimpleCNN()
    assert m.arena.numel == 520586
    assert [s.offset for s in m.arena.specs] == [0, 288, 320, 18752, 18816, 520576]
    b0, b1 = m.buckets()
    assert (b0.start, b0.numel) == (18816, 501770)      # fl.weight + fl.bias: 2,007,080 B
    assert (b1.start, b1.numel) == (0, 18816)           # conv params: 75,264 B
    assert b0.numel * 4 == 2007080 and b1.numel * 4 == 75264


def test_init_matches_torch_default_under_same_seed():
    torch.manual_seed(42)
    m = SimpleCNN()
    torch.manual_seed(42)
    net, fl = _torch_modules()
    sd = m.state_dict()
    assert torch.equal(sd["net.0.weight"], net[0].weight)
    assert torch.equal(sd["net.0.bias"], net[0].bias)
    assert torch.equal(sd["net.2.weight"], net[2].weight)
    assert torch.equal(sd["net.2.bias"], net[2].bias)
    assert torch.equal(sd["fl.weight"], fl.weight)
    assert torch.equal(sd["fl.bias"], fl.bias)


def test_forward_backward_matches_autograd_modules():
    torch.manual_seed(0)
    m = SimpleCNN()
    torch.manual_seed(0)
    net, fl = _torch_modules()
    x = torch.rand(5, 1, 28, 28)
    y = torch.randint(0, 10, (5,))
    loss = m.forward_backward(x, y)
    ref = F.cross_entropy(fl(net(x)), y)
    ref.backward()
    assert torch.allclose(loss, ref.detach(), atol=1e-6)
    assert m.arena.params.grad is m.arena.grads
    for name, p in (("net.0.weight", net[0].weight), ("net.0.bias", net[0].bias),
                    ("net.2.weight", net[2].weight), ("net.2.bias", net[2].bias),
                    ("fl.weight", fl.weight), ("fl.bias", fl.bias)):
        assert torch.allclose(m.arena.grad_view(name), p.grad, atol=1e-6), name
    # second step: gradients are overwritten, not accumulated
    m.forward_backward(x, y)
    assert torch.allclose(m.arena.grad_view("fl.bias"), fl.bias.grad, atol=1e-6)


def test_sgd_step_matches_torch_sgd():
    torch.manual_seed(0)
    m = SimpleCNN()
    torch.manual_seed(0)
    net, fl = _torch_modules()
    opt = torch.optim.SGD(list(net.parameters()) + list(fl.parameters()), lr=0.01)
    x = torch.rand(4, 1, 28, 28)
    y = torch.randint(0, 10, (4,))
    for _ in range(3):
        m.forward_backward(x, y)
        m.sgd_step(0.01)
        opt.zero_grad()
        F.cross_entropy(fl(net(x)), y).backward()
        opt.step()
    assert torch.allclose(m.arena.view("fl.weight").detach(), fl.weight.detach(), atol=1e-6)
    assert torch.allclose(m.arena.view("net.2.weight").detach(), net[2].weight.detach(), atol=1e-6)


def test_state_dict_is_a_copy():
    m = SimpleCNN()
    sd = m.state_dict()
    sd["fl.bias"].add_(1.0)
    assert not torch.equal(sd["fl.bias"], m.arena.view("fl.bias").detach())


def test_header_declares_every_exported_entry_point():
    """include/ddp_b200.h is the C contract: every extern "C" ddpb_* function defined in csrc/ must
    be declared there (checked on the sources, so it runs without a GPU or a built library)."""
    import glob
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "ddp_b200.h")).read()
    declared = set(re.findall(r"\b(ddpb_[a-z0-9_]+)\s*\(", header))
    defined = set()
    for path in glob.glob(os.path.join(root, "csrc", "*.cu")):
        src = open(path).read()
        defined |= set(re.findall(r"^(?:extern \"C\" )?(?:const )?[a-z_ ]+\*?\s*\*?(ddpb_[a-z0-9_]+)\s*\(", src, re.M))
    assert len(defined) > 40, sorted(defined)
    missing = sorted(d for d in defined if d not in declared)
    assert not missing, f"not declared in include/ddp_b200.h: {missing}"


def test_staged_backward_equals_plain_backward():
    """The bucket-by-bucket backward (used to overlap the gloo all-reduce with the convolution
    backward) must give bit-identical gradients and report the buckets in gradient-ready order."""
    from ddp_b200.models.simple_cnn import BUCKET_GROUPS, SimpleCNN
    torch.manual_seed(4)
    m = SimpleCNN()
    x, y = torch.rand(6, 1, 28, 28), torch.randint(0, 10, (6,))
    l0 = m.forward_backward(x, y)
    g0 = m.arena.grads.clone()
    m.arena.grads.fill_(123.0)
    order = []
    seen_at_call = []

    def ready(i):
        order.append(i)
        b = m.buckets()[i]
        seen_at_call.append(torch.equal(m.arena.grads[b.start:b.start + b.numel], g0[b.start:b.start + b.numel]))

    l1 = m.forward_backward_staged(x, y, ready)
    assert order == list(range(len(BUCKET_GROUPS))) and all(seen_at_call)
    assert torch.equal(l0, l1)
    covered = torch.zeros_like(g0, dtype=torch.bool)
    for b in m.buckets():
        covered[b.start:b.start + b.numel] = True
    assert torch.equal(m.arena.grads[covered], g0[covered])


def test_every_kernel_launched_as_programmatic_dependent_waits_for_its_predecessor():
    """launch_k() (csrc/launch.cuh) launches with the programmatic-stream-serialization attribute: such
    a kernel may start while its stream predecessor is still running and MUST execute
    griddepcontrol.wait (pdl_sync / pdl_wait) before touching anything the predecessor wrote.  A kernel
    without it silently races (found once: the exact-fp32 conv1 weight gradient read dact1 early and
    the replicas drifted from the single-GPU trajectory after a few steps).  Checked on the sources."""
    import glob
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    srcs = {p: open(p).read() for p in glob.glob(os.path.join(root, "csrc", "*.cu"))}
    everything = "\n".join(srcs.values())
    launched = set(re.findall(r"launch_k\(\s*([A-Za-z_0-9]+)\s*[<,]", everything))
    # `auto kern = some_kernel<...>; launch_k(kern, ...)`
    launched |= set(re.findall(r"auto kern = ([A-Za-z_0-9]+)<", everything))
    launched.discard("kern")
    assert len(launched) >= 15, sorted(launched)
    missing = []
    for name in sorted(launched):
        m = re.search(r"__global__[^;{]*?\b" + name + r"\s*\([^)]*\)\s*\{", everything, re.S)
        assert m, f"definition of {name} not found"
        depth, i = 1, m.end()
        while depth and i < len(everything):
            depth += {"{": 1, "}": -1}.get(everything[i], 0)
            i += 1
        body = everything[m.end():i]
        if "pdl_sync()" not in body and "pdl_wait()" not in body:
            missing.append(name)
    assert not missing, f"launched through launch_k without griddepcontrol.wait: {missing}"


def test_linear_b200_and_wide_cnn_b200_fall_back_to_torch_ops_on_cpu():
    """models/linear.py on CPU tensors: same parameters, keys and results as nn.Linear / WideCNN."""
    import torch.nn as nn
    from ddp_b200.models.linear import LinearB200, ReLUMaxPool2dB200
    from ddp_b200.models.wide_cnn import WideCNN, WideCNNB200
    torch.manual_seed(0)
    a = LinearB200(20, 7, fused_relu=True)
    torch.manual_seed(0)
    b = nn.Linear(20, 7)
    assert list(a.state_dict()) == list(b.state_dict())
    x = torch.randn(5, 20)
    assert torch.equal(a(x), torch.relu(b(x)))
    p = torch.randn(2, 3, 8, 8)
    assert torch.equal(ReLUMaxPool2dB200()(p), nn.functional.max_pool2d(torch.relu(p), 2))
    torch.manual_seed(1)
    wa = WideCNNB200(hidden=16)
    torch.manual_seed(1)
    wb = WideCNN(hidden=16)
    sa, sb = wa.state_dict(), wb.state_dict()
    assert list(sa) == list(sb) and all(torch.equal(sa[k], sb[k]) for k in sa)   # same seed, same values
    # fc1's weight is STORED with its columns in (h, w, c) order; state_dict converts both ways
    assert not torch.equal(wa.classifier[0].weight, wb.classifier[0].weight)
    xi = torch.rand(2, 3, 32, 32)
    assert torch.allclose(wa(xi), wb(xi), atol=1e-5)
    wc = WideCNNB200(hidden=16)
    wc.load_state_dict(sb)
    assert torch.equal(wc.classifier[0].weight, wa.classifier[0].weight)


def test_split_row_planes_reference_roundtrip():
    """ops/layout.py restates the packed operand format of the general GEMM (csrc/gemm_tc.cu): two bf16
    terms carry a 16-bit significand, the padding is zero, element (r, k) sits at [set, k//8, r, k%8]."""
    from ddp_b200.ops.layout import GEMM_PAD, pack_rows_split, unpack_rows_split
    torch.manual_seed(0)
    x = torch.randn(37, 750)
    p = pack_rows_split(x, rows_cap=100)
    assert p.shape == (2, 768 // 8, GEMM_PAD, 8) and p.dtype == torch.bfloat16
    assert float(p[0, 750 // 8, 5, 750 % 8 - 1]) == float(x[5, 749].bfloat16())
    back = unpack_rows_split(p, 37, 750)
    assert float((back - x).abs().max() / x.abs().max()) < 2 ** -15
    assert float(p[:, :, 37:, :].float().abs().max()) == 0 and float(p[:, 94:, :, :].float().abs().max()) == 0
