"""ddpb_conv1_wgrad_noalloc: the conv1 weight-gradient entry point for embedders that cannot
allocate device memory (a host that owns all of HBM through its own pool, or a plug-in that is
not allowed to call cudaMalloc because it would stall the host's other streams).  Its scratch is
a fixed array inside the library instead of a caller buffer.

The test drives BOTH entry points through the raw C ABI (ctypes on the .so, device pointers and a
stream — exactly what such an embedder would do) and requires the same gradients from the
no-allocation variant as from ddpb_conv1_wgrad, on two calls in a row with different batches so a
stale partial from the first call would show up in the second.  It also checks the size limit and
that the call leaves the process' device-memory footprint unchanged.

The file name sorts last on purpose: a run that stops at the first failure has then already run
every other test.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pytorch-distributed-data-parallel-ddp-trainer_b200", "lib", "libddp_b200.so")
H = W = 28
DDPB_ERR_LIMIT = 9003
P, I = ctypes.c_void_p, ctypes.c_int


@pytest.fixture(scope="module")
def cabi():
    lib = ctypes.CDLL(LIB)
    lib.ddpb_conv1_wgrad.restype = I
    lib.ddpb_conv1_wgrad.argtypes = [P, P, P, P, I, I, I, I, I, I, P]
    lib.ddpb_conv1_wgrad_scratch_bytes.restype = ctypes.c_size_t
    lib.ddpb_conv1_wgrad_scratch_bytes.argtypes = [I, I, I, I, I]
    lib.ddpb_conv1_wgrad_noalloc.restype = I
    lib.ddpb_conv1_wgrad_noalloc.argtypes = [P, P, P, I, I, I, I, I, I, P]
    lib.ddpb_conv1_wgrad_noalloc_max_batch.restype = I
    lib.ddpb_conv1_wgrad_noalloc_max_batch.argtypes = [I, I]
    lib.ddpb_error_string.restype = ctypes.c_char_p
    lib.ddpb_error_string.argtypes = [I]
    return lib


def _inputs(batch, cap, seed):
    from ddp_b200.ops import layout
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(batch, 1, H, W, generator=g).cuda()
    dact = (torch.randn(batch, 32, H, W, generator=g) * 0.1).cuda()
    return x, layout.nchw_to_planes(dact, cap)


def _conv2d_oracle(batch, seed):
    """The same gradient from torch itself (fp32, TF32 off), on the bf16-rounded dact the planes hold."""
    import torch.nn.functional as F
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(batch, 1, H, W, generator=g).cuda()
    dact = (torch.randn(batch, 32, H, W, generator=g) * 0.1).cuda().to(torch.bfloat16).float()
    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        w = torch.zeros(32, 1, 3, 3, device="cuda", requires_grad=True)
        b = torch.zeros(32, device="cuda", requires_grad=True)
        F.conv2d(x, w, b, padding=1).backward(dact)
    finally:
        torch.backends.cudnn.allow_tf32 = prev
    return torch.cat([w.grad.reshape(-1), b.grad])


def _both(lib, batch, cap, seed):
    x, dact_p = _inputs(batch, cap, seed)
    stream = torch.cuda.current_stream().cuda_stream
    want = torch.zeros(320, device="cuda")
    scratch = torch.empty(lib.ddpb_conv1_wgrad_scratch_bytes(1, 32, cap, H, W) // 4 + 4, device="cuda")
    rc = lib.ddpb_conv1_wgrad(dact_p.data_ptr(), x.data_ptr(), want.data_ptr(), scratch.data_ptr(),
                              1, 32, batch, cap, H, W, stream)
    assert rc == 0, lib.ddpb_error_string(rc)
    got = torch.full((320,), float("nan"), device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.mem_get_info()[0]
    rc = lib.ddpb_conv1_wgrad_noalloc(dact_p.data_ptr(), x.data_ptr(), got.data_ptr(),
                                      1, 32, batch, cap, H, W, stream)
    assert rc == 0, lib.ddpb_error_string(rc)
    torch.cuda.synchronize()
    after = torch.cuda.mem_get_info()[0]
    return got, want, before, after


def test_noalloc_matches_allocating_entry_point_twice_in_a_row(cabi):
    # warm-up call: the first call on a device resolves the scratch array's address
    _both(cabi, 8, 8, seed=0)
    # two calls in a row with shrinking batches (a stale partial of the previous call would show), down
    # to batches of 2 and 1: fewer images than the minimum of four partial blocks, so some blocks own
    # no image and must still write all-zero partials for the reduction to read
    for batch, cap, seed in ((64, 64, 1), (37, 64, 2), (2, 2, 4), (1, 2, 5)):
        got, want, before, after = _both(cabi, batch, cap, seed)
        scale = want.abs().max().item()
        assert torch.isfinite(got).all()
        # same tolerance as tests/test_kernels_gpu.py::test_conv1_wgrad
        assert torch.allclose(got, want, atol=1e-3 * scale, rtol=1e-4), (got - want).abs().max().item()
        # and directly against torch's own convolution backward, not only against the allocating twin
        oracle = _conv2d_oracle(batch, seed)
        assert torch.allclose(got, oracle, atol=1e-3 * scale, rtol=1e-4), (got - oracle).abs().max().item()
        assert after >= before, "the no-allocation entry point must not shrink free device memory"


def test_noalloc_rejects_batches_beyond_its_fixed_scratch(cabi):
    limit = cabi.ddpb_conv1_wgrad_noalloc_max_batch(H, W)
    assert limit == 147
    cap = limit + 40
    x, dact_p = _inputs(cap, cap, seed=3)
    out = torch.zeros(320, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    ok = cabi.ddpb_conv1_wgrad_noalloc(dact_p.data_ptr(), x.data_ptr(), out.data_ptr(), 1, 32,
                                       limit, cap, H, W, stream)
    assert ok == 0, cabi.ddpb_error_string(ok)
    rc = cabi.ddpb_conv1_wgrad_noalloc(dact_p.data_ptr(), x.data_ptr(), out.data_ptr(), 1, 32,
                                       cap, cap, H, W, stream)
    assert rc == DDPB_ERR_LIMIT
    assert b"limit" in cabi.ddpb_error_string(rc)
    torch.cuda.synchronize()
