"""Time the native fully connected layer (csrc/gemm_tc.cu through models/linear.py) against torch's
nn.Linear (cuBLAS fp32, and cuBLAS with TF32 allowed) on one GPU, forward + backward, CUDA events.

    python tools/bench_linear.py [--batch 256] [--in_f 32768] [--out_f 750] [--iters 30]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn


def timed(fn, iters, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--in_f", type=int, default=32768)
    ap.add_argument("--out_f", type=int, default=750)
    ap.add_argument("--iters", type=int, default=30)
    a = ap.parse_args()
    from ddp_b200.ops import native as nat
    from ddp_b200.models.linear import LinearB200
    nat.lib()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    B, K, O = a.batch, a.in_f, a.out_f
    x = torch.randn(B, K, device=dev, requires_grad=True)
    g = torch.randn(B, O, device=dev)
    out = {"shape": {"batch": B, "in": K, "out": O}, "flops_fwd_bwd": 6.0 * B * K * O}

    lin = LinearB200(K, O, fused_relu=True, native=True).to(dev)

    def ours():
        y = lin(x)
        torch.autograd.grad(y, (x, lin.weight, lin.bias), g)

    out["ours_fwd_bwd_us"] = timed(ours, a.iters)

    # the pieces
    pw, px, pdy = nat.PackedMatrix(O, K, dev), nat.PackedMatrix(B, K, dev), nat.PackedMatrix(B, O, dev)
    w = lin.weight.detach()
    xd = x.detach()
    out["pack_w_us"] = timed(lambda: pw.pack(w), a.iters)
    out["pack_x_us"] = timed(lambda: px.pack(xd), a.iters)
    out["pack_dy_us"] = timed(lambda: pdy.pack(g), a.iters)
    s = nat.gemm_suggest_splits(O, B, K)
    part = torch.empty(s, B, O, device=dev)
    y = torch.empty(B, O, device=dev)
    out["fwd_splits"] = s
    out["fwd_gemm_us"] = timed(lambda: nat.gemm_split(pw, False, px, False, O, B, K, part, O, True, s, B * O), a.iters)
    out["fwd_reduce_us"] = timed(lambda: nat.splitk_reduce(part, s, B * O, B, O, O, lin.bias.detach(), True, y), a.iters)
    dx = torch.empty(B, K, device=dev)
    s2 = nat.gemm_suggest_splits(K, B, O)
    out["dgrad_splits"] = s2
    part2 = torch.empty(s2, B, K, device=dev)
    out["dgrad_gemm_us"] = timed(lambda: nat.gemm_split(pw, True, pdy, False, K, B, O, part2, K, True, s2, B * K), a.iters)
    dw = torch.empty(O, K, device=dev)
    s3 = nat.gemm_suggest_splits(O, K, B)
    out["wgrad_splits"] = s3
    part3 = torch.empty(s3, O, K, device=dev)
    out["wgrad_gemm_us"] = timed(lambda: nat.gemm_split(pdy, True, px, True, O, K, B, part3, K, False, s3, O * K), a.iters)
    del part, part2, part3, dx, dw

    tl = nn.Sequential(nn.Linear(K, O), nn.ReLU()).to(dev)

    def theirs():
        y = tl(x)
        torch.autograd.grad(y, (x, tl[0].weight, tl[0].bias), g)

    for name, flag in (("torch_fp32_fwd_bwd_us", False), ("torch_tf32_fwd_bwd_us", True)):
        old = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = flag
        try:
            out[name] = timed(theirs, a.iters)
        finally:
            torch.backends.cuda.matmul.allow_tf32 = old
    out["speedup_vs_fp32"] = out["torch_fp32_fwd_bwd_us"] / out["ours_fwd_bwd_us"]
    out["speedup_vs_tf32"] = out["torch_tf32_fwd_bwd_us"] / out["ours_fwd_bwd_us"]
    out["ours_tflops_effective"] = out["flops_fwd_bwd"] / out["ours_fwd_bwd_us"] / 1e6
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
