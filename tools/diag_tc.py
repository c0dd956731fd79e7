"""Bring-up probe for the tcgen05 convolution kernels: structured inputs that localise a wrong
descriptor / layout to a tap, a K-chunk or a TMEM lane mapping.  Prints a compact report.
Run on a B200:  python tools/diag_tc.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from ddp_b200.ops import layout as lay
from ddp_b200.ops import native as nat

H = W = 28
torch.backends.cudnn.allow_tf32 = False


def bf(x):
    return x.to(torch.bfloat16).float()


def summarize(name, got, want):
    err = (got - want).abs()
    print(f"  {name}: max_err={err.max().item():.4g} max_want={want.abs().max().item():.4g} "
          f"bad={(err > 2e-2 + 1e-2 * want.abs()).sum().item()}/{err.numel()}", flush=True)
    return err


def probe_fwd(batch):
    print(f"[fwd] batch={batch}")
    torch.manual_seed(0)
    a = torch.relu(torch.randn(batch, 32, H, W, device="cuda"))
    b = torch.zeros(64, device="cuda")
    inp = lay.nchw_to_planes(a, batch)
    for tap in range(9):
        w = torch.zeros(64, 32, 3, 3, device="cuda")
        for n in range(64):
            w[n, n % 32, tap // 3, tap % 3] = 1.0 if n < 32 else 0.5
        out = nat.alloc_planes(64, batch, H, W)
        nat.conv3x3_fwd(inp, w, b, out, batch, batch, H, W)
        torch.cuda.synchronize()
        got = lay.planes_to_nchw(out, batch, 64, H, W)
        want = F.relu(F.conv2d(bf(a), bf(w), b, padding=1))
        err = summarize(f"tap{tap} shift-identity", got, want)
        if err.max() > 0.05 and tap == 4:
            bad = (err > 0.05).nonzero()
            print("    first bad (b,c,h,w):", bad[:6].tolist())
            print("    bad per channel:", (err > 0.05).sum(dim=(0, 2, 3)).tolist())
            print("    bad per row h:", (err > 0.05).sum(dim=(0, 1, 3)).tolist())
    w = torch.randn(64, 32, 3, 3, device="cuda") * 0.06
    b = torch.randn(64, device="cuda") * 0.1
    out = nat.alloc_planes(64, batch, H, W)
    nat.conv3x3_fwd(inp, w, b, out, batch, batch, H, W)
    torch.cuda.synchronize()
    summarize("random", lay.planes_to_nchw(out, batch, 64, H, W),
              F.relu(F.conv2d(bf(a), bf(w), b, padding=1)))


def probe_wgrad(batch):
    print(f"[wgrad] batch={batch}")
    torch.manual_seed(1)
    act1 = torch.relu(torch.randn(batch, 32, H, W, device="cuda"))
    dout = torch.randn(batch, 64, H, W, device="cuda") * 0.1
    dw = torch.zeros(64, 32, 3, 3, device="cuda")
    db = torch.zeros(64, device="cuda")
    scr = nat.scratch(nat.conv3x3_wgrad_scratch_bytes(32, 64))
    nat.conv3x3_wgrad(lay.nchw_to_planes(dout, batch), lay.nchw_to_planes(act1, batch), dw, db, scr,
                      32, 64, batch, batch, H, W)
    torch.cuda.synchronize()
    wref = torch.zeros(64, 32, 3, 3, device="cuda", requires_grad=True)
    bref = torch.zeros(64, device="cuda", requires_grad=True)
    F.conv2d(bf(act1), wref, bref, padding=1).backward(bf(dout))
    err = summarize("dW", dw, wref.grad)
    print("    err per tap:", [round(v, 4) for v in err.amax(dim=(0, 1)).flatten().tolist()])
    print("    err per co (first 16):", [round(v, 3) for v in err.amax(dim=(1, 2, 3))[:16].tolist()])
    summarize("db", db, bref.grad)


def probe_dgrad(batch):
    print(f"[dgrad] batch={batch}")
    torch.manual_seed(2)
    act1 = torch.ones(batch, 32, H, W, device="cuda")
    dout = torch.randn(batch, 64, H, W, device="cuda") * 0.1
    w = torch.randn(64, 32, 3, 3, device="cuda") * 0.06
    din = nat.alloc_planes(32, batch, H, W)
    nat.conv3x3_dgrad(lay.nchw_to_planes(dout, batch), w, lay.nchw_to_planes(act1, batch), din,
                      batch, batch, H, W)
    torch.cuda.synchronize()
    summarize("random", lay.planes_to_nchw(din, batch, 32, H, W),
              F.conv_transpose2d(bf(dout), bf(w), padding=1))


if __name__ == "__main__":
    print("device:", torch.cuda.get_device_name(0), "SMs:", nat.lib().ddpb_num_sms())
    for fn, b in ((probe_fwd, 2), (probe_dgrad, 2), (probe_wgrad, 2), (probe_fwd, 64)):
        try:
            fn(b)
        except Exception as e:  # keep going: one broken kernel must not hide the others
            print(f"  !! {fn.__name__} raised {type(e).__name__}: {e}", flush=True)
            if "CUDA" in str(e) or "illegal" in str(e) or "launch" in str(e):
                print("  (CUDA context is poisoned; stopping)")
                break
