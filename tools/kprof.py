"""In-kernel phase timing of the tcgen05 conv kernels (uses ddpb_set_prof_buffer).
Stamps (SM clock): 0 start, 1 weights staged, 2 setup done, 4 first tile landed (MMA thread),
5 last MMA issued, 6 first accumulator ready (epilogue), 7 epilogue done, 8 kernel end.
    python tools/kprof.py [batch ...]
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from ddp_b200.ops import layout as lay
from ddp_b200.ops import native as nat

H = W = 28
MHZ = 1965.0


def report(name, prof, nblocks):
    p = prof[: nblocks * 16].view(nblocks, 16).cpu().double()
    t0 = p[:, 0].min()
    def span(a, b):
        m = (p[:, a] > 0) & (p[:, b] > 0)
        if not m.any():
            return "   n/a"
        d = (p[m, b] - p[m, a]) / MHZ
        return f"{d.mean():6.2f}/{d.max():6.2f}"
    end = ((p[:, 8].max() - t0) / MHZ).item()
    print(f"{name:28s} blocks={nblocks:3d} total={end:7.2f}us | start-skew(max)={(p[:,0].max()-t0).item()/MHZ:5.2f} "
          f"| 0-1 {span(0,1)} | 1-2 {span(1,2)} | 2-4 {span(2,4)} | 4-5 {span(4,5)} | 5-6' {span(5,7)} | 6-7 {span(6,7)} | 7-8 {span(7,8)}  (mean/max us)",
          flush=True)


def main():
    lib = nat.lib()
    lib.ddpb_set_prof_buffer.argtypes = [ctypes.c_void_p]
    lib.ddpb_set_prof_buffer.restype = None
    sms = lib.ddpb_num_sms()
    prof = torch.zeros(sms * 16, dtype=torch.int64, device="cuda")
    batches = [int(a) for a in sys.argv[1:]] or [64, 4096]
    for b in batches:
        torch.manual_seed(0)
        a1 = lay.nchw_to_planes(torch.relu(torch.randn(b, 32, H, W, device="cuda")), b)
        d2 = lay.nchw_to_planes(torch.randn(b, 64, H, W, device="cuda") * 0.1, b)
        w = torch.randn(64, 32, 3, 3, device="cuda") * 0.06
        bias = torch.zeros(64, device="cuda")
        out2 = nat.alloc_planes(64, b, H, W)
        out1 = nat.alloc_planes(32, b, H, W)
        dw = torch.zeros(64, 32, 3, 3, device="cuda")
        db = torch.zeros(64, device="cuda")
        scr = nat.scratch(nat.conv3x3_wgrad_scratch_bytes(32, 64))
        nblocks = min(sms, lay.tiles_for(b, H, W))
        calls = {
            "fwd": lambda: nat.conv3x3_fwd(a1, w, bias, out2, b, b, H, W),
            "dgrad": lambda: nat.conv3x3_dgrad(d2, w, a1, out1, b, b, H, W),
            "wgrad": lambda: nat.conv3x3_wgrad(d2, a1, dw, db, scr, 32, 64, b, b, H, W),
        }
        for name, fn in calls.items():
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            for _ in range(10):
                fn()
            ev1.record()
            torch.cuda.synchronize()
            prof.zero_()
            lib.ddpb_set_prof_buffer(prof.data_ptr())
            fn()
            torch.cuda.synchronize()
            lib.ddpb_set_prof_buffer(None)
            report(f"{name} b={b} ({ev0.elapsed_time(ev1) * 100:.1f}us/call)", prof, nblocks)


if __name__ == "__main__":
    main()
