"""Per-kernel device time of one widened-CNN training step (BASELINE.json config 4), native and torch
models, with torch.profiler (CUPTI).  Diagnostic only: numbers taken under a profiler are not bench values.

    python tools/profile_wide.py [--batch 256]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from torch.profiler import ProfilerActivity, profile


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--rows", type=int, default=22)
    a = ap.parse_args()
    from ddp_b200.models.wide_cnn import WideCNN, WideCNNB200
    dev = torch.device("cuda", 0)
    x = torch.rand(a.batch, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (a.batch,), device=dev)
    for name, ctor in (("native", lambda: WideCNNB200(native=True)), ("torch", WideCNN)):
        torch.manual_seed(0)
        m = ctor().to(dev)
        opt = torch.optim.SGD(m.parameters(), lr=0.01)

        def step():
            opt.zero_grad()
            F.cross_entropy(m(x), y).backward()
            opt.step()

        for _ in range(5):
            step()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(5):
                step()
            torch.cuda.synchronize()
        print(f"==== {name}: per-kernel device time over 5 steps")
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=a.rows, max_name_column_width=70))
        del m, opt


if __name__ == "__main__":
    main()
