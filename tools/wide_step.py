"""A few eager training steps of the widened CNN with the native layers (models/wide_cnn.py WideCNNB200)
on one GPU: the program tools/run_ncu_wide.sh profiles.    python tools/wide_step.py [--batch 256] [--steps 4]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=4)
    a = ap.parse_args()
    from ddp_b200.models.wide_cnn import WideCNNB200
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    m = WideCNNB200(native=True).to(dev)
    opt = torch.optim.SGD(m.parameters(), lr=0.01)
    x = torch.rand(a.batch, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (a.batch,), device=dev)
    for _ in range(a.steps):
        opt.zero_grad()
        loss = F.cross_entropy(m(x), y)
        loss.backward()
        opt.step()
    torch.cuda.synchronize()
    print(f"ok loss {float(loss):.4f}")


if __name__ == "__main__":
    main()
