"""Localise a data-parallel mismatch: one step of the N-rank engine vs the one-GPU engine on the
concatenated batch, per parameter.  torchrun --nproc-per-node N tools/dbg_dp.py [precision] [steps]"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddp_b200.parallel.engine import B200DataParallel  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
use_graph = (sys.argv[3] == "graph") if len(sys.argv) > 3 else False
B = 64
eng = B200DataParallel(rank, world, dev, batch_cap=B, lr=0.01, use_graph=use_graph, precision=prec)
p0 = eng.flat_params.detach().clone()
g = torch.Generator().manual_seed(7)
xs = torch.rand(steps, world, B, 1, 28, 28, generator=g).to(dev)
ys = torch.randint(0, 10, (steps, world, B), generator=g).to(dev)
for i in range(steps):
    eng.step(xs[i, rank], ys[i, rank])
eng.synchronize()
solo = B200DataParallel(0, 1, dev, batch_cap=B * world, lr=0.01, use_graph=False, precision=prec)
solo.flat_params.copy_(p0)
solo.model.pack_weights()
for i in range(steps):
    solo.step(xs[i].reshape(-1, 1, 28, 28), ys[i].reshape(-1))
solo.synchronize()
if rank == 0:
    for s in eng.model.arena.specs:
        a, b = eng.model.arena.view(s.name), solo.model.arena.view(s.name)
        ref = p0[s.offset:s.offset + s.numel].view(s.shape)
        upd = (b - ref).abs().max().item()
        print(f"{prec} W={world} steps={steps} {s.name:14s} max|dp - solo| = {(a - b).abs().max().item():.3e}   max|update| = {upd:.3e}")
solo.close(sync_peers=False)
eng.close()
dist.destroy_process_group()
