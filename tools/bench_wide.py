"""BASELINE.json config 4: widened CNN (~25 M params, 3x32x32 synthetic), DP on N GPUs.
native_graph = the native arm with the whole step replayed from a CUDA graph (BucketedDataParallel use_graph)
native = WideCNNB200 (fully connected layers on the tcgen05 GEMM of csrc/gemm_tc.cu, fused ReLU+max-pool;
         convolutions through cuDNN) under BucketedDataParallel, fc1's weight gradient written straight
         into the gradient arena
ours   = the torch module under BucketedDataParallel (fused all-reduce+SGD per bucket on side streams,
         our communicator): isolates the data-parallel layer
torch  = DistributedDataParallel + NCCL + torch.optim.SGD on the torch module and the same inputs.

    python -m torch.distributed.run --nproc-per-node 8 tools/bench_wide.py [--batch 256] [--steps 30]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F

from ddp_b200.models.wide_cnn import WideCNN, WideCNNB200, param_count


def timed_steps(step, xs, ys, steps, warmup, world, dev):
    for i in range(warmup):
        step(xs[i % len(xs)], ys[i % len(ys)])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        loss = step(xs[i % len(xs)], ys[i % len(ys)])
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), float(loss)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=6)
    ap.add_argument("--bucket_mb", type=float, default=25.0)
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    g = torch.Generator().manual_seed(1)
    xs = [torch.rand(a.batch, 3, 32, 32, generator=g).to(dev) for _ in range(4)]
    ys = [torch.randint(0, 10, (a.batch,), generator=g).to(dev) for _ in range(4)]
    out = {"config": "wide_cnn", "world": world, "per_rank_batch": a.batch}

    from ddp_b200.parallel.bucketed import BucketedDataParallel
    torch.manual_seed(0)
    mn = WideCNNB200(native=True).to(dev)
    eng = BucketedDataParallel(mn, rank, world, lr=0.01, bucket_mb=a.bucket_mb)
    ms, loss = timed_steps(lambda x, y: eng.step(x, y, F.cross_entropy), xs, ys, a.steps, a.warmup, world, dev)
    out["native_ms_per_step"], out["native_img_s"], out["native_loss"] = ms, world * a.batch / ms * 1e3, loss
    eng.close()
    del eng, mn
    torch.cuda.empty_cache()

    # the same arm with the whole step (forward, backward, bucket exchanges) replayed from a CUDA graph
    torch.manual_seed(0)
    mg = WideCNNB200(native=True).to(dev)
    eng = BucketedDataParallel(mg, rank, world, lr=0.01, bucket_mb=a.bucket_mb, use_graph=True)
    ms, loss = timed_steps(lambda x, y: eng.step(x, y, F.cross_entropy), xs, ys, a.steps, max(a.warmup, 6), world, dev)  # >= 6: two eager steps and one capture per parity
    out["native_graph_ms_per_step"], out["native_graph_img_s"] = ms, world * a.batch / ms * 1e3
    out["native_graph_loss"] = loss
    eng.close()
    del eng, mg
    torch.cuda.empty_cache()

    torch.manual_seed(0)
    m = WideCNN().to(dev)
    out["params"] = param_count(m)
    eng = BucketedDataParallel(m, rank, world, lr=0.01, bucket_mb=a.bucket_mb)
    out["buckets_mb"] = [round(b["n"] * 4 / 2**20, 2) for b in eng.buckets]
    ms, loss = timed_steps(lambda x, y: eng.step(x, y, F.cross_entropy), xs, ys, a.steps, a.warmup, world, dev)
    out["ours_ms_per_step"], out["ours_img_s"], out["ours_loss"] = ms, world * a.batch / ms * 1e3, loss
    eng.close()
    del eng, m
    torch.cuda.empty_cache()

    from torch.nn.parallel import DistributedDataParallel as DDP
    torch.manual_seed(0)
    m2 = WideCNN().to(dev)
    ddp = DDP(m2, device_ids=[local]) if world > 1 else m2
    opt = torch.optim.SGD(ddp.parameters(), lr=0.01)

    def torch_step(x, y):
        opt.zero_grad()
        loss = F.cross_entropy(ddp(x), y)
        loss.backward()
        opt.step()
        return loss.detach()

    ms, loss = timed_steps(torch_step, xs, ys, a.steps, a.warmup, world, dev)
    out["torch_ddp_ms_per_step"], out["torch_ddp_img_s"], out["torch_loss"] = ms, world * a.batch / ms * 1e3, loss
    out["native_speedup_vs_torch_ddp"] = out["torch_ddp_ms_per_step"] / out["native_ms_per_step"]
    out["native_graph_speedup_vs_torch_ddp"] = out["torch_ddp_ms_per_step"] / out["native_graph_ms_per_step"]
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
