#!/usr/bin/env python
"""Headline benchmark: data-parallel training throughput of the reference's SimpleCNN
(BASELINE.json: "train images/sec + step latency at 1/2/4/8 B200 vs ref NCCL DDP").

    python bench.py --gpus N --steps K --warmup W [--batch_size B] [--impl reference]

N > 1 is launched by the driver under ``python -m torch.distributed.run`` (one rank per GPU,
RANK/LOCAL_RANK/WORLD_SIZE from the environment); run directly with --gpus N it re-launches
itself that way.  Rank 0 prints ONE JSON line.

What is timed (both arms identically):
  * ``value`` / ``ms_per_step``: K training steps (forward, backward, gradient all-reduce,
    optimizer update) with inputs already on the device, each step bracketed by CUDA events on
    the launching stream, an L2 flush (256 MiB write) between timed steps, barrier +
    synchronize before and after the loop, max over ranks.  Weak scaling: per-rank batch fixed.
  * ``e2e``: the same K steps through the public API a user calls (ours: engine.step(x, y);
    reference: its loop body) where EVERY step copies that step's fp32 images + int64 labels
    from pinned host memory and reads the loss back to the host; wall clock between
    synchronised barriers, max over ranks.

``--impl reference`` runs the UNMODIFIED reference modules installed in baseline/_ref
(model.SimpleCNN, utils.setup/cleanup) through the reference's own stock path — torch DDP over
NCCL, cuDNN/cuBLAS kernels, torch.optim.SGD — with only the device-placement edits the reference
needs to run on CUDA at all.  Nothing of this repo is imported in that arm.
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(ROOT, "baseline", "_ref")
L2_FLUSH_BYTES = 256 << 20
POOL = 16  # distinct synthetic batches cycled through


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi sampled every 200 ms while the timed region runs (B200_PROFILING.md recipe)."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(prefix="clocks_", suffix=".csv")
            os.close(fd)
            self.out = open(self.path, "w")
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "200", "-i", str(self.gpu_index)], stdout=self.out, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.out.close()
        sm, smmax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        with open(self.path) as f:
            for line in f:
                p = [s.strip() for s in line.split(",")]
                if len(p) < 9:
                    continue
                try:
                    sm.append(float(p[1]))
                    smmax.append(float(p[2]))
                    power.append(float(p[3]))
                except ValueError:
                    continue
                for name, val in zip(names, p[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
        os.unlink(self.path)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(smmax) if smmax else None,
                "power_w_max": max(power) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def _physical_gpu_index(local_rank: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


# --------------------------------------------------------------------------- shared pieces
def _dist_env(args):
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    return rank, world, local


def _make_pool(batch: int, pinned: bool, rank: int = 0):
    """Synthetic batches of THIS rank (both arms): a different stream per rank, so the ranks push
    different gradients and a collective that skipped a peer would change the result."""
    import torch
    g = torch.Generator().manual_seed(1234 + rank)
    xs = torch.rand(POOL, batch, 1, 28, 28, generator=g)
    ys = torch.randint(0, 10, (POOL, batch), generator=g)
    if pinned:
        xs, ys = xs.pin_memory(), ys.pin_memory()
    return xs, ys


def _max_over_ranks(value: float, world: int, device) -> float:
    import torch
    import torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _barrier_sync(world: int):
    import torch
    import torch.distributed as dist
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def _emit(args, rank, world, impl, dev_ms_total, e2e_s_total, clocks, launches, dtype, extra_cfg):
    if rank != 0:
        return
    B, K = args.batch_size, args.steps
    ms = dev_ms_total / K
    value = world * B / (ms * 1e-3)
    e2e_value = world * B * K / e2e_s_total
    out = {
        "metric": "train_images_per_sec", "value": round(value, 1), "unit": "images/s",
        "n_gpus": world, "steps": K, "warmup": max(3, args.warmup), "ms_per_step": round(ms, 5),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": dtype, "data": "synthetic", "impl": impl,
        "config": dict({"model": "SimpleCNN (reference model.py: conv1-32,conv32-64,fc50176-10)",
                        "global_batch": B * world, "per_rank_batch": B, "seq_len": None,
                        "input": "1x28x28 fp32", "optimizer": "SGD lr=0.01",
                        "parallelism": f"dp{world}",
                        "l2": "flushed between timed steps (256 MiB write, outside the events)",
                        "timing": "CUDA events per step on the launching stream, max over ranks"},
                       **extra_cfg),
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 1), "unit": "images/s",
                "ms_per_step": round(e2e_s_total * 1e3 / K, 5),
                "h2d_bytes_per_step": B * 784 * 4 + B * 8, "d2h_bytes_per_step": 4,
                "timing": "wall clock between synchronised barriers, max over ranks"},
        "gpu_launches": launches,
    }
    print(json.dumps(out), flush=True)


# --------------------------------------------------------------------------- our arm
def run_ours(args):
    rank, world, local = _dist_env(args)
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    from ddp_b200.ops import native
    native.lib()  # fail loudly if the native library is missing
    from ddp_b200.parallel.engine import B200DataParallel
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    B, K, W = args.batch_size, args.steps, max(3, args.warmup)
    torch.manual_seed(0)
    eng = B200DataParallel(rank, world, device, batch_cap=B, lr=0.01, use_graph=not args.no_graph,
                           algo=args.algo, precision=args.dtype)
    xs_h, ys_h = _make_pool(B, pinned=True, rank=rank)
    xs_d, ys_d = xs_h.to(device), ys_h.to(device)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=device)

    def stage(i):  # put batch i into the engine's static inputs, outside the timed events
        xb, yb = eng.acquire_inputs()
        with torch.cuda.stream(eng.main):
            eng.main.wait_event(eng._ev_free[eng.step_count & 1])
            xb.copy_(xs_d[i % POOL], non_blocking=True)
            yb.copy_(ys_d[i % POOL], non_blocking=True)
        return xb, yb

    # ---- device-timed region
    for i in range(W):
        xb, yb = stage(i)
        eng.step(xb, yb)
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    stops = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    sampler = ClockSampler(_physical_gpu_index(local))
    _barrier_sync(world)
    sampler.start()
    for i in range(K):
        xb, yb = stage(W + i)
        with torch.cuda.stream(eng.main):
            flush.zero_()
            starts[i].record(eng.main)
        eng.step(xb, yb)
        stops[i].record(eng.main)
    _barrier_sync(world)
    clocks = sampler.stop()
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, stops))
    dev_ms = _max_over_ranks(dev_ms, world, device)
    loss_dev = eng.read_loss()

    # ---- end-to-end region: pinned host inputs every step, loss read back every step
    for i in range(W):
        eng.step(xs_h[i % POOL], ys_h[i % POOL])
    losses = []
    _barrier_sync(world)
    t0 = time.perf_counter()
    for i in range(K):
        eng.step(xs_h[i % POOL], ys_h[i % POOL])
        losses.append(eng.loss_host[(eng.step_count - 1) & 1])  # async D2H landed in pinned memory
    eng.synchronize()
    _barrier_sync(world)
    e2e_s = _max_over_ranks(time.perf_counter() - t0, world, device)
    loss_e2e = eng.read_loss()
    launches = K * eng.launches_per_step(B)
    checks = _replica_checks(eng, rank, world, device, B, args)
    extra = {"graph": not args.no_graph, **checks,
             "allreduce": "fused (split-K reduce +) all-reduce + SGD kernels over peer memory, in-kernel flags",
             "communicator": "vmm+multicast" if getattr(getattr(eng, "comm", None), "multicast", False) else "cuda-ipc",
             "tail_update": getattr(eng, "tail_update", None),
             "net2_order": getattr(eng, "net2_order", None),
             "final_loss": round(loss_e2e, 4), "loss_after_device_region": round(loss_dev, 4),
             "native_lib": os.path.relpath(native.LIB_PATH, ROOT)}
    _emit(args, rank, world, "ours", dev_ms, e2e_s, clocks, launches, DTYPE_LABEL[args.dtype], extra)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


DTYPE_LABEL = {
    # fp32-equivalent storage; conv GEMMs from split hi+lo bf16 operands (3 MMAs per K step, 16-bit
    # significand >= TF32's 11, fp32 accumulate); conv1 fwd TF32; linear layer fp32 FMA.  Error vs an
    # fp64 oracle <= the reference path's on every op: profiles/precision.md
    "bf16x3": "fp32 (bf16x3 split tensor-core operands >= TF32, fp32 accumulate; fp32 linear)",
    "bf16": "bf16",
}


def _replica_checks(eng, rank, world, device, B, args):
    """Outside every timed region: (1) all ranks hold bit-identical parameters after the run;
    (2) DP equivalence — two more steps of the N-rank engine against the SAME engine class run on one
    GPU over the concatenated global batch, from the same parameters; max |delta parameter|."""
    import torch
    import torch.distributed as dist
    from ddp_b200.parallel.engine import B200DataParallel
    eng.synchronize()
    flat = eng.flat_params.detach().clone()
    identical = True
    if world > 1:
        everyone = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(everyone, flat)
        identical = all(torch.equal(everyone[0], t) for t in everyone[1:])
    p0 = flat.clone()
    steps = 2
    for i in range(steps):
        xs, ys = _make_pool(B, pinned=False, rank=rank)
        eng.step(xs[i].to(device), ys[i].to(device))
    eng.synchronize()
    after_n = eng.flat_params.detach().clone()
    solo = B200DataParallel(0, 1, device, batch_cap=B * world, lr=0.01, use_graph=False, precision=args.dtype)
    solo.flat_params.copy_(p0)
    solo.model.pack_weights()
    pools = [_make_pool(B, pinned=False, rank=r) for r in range(world)]
    for i in range(steps):
        xg = torch.cat([pools[r][0][i] for r in range(world)]).to(device)
        yg = torch.cat([pools[r][1][i] for r in range(world)]).to(device)
        solo.step(xg, yg)
    solo.synchronize()
    n = solo.model.arena.numel
    err = float((solo.flat_params[:n] - after_n[:n]).abs().max())
    upd = float((after_n[:n] - p0[:n]).abs().max())
    solo.close(sync_peers=False)
    return {"replicas_bit_identical": bool(identical), "dp_equiv_max_err": err,
            "dp_equiv_max_update": upd, "dp_equiv_steps": steps}


# --------------------------------------------------------------------------- reference arm
def run_reference(args):
    if not os.path.exists(os.path.join(REF_DIR, "model.py")):
        print(json.dumps({"impl": "reference",
                          "unavailable": "baseline/_ref missing (pip install of /root/reference not done)"}))
        return
    rank, world, local = _dist_env(args)
    sys.path[:] = [p for p in sys.path if os.path.abspath(p or ".") != ROOT]
    sys.path.insert(0, REF_DIR)
    import torch
    import torch.nn as nn
    import torch.optim as optim
    from torch.nn.parallel import DistributedDataParallel as DDP
    import model as ref_model   # reference model.py, unmodified
    import utils as ref_utils   # reference utils.py, unmodified
    assert os.path.abspath(ref_model.__file__).startswith(REF_DIR), ref_model.__file__
    assert os.path.abspath(ref_utils.__file__).startswith(REF_DIR), ref_utils.__file__
    os.environ.setdefault("LOCAL_RANK", str(local))
    with contextlib.redirect_stdout(sys.stderr):
        ref_utils.setup(rank=rank, world_size=world)      # reference utils.py:5-14 (nccl, set_device(rank))
    device = torch.device("cuda", torch.cuda.current_device())
    B, K, W = args.batch_size, args.steps, max(3, args.warmup)
    torch.manual_seed(0)
    # reference train_ddp.py:33-42 with the device-placement edits of BASELINE.md §3
    model = ref_model.SimpleCNN().to(device)
    model = DDP(model, device_ids=[device.index])
    loss_fn = nn.CrossEntropyLoss()
    opt = optim.SGD(model.parameters(), lr=0.01)
    model.train()

    def step(images, labels):                              # reference train_ddp.py:196-200
        opt.zero_grad()
        output = model(images)
        loss = loss_fn(output, labels)
        loss.backward()
        opt.step()
        return loss

    xs_h, ys_h = _make_pool(B, pinned=True, rank=rank)
    xs_d, ys_d = xs_h.to(device), ys_h.to(device)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=device)
    for i in range(W):
        step(xs_d[i % POOL], ys_d[i % POOL])
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    stops = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    sampler = ClockSampler(_physical_gpu_index(local))
    _barrier_sync(world)
    sampler.start()
    for i in range(K):
        flush.zero_()
        starts[i].record()
        loss = step(xs_d[(W + i) % POOL], ys_d[(W + i) % POOL])
        stops[i].record()
    _barrier_sync(world)
    clocks = sampler.stop()
    dev_ms = _max_over_ranks(sum(s.elapsed_time(e) for s, e in zip(starts, stops)), world, device)

    loss_host = torch.zeros(K, dtype=torch.float32).pin_memory()
    for i in range(W):
        step(xs_h[i % POOL].to(device, non_blocking=True), ys_h[i % POOL].to(device, non_blocking=True))
    _barrier_sync(world)
    t0 = time.perf_counter()
    for i in range(K):
        images = xs_h[i % POOL].to(device, non_blocking=True)
        labels = ys_h[i % POOL].to(device, non_blocking=True)
        loss = step(images, labels)
        loss_host[i:i + 1].copy_(loss.detach().reshape(1), non_blocking=True)
    _barrier_sync(world)
    e2e_s = _max_over_ranks(time.perf_counter() - t0, world, device)
    extra = {"graph": False, "allreduce": "torch DDP Reducer + NCCL", "final_loss": round(float(loss_host[-1]), 4),
             "edits": "model.to(device); DDP(device_ids); inputs on device; env rank (BASELINE.md §3)",
             "tf32_conv": bool(torch.backends.cudnn.allow_tf32)}
    _emit(args, rank, world, "reference", dev_ms, e2e_s, clocks, 0, "fp32 (cuDNN TF32 conv, default)", extra)
    with contextlib.redirect_stdout(sys.stderr):
        ref_utils.cleanup()


# --------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--batch_size", type=int, default=64, help="per-rank batch (BASELINE.json config 2: 64; config 3: 4096)")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no_graph", action="store_true", help="ours: launch kernels eagerly instead of replaying a CUDA graph")
    ap.add_argument("--algo", type=int, default=0, help="ours: all-reduce algorithm 0=auto 1=one-shot 2=two-shot")
    ap.add_argument("--dtype", choices=["bf16x3", "bf16"], default="bf16x3",
                    help="ours: bf16x3 = the reference's precision or better (default, the headline); "
                         "bf16 = fast mode, reported on its own, never the default")
    args = ap.parse_args()
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        port = 29500 + (os.getpid() % 400)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
