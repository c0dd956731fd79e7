"""All-reduce sweep (BASELINE.json config 5): our peer-memory kernels vs NCCL, 4 KB .. 1 GB, fp32 and
bf16, plus the fused all-reduce + 1/W + SGD kernel vs the reference's sequence (NCCL all-reduce,
divide, foreach SGD — what DDP + torch.optim.SGD run).  Device-timed with CUDA events, max over ranks.

    python -m torch.distributed.run --nproc-per-node 8 tools/allreduce_sweep.py [--max-mb 1024] [--out file.json]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from ddp_b200.parallel.comm import ALGO_NVLS, ALGO_ONESHOT, ALGO_TWOSHOT, SymmetricComm


def timed(fn, iters, world, dev):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(iters):
        fn()
    ev1.record()
    torch.cuda.synchronize()
    t = torch.tensor([ev0.elapsed_time(ev1) * 1e3 / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--out", default=None)
    ap.add_argument("--multicast", action="store_true",
                    help="VMM + multicast communicator: adds the in-switch (multimem) columns")
    a = ap.parse_args()
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    max_bytes = a.max_mb << 20
    comm = SymmetricComm(rank, world, 2 * max_bytes + (1 << 20), multicast=a.multicast)
    nvls = comm.multicast
    if rank == 0:
        print(f"world {world}, multicast (NVLS) {'bound' if nvls else 'not in use'}", flush=True)
    sizes = [4 << 10, 16 << 10, 64 << 10, 75264, 256 << 10, 1 << 20, 2007088, 4 << 20, 16 << 20, 64 << 20,
             256 << 20, 1 << 30]
    sizes = [s for s in sizes if s <= max_bytes]
    rows = []
    for dtype, esize, name in ((torch.float32, 4, "fp32"), (torch.bfloat16, 2, "bf16")):
        for nbytes in sizes:
            n = nbytes // esize
            iters = 50 if nbytes <= (4 << 20) else (10 if nbytes <= (64 << 20) else 4)
            src = comm.tensor(0, n, dtype)
            src.normal_()
            nccl_buf = torch.randn(n, device=dev, dtype=torch.float32).to(dtype)
            row = {"dtype": name, "bytes": nbytes, "world": world}
            row["nccl_us"] = timed(lambda: dist.all_reduce(nccl_buf), iters, world, dev)
            row["oneshot_us"] = timed(lambda: comm.allreduce(0, max_bytes, nbytes, dtype, ALGO_ONESHOT, slot=2),
                                      iters, world, dev)
            row["twoshot_us"] = timed(lambda: comm.allreduce(0, max_bytes, nbytes, dtype, ALGO_TWOSHOT, slot=3),
                                      iters, world, dev)
            best = min(row["oneshot_us"], row["twoshot_us"])
            if nvls:
                row["nvls_us"] = timed(lambda: comm.allreduce(0, max_bytes, nbytes, dtype, ALGO_NVLS, slot=6),
                                       iters, world, dev)
                best = min(best, row["nvls_us"])
            row["best_us"] = best
            # bus bandwidth convention of nccl-tests: 2*(W-1)/W * bytes / time
            row["busbw_gbs_ours"] = 2 * (world - 1) / world * nbytes / best / 1e3
            row["busbw_gbs_nccl"] = 2 * (world - 1) / world * nbytes / row["nccl_us"] / 1e3
            rows.append(row)
            if rank == 0:
                print(f"{name} {nbytes:>11d} B  nccl {row['nccl_us']:9.1f} us | one-shot {row['oneshot_us']:9.1f} | "
                      f"two-shot {row['twoshot_us']:9.1f} | nvls {row.get('nvls_us', float('nan')):9.1f} | busbw ours {row['busbw_gbs_ours']:7.1f} nccl {row['busbw_gbs_nccl']:7.1f} GB/s",
                      flush=True)
    # fused all-reduce + /W + SGD (fp32) vs NCCL all-reduce + div + foreach SGD on the two real buckets and up
    fused = []
    lr_dev = torch.tensor([0.01], device=dev)
    for nbytes in [s for s in (75264, 2007088, 16 << 20, 64 << 20, 256 << 20) if 2 * s <= 2 * max_bytes]:
        n = nbytes // 4
        iters = 50 if nbytes <= (4 << 20) else 10
        p_ref = [torch.randn(n, device=dev)]
        g_ref = [torch.randn(n, device=dev)]

        def reference_sequence():
            dist.all_reduce(g_ref[0])
            g_ref[0].div_(world)
            torch._foreach_add_(p_ref, g_ref, alpha=-0.01)

        row = {"bytes": nbytes, "world": world}
        row["nccl_div_sgd_us"] = timed(reference_sequence, iters, world, dev)
        row["fused_oneshot_us"] = timed(
            lambda: comm.allreduce_sgd(max_bytes, 0, n, lr_dev, ALGO_ONESHOT, slot=4), iters, world, dev)
        row["fused_twoshot_us"] = timed(
            lambda: comm.allreduce_sgd(max_bytes, 0, n, lr_dev, ALGO_TWOSHOT, slot=5), iters, world, dev)
        best = min(row["fused_oneshot_us"], row["fused_twoshot_us"])
        if nvls:
            row["fused_nvls_us"] = timed(
                lambda: comm.allreduce_sgd(max_bytes, 0, n, lr_dev, ALGO_NVLS, slot=7), iters, world, dev)
            best = min(best, row["fused_nvls_us"])
        # roofline: bytes that must cross NVLink per GPU (two-shot: read (W-1)/W*S, write (W-1)/W*S)
        # over the measured 770 GB/s per direction (B200_PROFILING.md)
        wire = (world - 1) / world * nbytes
        row["roofline_us"] = wire / 770e3
        row["fraction_of_roofline"] = row["roofline_us"] / best
        fused.append(row)
        if rank == 0:
            print(f"fused {nbytes:>11d} B  nccl+div+sgd {row['nccl_div_sgd_us']:9.1f} us | fused one-shot "
                  f"{row['fused_oneshot_us']:9.1f} | fused two-shot {row['fused_twoshot_us']:9.1f} | "
                  f"fused nvls {row.get('fused_nvls_us', float('nan')):9.1f} | "
                  f"roofline {row['roofline_us']:7.1f} us ({row['fraction_of_roofline']:.2f})", flush=True)
    if rank == 0 and a.out:
        with open(a.out, "w") as f:
            json.dump({"allreduce": rows, "fused": fused, "multicast": nvls}, f, indent=1)
    comm.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
