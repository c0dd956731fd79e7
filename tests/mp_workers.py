"""Multi-rank GPU scenarios, launched by tests/test_multi_gpu.py under torch.distributed.run
(one rank per GPU).  Each scenario asserts internally; a failing rank exits non-zero.

    python -m torch.distributed.run --nproc-per-node N tests/mp_workers.py <scenario>
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist


def _init():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    return rank, world, dev


def _rank_data(seed, r, n, dtype, dev):
    g = torch.Generator(device="cpu").manual_seed(seed * 131 + r)
    return (torch.randn(n, generator=g) * (r + 1)).to(dtype).to(dev)


def scenario_allreduce():
    """Bit-compare the P2P all-reduce with a host-ordered sum (rank 0..W-1), both algorithms,
    fp32 and bf16, 4 KB .. 64 MB (SURVEY.md §4.2 "all-reduce kernel")."""
    from ddp_b200.parallel.comm import ALGO_ONESHOT, ALGO_TWOSHOT, SymmetricComm
    rank, world, dev = _init()
    max_bytes = 64 << 20
    comm = SymmetricComm(rank, world, 2 * max_bytes)
    for dtype, esize in ((torch.float32, 4), (torch.bfloat16, 2)):
        for nbytes in (4096, 75264, 2007088, 16 << 20, max_bytes):
            n = nbytes // esize
            src = comm.tensor(0, n, dtype)
            dst = comm.tensor(max_bytes, n, dtype)
            for it, algo in enumerate((ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_ONESHOT, ALGO_TWOSHOT)):
                seed = nbytes % 9973 + it
                src.copy_(_rank_data(seed, rank, n, dtype, dev))
                dst.zero_()
                comm.allreduce(0, max_bytes, nbytes, dtype, algo=algo, slot=2)
                torch.cuda.synchronize()
                want = _rank_data(seed, 0, n, dtype, dev)
                for r in range(1, world):
                    want = (want.float() + _rank_data(seed, r, n, dtype, dev).float()).to(dtype)
                assert torch.equal(dst, want), (str(dtype), nbytes, algo, (dst.float() - want.float()).abs().max().item())
    comm.close()
    dist.barrier()
    if rank == 0:
        print("allreduce ok", flush=True)


def scenario_fused_sgd():
    """Fused all-reduce + /W + SGD vs p - lr*mean(g); every rank must end bit-identical."""
    from ddp_b200.parallel.comm import ALGO_ONESHOT, ALGO_TWOSHOT, SymmetricComm
    rank, world, dev = _init()
    n = 501772
    region = n * 4 + 256 - (n * 4) % 256
    comm = SymmetricComm(rank, world, 2 * region)
    p = comm.tensor(0, n)
    gr = comm.tensor(region, n)
    lr = torch.tensor([0.01], device=dev)
    p.copy_(_rank_data(7, 0, n, torch.float32, dev))          # identical on all ranks
    want = p.clone()
    for it, algo in enumerate((ALGO_ONESHOT, ALGO_TWOSHOT) * 3):
        gr.copy_(_rank_data(100 + it, rank, n, torch.float32, dev))
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce_sgd(region, 0, n, lr, algo=algo, slot=it % 2)
        torch.cuda.synchronize()
        s = _rank_data(100 + it, 0, n, torch.float32, dev)
        for r in range(1, world):
            s = s + _rank_data(100 + it, r, n, torch.float32, dev)
        want = want - 0.01 * (s / world)
        assert torch.allclose(p, want, rtol=1e-5, atol=1e-6), (it, (p - want).abs().max().item())
        gathered = [torch.empty_like(p) for _ in range(world)]
        dist.all_gather(gathered, p.clone())
        for r in range(world):
            assert torch.equal(gathered[r], gathered[0]), f"rank {r} diverged bitwise at it {it}"
        dist.barrier()
    # split-K reduce + all-reduce + SGD in one kernel (the conv1 bucket path of the engine)
    from ddp_b200.parallel.comm import ALGO_PUSH
    n2, nparts = 320, 37
    p2 = comm.tensor(0, n2)
    g2 = comm.tensor(region, n2)

    def parts_of(seed, r):
        return _rank_data(seed, r, nparts * n2, torch.float32, dev).view(nparts, n2).contiguous()

    for ai, algo in enumerate((ALGO_PUSH, ALGO_ONESHOT)):
        for it in range(3):
            p2.copy_(_rank_data(50 + it, 0, n2, torch.float32, dev))
            start = p2.clone()
            parts = parts_of(200 + it, rank)
            torch.cuda.synchronize()
            dist.barrier()
            comm.reduce_allreduce_sgd(parts, nparts, n2, region, 0, lr, algo=algo, slot=3 + ai)
            torch.cuda.synchronize()
            assert torch.allclose(g2, parts.sum(0), rtol=1e-5, atol=1e-5), "local gradient not published"
            tot = torch.zeros(n2, device=dev)
            for r in range(world):
                tot += parts_of(200 + it, r).sum(0)
            want2 = start - 0.01 * tot / world
            assert torch.allclose(p2, want2, rtol=1e-4, atol=1e-4), (algo, it, (p2 - want2).abs().max().item())
            gathered = [torch.empty_like(p2) for _ in range(world)]
            dist.all_gather(gathered, p2.clone())
            assert all(torch.equal(t, gathered[0]) for t in gathered), "reduce_allreduce_sgd diverged bitwise"
            dist.barrier()
        # back-to-back calls with no host synchronisation in between: exercises the epoch double
        # buffer of the push variant (ranks run ahead of each other by up to one call)
        reps = 64
        p2.copy_(_rank_data(60, 0, n2, torch.float32, dev))
        start = p2.clone()
        all_parts = [parts_of(300 + k, rank) for k in range(reps)]
        torch.cuda.synchronize()
        dist.barrier()
        for k in range(reps):
            comm.reduce_allreduce_sgd(all_parts[k], nparts, n2, region, 0, lr, algo=algo, slot=3 + ai)
            if rank == k % world:
                torch.cuda._sleep(200000)  # skew the ranks against each other
        torch.cuda.synchronize()
        tot = torch.zeros(n2, device=dev, dtype=torch.float64)
        for k in range(reps):
            for r in range(world):
                tot += parts_of(300 + k, r).sum(0).double()
        want2 = start.double() - 0.01 * tot / world
        assert torch.allclose(p2.double(), want2, rtol=1e-4, atol=1e-3), (algo, (p2 - want2).abs().max().item())
        gathered = [torch.empty_like(p2) for _ in range(world)]
        dist.all_gather(gathered, p2.clone())
        assert all(torch.equal(t, gathered[0]) for t in gathered), "back-to-back calls diverged bitwise"
        dist.barrier()
    # the conv2 bucket path: partials are weight-gradient accumulator images, the kernel permutes
    from ddp_b200.parallel.comm import CONV3X3_WGRAD_IMAGE, LAYOUT_CONV3X3_WGRAD, conv3x3_wgrad_image_dest
    n_src, n_dst, np3 = CONV3X3_WGRAD_IMAGE, 64 * 32 * 9 + 64, 5
    dest = conv3x3_wgrad_image_dest().to(dev)
    keep = dest >= 0
    p3 = comm.tensor(0, n_dst)
    g3 = comm.tensor(region, n_dst)

    def image_of(seed, r):
        return _rank_data(seed, r, np3 * n_src, torch.float32, dev).view(np3, n_src).contiguous()

    p3.copy_(_rank_data(70, 0, n_dst, torch.float32, dev))
    start = p3.clone()
    reps = 6
    imgs = [image_of(400 + k, rank) for k in range(reps)]
    torch.cuda.synchronize()
    dist.barrier()
    for k in range(reps):  # back to back, no host sync
        comm.reduce_allreduce_sgd(imgs[k], np3, n_src, region, 0, lr, algo=ALGO_PUSH,
                                  layout=LAYOUT_CONV3X3_WGRAD, slot=6)
    torch.cuda.synchronize()
    tot = torch.zeros(n_dst, device=dev, dtype=torch.float64)
    for k in range(reps):
        for r in range(world):
            tot[dest[keep]] += image_of(400 + k, r).sum(0).double()[keep]
    want3 = start.double() - 0.01 * tot / world
    assert torch.allclose(p3.double(), want3, rtol=1e-4, atol=1e-3), (p3 - want3).abs().max().item()
    mine = torch.zeros(n_dst, device=dev)
    mine[dest[keep]] = imgs[-1].sum(0)[keep]
    assert torch.allclose(g3, mine, rtol=1e-5, atol=1e-4), "local conv2 gradient not published"
    gathered = [torch.empty_like(p3) for _ in range(world)]
    dist.all_gather(gathered, p3.clone())
    assert all(torch.equal(t, gathered[0]) for t in gathered), "conv3x3-layout update diverged bitwise"
    dist.barrier()
    comm.close()
    if rank == 0:
        print("fused_sgd ok", flush=True)


def scenario_stress():
    """Flag-protocol stress: thousands of back-to-back fused steps with per-rank random delays
    and changing payloads (compute-sanitizer cannot see cross-GPU races)."""
    from ddp_b200.parallel.comm import SymmetricComm
    rank, world, dev = _init()
    n = 18816
    region = 128 << 10
    comm = SymmetricComm(rank, world, 3 * region)
    p = comm.tensor(0, n)
    g = [comm.tensor(region, n), comm.tensor(2 * region, n)]
    lr = torch.tensor([1.0], device=dev)
    p.zero_()
    rng = torch.Generator().manual_seed(rank)
    total = 0.0
    iters = 3000
    for it in range(iters):
        par = it & 1
        g[par].fill_(float((it % 7) + rank))                 # mean over ranks = it%7 + (W-1)/2
        if int(torch.randint(0, 4, (1,), generator=rng)) == 0:
            torch.cuda._sleep(int(torch.randint(1000, 200000, (1,), generator=rng)))
        comm.allreduce_sgd(region * (1 + par), 0, n, lr, algo=1 + (it % 2), slot=it % 2)
        total += (it % 7) + (world - 1) / 2.0
    torch.cuda.synchronize()
    assert torch.allclose(p, torch.full_like(p, -total), rtol=1e-5), (p[:4].tolist(), -total)
    assert comm.status() == 0
    comm.close()
    dist.barrier()
    if rank == 0:
        print("stress ok", flush=True)


def scenario_dp_equiv():
    """ws=k with per-rank batch b == one process with batch k*b on the concatenated shards
    (SURVEY.md §4.2 "DP equivalence"), CUDA graph on, plus a short tail batch."""
    from ddp_b200.models.simple_cnn import SimpleCNN
    from ddp_b200.parallel.engine import B200DataParallel
    rank, world, dev = _init()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    b = 16
    torch.manual_seed(1000 + rank)                  # ranks start DIFFERENT; the engine must sync them
    eng = B200DataParallel(rank, world, dev, batch_cap=b, lr=0.01)
    torch.manual_seed(5)
    oracle = SimpleCNN(device=dev, init=False)
    sd = eng.state_dict()
    oracle.load_state_dict(sd)
    gathered = [torch.empty_like(eng.flat_params) for _ in range(world)]
    dist.all_gather(gathered, eng.flat_params.clone())
    assert all(torch.equal(t, gathered[0]) for t in gathered), "initial broadcast failed"
    g = torch.Generator().manual_seed(99)
    xs = torch.rand(12, world, b, 1, 28, 28, generator=g).to(dev)
    ys = torch.randint(0, 10, (12, world, b), generator=g).to(dev)
    for it in range(12):
        bb = b if it % 4 != 3 else 9
        lo = oracle.forward_backward(xs[it, :, :bb].reshape(-1, 1, 28, 28), ys[it, :, :bb].reshape(-1))
        oracle.sgd_step(0.01)
        eng.step(xs[it, rank, :bb], ys[it, rank, :bb])
        lg = torch.tensor([eng.read_loss()], device=dev)
        dist.all_reduce(lg)
        assert abs(lg.item() / world - lo.item()) < 2e-3 * max(1.0, abs(lo.item())), (it, lg.item() / world, lo.item())
    eng.synchronize()
    dist.all_gather(gathered, eng.flat_params.clone())
    assert all(torch.equal(t, gathered[0]) for t in gathered), "replicas diverged bitwise"
    diff = (eng.flat_params[:520586] - oracle.arena.params.detach()[:520586]).abs().max().item()
    if rank == 0:
        print(f"dp_equiv: world {world} max |param - fp32 oracle| after 12 steps = {diff:.3e}", flush=True)
    assert diff < 5e-4, diff  # reference-precision (bf16x3) path vs the full-fp32 oracle after 12 steps
    eng.close()
    dist.barrier()
    if rank == 0:
        print("dp_equiv ok", flush=True)


def scenario_bucketed():
    """BucketedDataParallel (widened CNN, several buckets, hooks + side stream) must match plain
    SGD on the concatenated batch and keep the replicas bit-identical."""
    import copy
    import torch.nn.functional as F
    from ddp_b200.models.wide_cnn import WideCNN
    from ddp_b200.parallel.bucketed import BucketedDataParallel
    rank, world, dev = _init()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(3)
    m = WideCNN(hidden=96).to(dev)                      # smaller head keeps the test quick
    ref = copy.deepcopy(m)
    eng = BucketedDataParallel(m, rank, world, lr=0.05, bucket_mb=0.25)
    assert len(eng.buckets) >= 3, len(eng.buckets)
    opt = torch.optim.SGD(ref.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(17)
    b = 8
    xs = torch.rand(4, world, b, 3, 32, 32, generator=g).to(dev)
    ys = torch.randint(0, 10, (4, world, b), generator=g).to(dev)
    for it in range(4):
        eng.step(xs[it, rank], ys[it, rank], F.cross_entropy)
        opt.zero_grad()
        F.cross_entropy(ref(xs[it].reshape(-1, 3, 32, 32)), ys[it].reshape(-1)).backward()
        opt.step()
    torch.cuda.synchronize()
    for (n1, p1), (n2, p2) in zip(m.named_parameters(), ref.named_parameters()):
        d = (p1 - p2).abs().max().item()
        assert d < 2e-4 * max(1.0, p2.abs().max().item()), (n1, d)
    gathered = [torch.empty_like(eng.flat_params) for _ in range(world)]
    dist.all_gather(gathered, eng.flat_params.clone())
    assert all(torch.equal(t, gathered[0]) for t in gathered), "replicas diverged bitwise"
    eng.close()
    dist.barrier()
    if rank == 0:
        print("bucketed ok", flush=True)


def scenario_bucketed_native_graph():
    """Same as bucketed_native with the step replayed from CUDA graphs (captured per step parity)."""
    scenario_bucketed_native(use_graph=True, steps=8, tag="bucketed_native_graph")


def scenario_bucketed_native(use_graph=False, steps=4, tag="bucketed_native"):
    """The widened CNN on the native kernels (models/wide_cnn.py WideCNNB200: tcgen05 linear layers whose
    weight gradient is written straight into the gradient arena, channels-last conv1 / pooling kernels)
    under BucketedDataParallel must match plain torch SGD on the torch model over the concatenated batch,
    and keep the replicas bit-identical."""
    import torch.nn.functional as F
    from ddp_b200.models.wide_cnn import WideCNN, WideCNNB200
    from ddp_b200.parallel.bucketed import BucketedDataParallel
    rank, world, dev = _init()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(3)
    m = WideCNNB200(hidden=96, native=True).to(dev)
    ref = WideCNN(hidden=96).to(dev)
    ref.load_state_dict(m.state_dict())
    eng = BucketedDataParallel(m, rank, world, lr=0.05, bucket_mb=0.25, use_graph=use_graph)
    assert len(eng.buckets) >= 3 and len(eng._direct) == 2, (len(eng.buckets), eng._direct)
    opt = torch.optim.SGD(ref.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(17)
    b = 8
    xs = torch.rand(steps, world, b, 3, 32, 32, generator=g).to(dev)
    ys = torch.randint(0, 10, (steps, world, b), generator=g).to(dev)
    for it in range(steps):
        eng.step(xs[it, rank], ys[it, rank], F.cross_entropy)
        opt.zero_grad()
        F.cross_entropy(ref(xs[it].reshape(-1, 3, 32, 32)), ys[it].reshape(-1)).backward()
        opt.step()
    torch.cuda.synchronize()
    sd, rd = m.state_dict(), ref.state_dict()     # state_dict() brings fc1's columns back to (c, h, w) order
    for k in rd:
        d = (sd[k] - rd[k]).abs().max().item()
        assert d < 2e-4 * max(1.0, rd[k].abs().max().item()), (k, d)
    gathered = [torch.empty_like(eng.flat_params) for _ in range(world)]
    dist.all_gather(gathered, eng.flat_params.clone())
    assert all(torch.equal(t, gathered[0]) for t in gathered), "replicas diverged bitwise"
    if use_graph:
        assert len(eng._graphs) == 2, len(eng._graphs)
    eng.close()
    dist.barrier()
    if rank == 0:
        print(f"{tag} ok", flush=True)


def scenario_nvls():
    """VMM + multicast communicator: descriptor exchange, multicast bind, then the in-switch
    (multimem.ld_reduce / multimem.st) all-reduce and fused all-reduce+SGD.  Prints
    "nvls unavailable" (the test then skips) when the box cannot create multicast objects."""
    import warnings
    from ddp_b200.parallel.comm import ALGO_NVLS, ALGO_TWOSHOT, SymmetricComm
    rank, world, dev = _init()
    n = 1 << 20
    region = n * 4
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter("always")
        comm = SymmetricComm(rank, world, 3 * region, multicast=True)
    if not comm.multicast:
        if rank == 0:
            why = "; ".join(str(w.message) for w in caught) or "no reason given"
            print(f"nvls unavailable: {why}", flush=True)
        # the fallback communicator must still work
        lr = torch.tensor([0.01], device=dev)
        comm.tensor(0, n).fill_(1.0)
        comm.tensor(region, n).fill_(float(rank + 1))
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce_sgd(region, 0, n, lr, algo=ALGO_TWOSHOT, slot=0)
        torch.cuda.synchronize()
        want = 1.0 - 0.01 * sum(range(1, world + 1)) / world
        assert torch.allclose(comm.tensor(0, n), torch.full((n,), want, device=dev), atol=1e-6)
        comm.close()
        return
    # plain all-reduce through the switch, fp32 and bf16, several sizes incl. a ragged one
    for dtype, tol in ((torch.float32, 1e-5), (torch.bfloat16, 2e-2)):
        esize = 4 if dtype == torch.float32 else 2
        for nbytes in (4096, 75264, 1 << 20, region):
            m = nbytes // esize
            src = comm.tensor(0, m, dtype)
            out = comm.tensor(region, m, dtype)
            for it in range(2):
                src.copy_(_rank_data(11 + it, rank, m, dtype, dev))
                torch.cuda.synchronize()
                dist.barrier()
                comm.allreduce(0, region, nbytes, dtype, ALGO_NVLS, slot=1)
                torch.cuda.synchronize()
                want = torch.zeros(m, device=dev, dtype=torch.float32)
                for r in range(world):
                    want += _rank_data(11 + it, r, m, dtype, dev).float()
                err = (out.float() - want).abs().max().item()
                assert err <= tol * max(1.0, want.abs().max().item()), (dtype, nbytes, err)
                gathered = [torch.empty_like(out) for _ in range(world)]
                dist.all_gather(gathered, out.clone())
                assert all(torch.equal(t, gathered[0]) for t in gathered), "replicas differ after NVLS all-reduce"
                dist.barrier()
    # fused all-reduce + 1/W + SGD through the switch
    p = comm.tensor(0, n)
    gr = comm.tensor(region, n)
    lr = torch.tensor([0.01], device=dev)
    p.copy_(_rank_data(7, 0, n, torch.float32, dev))
    want = p.clone()
    for it in range(4):
        gr.copy_(_rank_data(100 + it, rank, n, torch.float32, dev))
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce_sgd(region, 0, n, lr, algo=ALGO_NVLS, slot=2)
        torch.cuda.synchronize()
        tot = torch.zeros(n, device=dev)
        for r in range(world):
            tot += _rank_data(100 + it, r, n, torch.float32, dev)
        want = want - 0.01 * (tot / world)
        assert torch.allclose(p, want, rtol=1e-5, atol=1e-5), (it, (p - want).abs().max().item())
        gathered = [torch.empty_like(p) for _ in range(world)]
        dist.all_gather(gathered, p.clone())
        assert all(torch.equal(t, gathered[0]) for t in gathered), f"replicas differ at it {it}"
        dist.barrier()
    comm.close()
    if rank == 0:
        print("nvls ok", flush=True)


def scenario_dead_peer():
    """Failure detection: the last rank never joins the collective; every other
    rank's kernel must give up after the communicator's timeout and surface a CUDA error on the
    host instead of spinning forever.  The trapped contexts are unusable afterwards, so every
    rank leaves with os._exit."""
    import time
    from ddp_b200.parallel.comm import SymmetricComm
    rank, world, dev = _init()
    n = 4096
    comm = SymmetricComm(rank, world, 2 * n * 4, timeout_ms=300)
    lr = torch.tensor([0.01], device=dev)
    comm.allreduce_sgd(n * 4, 0, n, lr, slot=0)       # a healthy round first
    torch.cuda.synchronize()
    dist.barrier()
    if rank == world - 1:
        time.sleep(6.0)                                # "dead": never launches the second round
        os._exit(0)
    t0 = time.perf_counter()
    failed = False
    try:
        comm.allreduce_sgd(n * 4, 0, n, lr, slot=0)
        torch.cuda.synchronize()
    except Exception as e:  # torch.AcceleratorError / RuntimeError depending on the torch version
        failed = True
        msg = str(e).splitlines()[0][:120]
    dt = time.perf_counter() - t0
    if not failed:
        print(f"rank {rank}: collective with a dead peer returned normally: assert failed", flush=True)
        os._exit(1)
    if dt > 5.0:
        print(f"rank {rank}: timeout took {dt:.1f}s: assert failed", flush=True)
        os._exit(1)
    if rank == 0:
        print(f"dead_peer ok ({dt * 1e3:.0f} ms, {msg})", flush=True)
    os._exit(0)


if __name__ == "__main__":
    {"dead_peer": scenario_dead_peer, "nvls": scenario_nvls, "bucketed": scenario_bucketed, "bucketed_native": scenario_bucketed_native, "bucketed_native_graph": scenario_bucketed_native_graph, "allreduce": scenario_allreduce, "fused_sgd": scenario_fused_sgd, "stress": scenario_stress,
     "dp_equiv": scenario_dp_equiv}[sys.argv[1]]()
    if dist.is_initialized():
        dist.destroy_process_group()
