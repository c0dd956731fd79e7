"""Worker functions for the CPU (gloo) multi-process tests; importable by spawn."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def dp_worker(rank, world_size, out_path, steps, batch, seed):
    import torch
    from ddp_b200.parallel.engine import CpuDataParallel
    from ddp_b200.parallel.process_group import cleanup, setup
    torch.set_num_threads(2)
    setup(rank, world_size, backend="gloo")
    torch.manual_seed(1000 + rank)           # different init per rank: the engine must sync from rank 0
    eng = CpuDataParallel(rank, world_size, lr=0.05)
    if rank == 0:
        init = eng.flat_params.detach().clone()
    g = torch.Generator().manual_seed(seed)
    xs = torch.rand(steps, world_size, batch, 1, 28, 28, generator=g)
    ys = torch.randint(0, 10, (steps, world_size, batch), generator=g)
    losses = []
    for it in range(steps):
        losses.append(float(eng.step(xs[it, rank], ys[it, rank])))
    torch.save({"params": eng.flat_params.detach().clone(), "losses": losses,
                "init": init if rank == 0 else None}, f"{out_path}.{rank}")
    cleanup()


def ours_loop_worker(rank, world_size, out_path, epochs, batch, seed, data_root):
    """Ours, same schedule as baseline/ref_oracle.py `loop` (for per-step comparison)."""
    import torch
    from ddp_b200.parallel.engine import CpuDataParallel
    from ddp_b200.parallel.process_group import cleanup, setup
    from ddp_b200.utils.data import get_dataloader
    torch.set_num_threads(2)
    setup(rank, world_size, backend="gloo")
    torch.manual_seed(seed)
    eng = CpuDataParallel(rank, world_size, lr=0.01)
    loader, sampler = get_dataloader(batch, world_size, rank, root=data_root)
    losses = []
    for epoch in range(epochs):
        sampler.set_epoch(epoch)
        for x, y in loader:
            losses.append(float(eng.step(x, y)))
    if rank == 0:
        torch.save({"losses": losses, "model": eng.state_dict()}, out_path)
    cleanup()


def fd_exchange_worker(rank, world_size, out_path):
    """Each rank hands every peer the read end of a pipe it has written its rank into."""
    import torch.distributed as dist
    from ddp_b200.parallel.comm import exchange_fds
    from ddp_b200.parallel.process_group import cleanup, setup
    setup(rank, world_size, backend="gloo")
    tok = ["tok-%d" % os.getpid() if rank == 0 else None]
    dist.broadcast_object_list(tok, src=0)
    r, w = os.pipe()
    os.write(w, f"hello from {rank}".encode() * (world_size - 1))
    os.close(w)
    got = exchange_fds(rank, world_size, [r], tok[0])
    os.close(r)
    seen = {}
    for src, fds in enumerate(got):
        assert (len(fds) == 1) == (src != rank), (src, fds)
        for fd in fds:
            seen[src] = os.read(fd, len(f"hello from {src}")).decode()
            os.close(fd)
    # a second round on a fresh token where only rank 0 has something to give (the multicast case)
    r2, w2 = os.pipe()
    os.write(w2, b"mc" * world_size)
    os.close(w2)
    got = exchange_fds(rank, world_size, [r2] if rank == 0 else [], tok[0] + "c")
    os.close(r2)
    for src, fds in enumerate(got):
        assert len(fds) == (1 if (src == 0 and rank != 0) else 0), (rank, src, fds)
        for fd in fds:
            seen["mc"] = os.read(fd, 2).decode()
            os.close(fd)
    with open(f"{out_path}.{rank}", "w") as f:
        f.write(repr(sorted((str(k), v) for k, v in seen.items())))
    cleanup()


def verify_replicas_worker(rank, world_size, out_path):
    """Matching settings pass; a rank configured differently is named in the error on every rank."""
    from ddp_b200.parallel.engine import CpuDataParallel
    from ddp_b200.parallel.process_group import cleanup, setup
    setup(rank, world_size, backend="gloo")
    eng = CpuDataParallel(rank, world_size)           # constructor runs the check with equal settings
    eng.verify_replicas(rank, world_size, tail_update="push")
    try:
        eng.verify_replicas(rank, world_size, tail_update="push" if rank == 0 else "split")
        msg = "no error"
    except RuntimeError as e:
        msg = str(e)
    with open(f"{out_path}.{rank}", "w") as f:
        f.write(msg)
    cleanup()
