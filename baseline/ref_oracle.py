"""Correctness oracle built from the UNMODIFIED reference modules in baseline/_ref
(SURVEY.md §4.3).  Used only by tests and BASELINE.md measurements — never imported by the
product package.

    python baseline/ref_oracle.py loop  <out.pt> <epochs> <batch>     # under MASTER_ADDR/PORT, spawns ws=2
    python baseline/ref_oracle.py patch <dest_dir>                    # write the resume-patched train_ddp.py

`loop` runs reference model.SimpleCNN + data.get_dataloader + utils.setup/cleanup with the
reference's own loop body (train_ddp.py:192-202), seeded so the initial weights are
reproducible, and records rank 0's per-step losses and final weights.
`patch` applies the four one-line fixes without which the reference's resume path crashes
(SURVEY.md §5.4 bugs B1-B3) to a copy of its train_ddp.py.
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")

PATCHES = (
    ("            dist.broadcast(key_len_tensor)\n", "            dist.broadcast(key_len_tensor, src=0)\n"),
    ("            dist.broadcast(shape_tensor)\n", "            dist.broadcast(shape_tensor, src=0)\n"),
    ("        dist.broadcast(num_key_tensor, src=0)\n",
     "        dist.broadcast(num_key_tensor, src=0)\n        num_keys = int(num_key_tensor.item())\n"),
    ("            numel = model_state[key].numel()\n",
     "            numel = int(torch.tensor(shape).prod().item()) if len(shape) else 1\n"),
)


def write_patched(dest_dir: str) -> str:
    with open(os.path.join(REF, "train_ddp.py")) as f:
        src = f.read()
    for old, new in PATCHES:
        assert src.count(old) == 1, f"reference changed? {old!r}"
        src = src.replace(old, new)
    os.makedirs(dest_dir, exist_ok=True)
    for name in ("model.py", "data.py", "utils.py"):
        with open(os.path.join(REF, name)) as f, open(os.path.join(dest_dir, name), "w") as g:
            g.write(f.read())
    path = os.path.join(dest_dir, "train_ddp.py")
    with open(path, "w") as f:
        f.write(src)
    return path


def _loop_worker(rank, world_size, out_path, epochs, batch_size, seed):
    sys.path.insert(0, REF)
    import torch
    import torch.nn as nn
    import torch.optim as optim
    from torch.nn.parallel import DistributedDataParallel as DDP
    from data import get_dataloader
    from model import SimpleCNN
    from utils import cleanup, setup
    torch.set_num_threads(2)
    setup(rank=rank, world_size=world_size)
    torch.manual_seed(seed)
    model = DDP(SimpleCNN())
    dataloader, sampler = get_dataloader(batch_size, world_size, rank)
    loss_fn = nn.CrossEntropyLoss()
    opt = optim.SGD(model.parameters(), lr=0.01)
    model.train()
    losses = []
    for epoch in range(epochs):
        sampler.set_epoch(epoch)
        for batch_idx, (images, labels) in enumerate(dataloader):
            opt.zero_grad()
            output = model(images)
            loss = loss_fn(output, labels)
            loss.backward()
            opt.step()
            losses.append(float(loss.item()))
    if rank == 0:
        torch.save({"losses": losses, "model": model.module.state_dict(),
                    "optimizer": opt.state_dict()}, out_path)
    cleanup()


def main():
    mode = sys.argv[1]
    if mode == "patch":
        print(write_patched(sys.argv[2]))
    elif mode == "loop":
        from torch.multiprocessing.spawn import spawn
        out, epochs, batch = sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
        seed = int(sys.argv[5]) if len(sys.argv) > 5 else 0
        spawn(_loop_worker, args=(2, out, epochs, batch, seed), nprocs=2)
    else:
        raise SystemExit(__doc__)


if __name__ == "__main__":
    main()
