"""CPU / gloo integration (BASELINE.json config 1): launch modes, DP
equivalence, save + resume, and agreement with the unmodified reference modules."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
HAVE_REF = os.path.exists(os.path.join(REF, "train_ddp.py"))


def _env(**kw):
    e = {k: v for k, v in os.environ.items() if k not in ("MASTER_ADDR", "MASTER_PORT", "RANK", "WORLD_SIZE", "LOCAL_RANK")}
    e.update({"DDP_B200_SYNTHETIC_SIZE": "256", "OMP_NUM_THREADS": "2", "CUDA_VISIBLE_DEVICES": ""})
    e.update(kw)
    return e


def _run(cmd, cwd, env, timeout=600):
    res = subprocess.run(cmd, cwd=cwd, env=env, capture_output=True, text=True, timeout=timeout)
    assert res.returncode == 0, (res.stdout[-3000:], res.stderr[-3000:])
    return res.stdout


def _free_port():
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_self_spawn_trains_saves_and_resumes(tmp_path):
    """`python train_ddp.py --world_size 2` with NO rendezvous variables set (the reference needs
    them exported); rerunning resumes at the next epoch."""
    cmd = [sys.executable, os.path.join(ROOT, "train_ddp.py"), "--batch_size", "64", "--world_size", "2"]
    out = _run(cmd + ["--epochs", "1"], str(tmp_path), _env())
    assert out.count("has initialized its process group with world size 2") == 2
    assert out.count("No checkpoint found, starting from scratch.") == 2
    assert "Epoch 0 | Batch 0 | Loss:" in out and out.count("cleaned up.") == 2
    p0 = tmp_path / "checkpoints" / "epoch_0.pt"
    assert p0.exists()
    out = _run(cmd + ["--epochs", "2"], str(tmp_path), _env())
    assert out.count("Starting epoch 1") == 2 and "Starting epoch 0" not in out
    assert (tmp_path / "checkpoints" / "epoch_1.pt").exists()
    a = torch.load(p0, weights_only=True)
    b = torch.load(tmp_path / "checkpoints" / "epoch_1.pt", weights_only=True)
    assert a["epoch"] == 0 and b["epoch"] == 1
    assert not torch.equal(a["model"]["fl.weight"], b["model"]["fl.weight"])
    # nothing left to do: loop body never runs, process just cleans up (reference train_ddp.py:192)
    out = _run(cmd + ["--epochs", "2"], str(tmp_path), _env())
    assert "Starting epoch" not in out and out.count("cleaned up.") == 2


def test_torchrun_does_not_double_spawn(tmp_path):
    """README headline command; the reference launches 4 processes here."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "train_ddp.py"), "--epochs", "1", "--batch_size", "64"]
    out = _run(cmd, str(tmp_path), _env())
    assert out.count("has initialized its process group with world size 2") == 2
    assert out.count("Starting epoch 0") == 2
    assert (tmp_path / "checkpoints" / "epoch_0.pt").exists()


def test_single_process(tmp_path):
    cmd = [sys.executable, os.path.join(ROOT, "train_ddp.py"), "--epochs", "1", "--batch_size", "32",
           "--world_size", "1"]
    out = _run(cmd, str(tmp_path), _env(DDP_B200_SYNTHETIC_SIZE="64"))
    assert "world size 1" in out and (tmp_path / "checkpoints" / "epoch_0.pt").exists()


def test_resume_from_reference_shipped_checkpoint(tmp_path):
    src = "/root/reference/checkpoints/epoch_1.pt"
    if not os.path.exists(src):
        pytest.skip("reference checkpoints not mounted")
    os.makedirs(tmp_path / "checkpoints")
    with open(src, "rb") as f, open(tmp_path / "checkpoints" / "epoch_1.pt", "wb") as g:
        g.write(f.read())
    cmd = [sys.executable, os.path.join(ROOT, "train_ddp.py"), "--epochs", "3", "--batch_size", "64",
           "--world_size", "2"]
    out = _run(cmd, str(tmp_path), _env(DDP_B200_SYNTHETIC_SIZE="128"))
    assert out.count("Starting epoch 2") == 2 and "Starting epoch 1" not in out
    assert (tmp_path / "checkpoints" / "epoch_2.pt").exists()


def test_dp_equivalence_cpu(tmp_path):
    """ws=2, per-rank batch b == ws=1, batch 2b on the concatenated shards; replicas end
    bit-identical; rank 0's initial weights win."""
    from torch.multiprocessing.spawn import spawn
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import cpu_workers
    from ddp_b200.models.simple_cnn import SimpleCNN
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(_free_port())
    out = str(tmp_path / "dp")
    steps, b, seed = 4, 8, 11
    spawn(cpu_workers.dp_worker, args=(2, out, steps, b, seed), nprocs=2)
    r0, r1 = torch.load(out + ".0"), torch.load(out + ".1")
    assert torch.equal(r0["params"], r1["params"])
    single = SimpleCNN(init=False)
    with torch.no_grad():
        single.arena.params.copy_(r0["init"])
    g = torch.Generator().manual_seed(seed)
    xs = torch.rand(steps, 2, b, 1, 28, 28, generator=g)
    ys = torch.randint(0, 10, (steps, 2, b), generator=g)
    for it in range(steps):
        single.forward_backward(xs[it].reshape(-1, 1, 28, 28), ys[it].reshape(-1))
        single.sgd_step(0.05)
    diff = (single.arena.params.detach() - r0["params"]).abs().max().item()
    assert diff < 2e-4, diff  # fp32 summation order differs between one 2b batch and two b batches


@pytest.mark.skipif(not HAVE_REF, reason="baseline/_ref not installed")
def test_matches_reference_step_for_step(tmp_path):
    """Identical synthetic IDX files + identical seed: our gloo path and the reference's own
    modules (model/data/utils + DDP + SGD) must produce the same per-step losses and weights."""
    from torch.multiprocessing.spawn import spawn
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import cpu_workers
    from ddp_b200.utils.data import synthetic_mnist, write_mnist_idx
    ds = synthetic_mnist(n=200, seed=3)
    write_mnist_idx(str(tmp_path / "data"), ds, train=True)
    write_mnist_idx(str(tmp_path / "data"), synthetic_mnist(n=16, seed=4), train=False)
    env = _env(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()))
    ref_out = str(tmp_path / "ref.pt")
    _run([sys.executable, os.path.join(ROOT, "baseline", "ref_oracle.py"), "loop", ref_out, "2", "32", "7"],
         str(tmp_path), env)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(_free_port())
    ours_out = str(tmp_path / "ours.pt")
    spawn(cpu_workers.ours_loop_worker, args=(2, ours_out, 2, 32, 7, str(tmp_path / "data")), nprocs=2)
    ref, ours = torch.load(ref_out), torch.load(ours_out)
    assert len(ref["losses"]) == len(ours["losses"]) == 8     # 100 samples/rank, b=32 -> 4 steps x 2 epochs
    for a, b in zip(ref["losses"], ours["losses"]):
        assert abs(a - b) < 1e-4, (ref["losses"], ours["losses"])
    for k, v in ref["model"].items():
        assert torch.allclose(v, ours["model"][k], atol=1e-5), k


@pytest.mark.skipif(not HAVE_REF, reason="baseline/_ref not installed")
def test_patched_reference_resumes_from_our_checkpoint(tmp_path):
    """Cross-load: the reference (with the 4 one-line resume fixes, oracle only) loads a
    checkpoint written by us and continues at epoch+1."""
    from ddp_b200.models.simple_cnn import SimpleCNN
    from ddp_b200.utils import checkpoint as ck
    from ddp_b200.utils.data import synthetic_mnist, write_mnist_idx
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_oracle
    work = tmp_path / "run"
    os.makedirs(work)
    ref_oracle.write_patched(str(tmp_path / "patched"))
    write_mnist_idx(str(work / "data"), synthetic_mnist(n=64, seed=3), train=True)
    write_mnist_idx(str(work / "data"), synthetic_mnist(n=16, seed=4), train=False)
    torch.manual_seed(0)
    ck.save_checkpoint(4, SimpleCNN().state_dict(), ck.sgd_state_dict(0.01, 6), str(work / "checkpoints"))
    env = _env(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()))
    out = _run([sys.executable, str(tmp_path / "patched" / "train_ddp.py"), "--epochs", "6", "--batch_size", "32"],
               str(work), env)
    assert out.count("Starting epoch 5") == 2 and "Starting epoch 4" not in out
    assert (work / "checkpoints" / "epoch_5.pt").exists()


def test_fd_exchange_between_rank_processes(tmp_path):
    """The descriptor hand-over used by the VMM/multicast communicator (AF_UNIX + SCM_RIGHTS),
    exercised on CPU with pipes standing in for the exported allocations."""
    from torch.multiprocessing.spawn import spawn
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import cpu_workers
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(_free_port())
    out = str(tmp_path / "fds")
    spawn(cpu_workers.fd_exchange_worker, args=(3, out), nprocs=3)
    for rank in range(3):
        seen = dict(eval(open(f"{out}.{rank}").read()))
        for src in range(3):
            if src != rank:
                assert seen[str(src)] == f"hello from {src}"
        assert (seen.get("mc") == "mc") == (rank != 0)


def test_replica_mismatch_is_reported_at_construction(tmp_path):
    """ranks that would run different collective sequences must fail fast."""
    from torch.multiprocessing.spawn import spawn
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import cpu_workers
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(_free_port())
    out = str(tmp_path / "verify")
    spawn(cpu_workers.verify_replicas_worker, args=(2, out), nprocs=2)
    m0, m1 = open(out + ".0").read(), open(out + ".1").read()
    assert "replica mismatch with rank 1" in m0 and "tail_update" in m0
    assert "replica mismatch with rank 0" in m1


def test_launcher_restarts_after_a_rank_dies_and_resumes(tmp_path):
    """the reference's only recovery is "rerun the command"; with --max_restarts the
    launcher does it.  Rank 1 is made to die once at the start of epoch 1: the group is restarted,
    resumes from epoch_0.pt and finishes both epochs; without restarts the same failure is fatal."""
    cmd = [sys.executable, os.path.join(ROOT, "train_ddp.py"), "--batch_size", "64", "--world_size", "2",
           "--epochs", "2"]
    env = _env()
    marker = tmp_path / "failed_once"
    env["DDP_B200_FAIL_ONCE"] = f"1:1:{marker}"
    out = _run(cmd + ["--max_restarts", "1"], str(tmp_path), env)
    assert marker.exists()
    # (which rank's exception the parent sees first is a race: the injected one or its peer's broken pipe)
    assert "restart 1/1" in out
    assert out.count("Starting epoch 0") == 2            # first attempt only
    assert out.count("Resumed from checkpoint, next epoch 1") == 2
    assert (tmp_path / "checkpoints" / "epoch_1.pt").exists()
    # the same kind of failure with no restarts allowed is fatal
    fresh = tmp_path / "fresh"
    fresh.mkdir()
    env["DDP_B200_FAIL_ONCE"] = f"0:0:{fresh / 'failed_once'}"
    res = subprocess.run(cmd, cwd=str(fresh), env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode != 0 and "injected failure" in (res.stdout + res.stderr)
