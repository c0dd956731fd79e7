"""Multi-GPU tests (need >= 2 visible GPUs; skipped otherwise): the peer-memory all-reduce
family, the fused all-reduce+SGD, flag-protocol stress, DP equivalence and the CLI under
torchrun.  Each runs tests/mp_workers.py (or train_ddp.py) under torch.distributed.run."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0
WORLDS = [w for w in (2, 3, 4, 6, 8) if w <= NGPU]  # odd and non-power-of-two sizes included


def _torchrun(world, script_args, cwd=None, env=None, timeout=600):
    port = 29600 + (os.getpid() + world * 7) % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port)] + script_args
    e = dict(os.environ)
    e.update(env or {})
    res = subprocess.run(cmd, cwd=cwd or ROOT, env=e, capture_output=True, text=True, timeout=timeout)
    if res.returncode != 0:
        # keep the interesting part: the first exception of any rank, not torchrun's epilogue
        lines = [l for l in (res.stdout + "\n" + res.stderr).splitlines()
                 if any(k in l for k in ("Error", "assert", "Assert", "error:", "diverged", "failed"))]
        pytest.fail("rank failure:\n" + "\n".join(lines[:25])[-4000:], pytrace=False)
    return res.stdout


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("world", WORLDS or [2])
@pytest.mark.parametrize("scenario", ["allreduce", "fused_sgd", "stress", "dp_equiv", "bucketed", "bucketed_native", "bucketed_native_graph"])
def test_scenario(world, scenario):
    out = _torchrun(world, [os.path.join(ROOT, "tests", "mp_workers.py"), scenario])
    assert f"{scenario} ok" in out


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("world", WORLDS or [2])
def test_multicast_communicator(world):
    out = _torchrun(world, [os.path.join(ROOT, "tests", "mp_workers.py"), "nvls"], timeout=300)
    if "nvls unavailable" in out:
        pytest.skip(next(l for l in out.splitlines() if "nvls unavailable" in l)[:300])
    assert "nvls ok" in out


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
def test_dead_peer_times_out_instead_of_hanging():
    out = _torchrun(2, [os.path.join(ROOT, "tests", "mp_workers.py"), "dead_peer"], timeout=120)
    assert "dead_peer ok" in out


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
def test_cli_under_torchrun_trains_and_resumes(tmp_path):
    env = {"DDP_B200_SYNTHETIC_SIZE": "2048"}
    out = _torchrun(2, [os.path.join(ROOT, "train_ddp.py"), "--epochs", "1", "--batch_size", "64"],
                    cwd=str(tmp_path), env=env)
    assert out.count("Starting epoch 0") == 2 and "No checkpoint found" in out
    assert os.path.exists(tmp_path / "checkpoints" / "epoch_0.pt")
    out = _torchrun(2, [os.path.join(ROOT, "train_ddp.py"), "--epochs", "2", "--batch_size", "64"],
                    cwd=str(tmp_path), env=env)
    assert out.count("Starting epoch 1") == 2 and "Starting epoch 0" not in out
    assert os.path.exists(tmp_path / "checkpoints" / "epoch_1.pt")
