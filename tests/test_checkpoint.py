"""Checkpoint contract (SURVEY.md §4.2 "Unit, CPU: checkpoint contract", §5.4): what we write
is what the reference writes, and we load what the reference shipped."""
import os
import time
from collections import OrderedDict

import pytest
import torch

from ddp_b200.models.simple_cnn import SimpleCNN
from ddp_b200.utils import checkpoint as ck

REF_CKPTS = "/root/reference/checkpoints"


def _expected_param_groups():
    return [{"lr": 0.01, "momentum": 0, "dampening": 0, "weight_decay": 0, "nesterov": False,
             "maximize": False, "foreach": None, "differentiable": False, "fused": None,
             "params": [0, 1, 2, 3, 4, 5]}]


def test_saved_file_matches_contract(tmp_path):
    torch.manual_seed(0)
    m = SimpleCNN()
    path = ck.save_checkpoint(3, m.state_dict(), ck.sgd_state_dict(0.01, 6), str(tmp_path))
    assert os.path.basename(path) == "epoch_3.pt"
    loaded = torch.load(path, weights_only=True)  # only tensors / OrderedDict / builtins inside
    assert list(loaded.keys()) == ["epoch", "model", "optimizer"]
    assert loaded["epoch"] == 3 and isinstance(loaded["epoch"], int)
    sd = loaded["model"]
    assert isinstance(sd, OrderedDict)
    assert list(sd.keys()) == ["net.0.weight", "net.0.bias", "net.2.weight", "net.2.bias",
                               "fl.weight", "fl.bias"]
    shapes = [(32, 1, 3, 3), (32,), (64, 32, 3, 3), (64,), (10, 50176), (10,)]
    for (k, v), shp in zip(sd.items(), shapes):
        assert tuple(v.shape) == shp and v.dtype == torch.float32 and v.is_contiguous()
        assert v.device.type == "cpu" and not v.requires_grad
        assert not k.startswith("module.")
    assert list(sd._metadata.keys()) == ["", "net", "net.0", "net.1", "net.2", "net.3", "net.4", "fl"]
    assert all(v == {"version": 1} for v in sd._metadata.values())
    assert loaded["optimizer"] == {"state": {}, "param_groups": _expected_param_groups()}
    # same dict torch's own SGD emits
    opt = torch.optim.SGD([torch.nn.Parameter(torch.zeros(1)) for _ in range(6)], lr=0.01)
    assert loaded["optimizer"] == opt.state_dict()
    assert abs(os.path.getsize(path) - 2085397) < 4096  # reference files are 2,085,397 B


def test_torch_module_loads_our_checkpoint(tmp_path):
    """A stock nn.Module with the reference's structure loads our state_dict strictly."""
    import torch.nn as nn
    torch.manual_seed(1)
    m = SimpleCNN()
    path = ck.save_checkpoint(0, m.state_dict(), ck.sgd_state_dict(0.01, 6), str(tmp_path))

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.net = nn.Sequential(nn.Conv2d(1, 32, 3, padding=1), nn.ReLU(),
                                     nn.Conv2d(32, 64, 3, padding=1), nn.ReLU(), nn.Flatten())
            self.fl = nn.Linear(50176, 10)

        def forward(self, x):
            return self.fl(self.net(x))

    net = Net()
    net.load_state_dict(torch.load(path, weights_only=True)["model"], strict=True)
    x = torch.rand(3, 1, 28, 28)
    assert torch.allclose(net(x), m(x), atol=1e-5)
    opt = torch.optim.SGD(net.parameters(), lr=0.5)
    opt.load_state_dict(torch.load(path, weights_only=True)["optimizer"])
    assert opt.param_groups[0]["lr"] == 0.01


@pytest.mark.skipif(not os.path.isdir(REF_CKPTS), reason="reference checkpoints not mounted")
@pytest.mark.parametrize("name,epoch", [("epoch_0.pt", 0), ("epoch_1.pt", 1)])
def test_load_reference_checkpoints(name, epoch):
    c = ck.load_checkpoint(os.path.join(REF_CKPTS, name))
    assert c["epoch"] == epoch
    m = SimpleCNN(init=False)
    m.load_state_dict(c["model"])
    for k, v in c["model"].items():
        assert torch.equal(m.arena.view(k).detach(), v)
    # DDP-prefixed keys (what the reference's resume path builds, train_ddp.py:182) load too
    m2 = SimpleCNN(init=False)
    m2.load_state_dict({"module." + k: v for k, v in c["model"].items()})
    assert torch.equal(m2.arena.params.detach(), m.arena.params.detach())
    out = m(torch.rand(2, 1, 28, 28))
    assert out.shape == (2, 10) and torch.isfinite(out).all()


def test_strict_loading_errors():
    m = SimpleCNN()
    sd = m.state_dict()
    del sd["fl.bias"]
    with pytest.raises(RuntimeError, match="Missing key"):
        m.load_state_dict(sd)
    sd = m.state_dict()
    sd["fl.weight"] = torch.zeros(10, 5)
    with pytest.raises(RuntimeError, match="size mismatch"):
        m.load_state_dict(sd)


def test_find_latest(tmp_path):
    d = str(tmp_path / "checkpoints")
    assert ck.find_latest(d) is None
    os.makedirs(d)
    assert ck.find_latest(d) is None
    for e in (0, 1):
        with open(os.path.join(d, f"epoch_{e}.pt"), "wb") as f:
            f.write(b"x")
        time.sleep(0.02)
    with open(os.path.join(d, ".DS_Store"), "wb") as f:  # newest file, but not a checkpoint
        f.write(b"junk")
    assert ck.find_latest(d).name == "epoch_1.pt"
    time.sleep(0.02)
    os.utime(os.path.join(d, "epoch_0.pt"))          # mtime does not matter, ctime does
    with open(os.path.join(d, "epoch_0.pt"), "ab") as f:
        f.write(b"y")                               # rewrite -> newest ctime wins, as in the reference
    assert ck.find_latest(d).name == "epoch_0.pt"


def test_atomic_save_leaves_no_temp(tmp_path):
    m = SimpleCNN()
    ck.save_checkpoint(0, m.state_dict(), ck.sgd_state_dict(0.01, 6), str(tmp_path))
    assert sorted(os.listdir(tmp_path)) == ["epoch_0.pt"]
