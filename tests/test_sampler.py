"""Sharding parity: ShardedSampler must reproduce torch's DistributedSampler index-for-index
(reference data.py:16-19; SURVEY.md §4.2 "Unit, CPU: sharding parity", §5.7)."""
import pytest
import torch
from torch.utils.data import DistributedSampler

from ddp_b200.parallel.sampler import ShardedSampler


class _Len:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


@pytest.mark.parametrize("n", [1, 7, 10, 64, 1000, 60000])
@pytest.mark.parametrize("ws", [1, 2, 3, 8])
@pytest.mark.parametrize("shuffle", [True, False])
@pytest.mark.parametrize("drop_last", [False, True])
def test_matches_distributed_sampler(n, ws, shuffle, drop_last):
    for rank in range(ws):
        ref = DistributedSampler(_Len(n), num_replicas=ws, rank=rank, shuffle=shuffle,
                                 drop_last=drop_last)
        ours = ShardedSampler(n, ws, rank, shuffle=shuffle, drop_last=drop_last)
        assert len(ours) == len(ref)
        for epoch in (0, 1, 5):
            ref.set_epoch(epoch)
            ours.set_epoch(epoch)
            assert list(ours) == list(ref)
            assert ours.indices().tolist() == list(ref)


def test_seed_and_epoch_change_permutation():
    a = ShardedSampler(100, 2, 0)
    first = a.indices().clone()
    a.set_epoch(1)
    assert not torch.equal(first, a.indices())
    b = ShardedSampler(100, 2, 0, seed=3)
    assert not torch.equal(first, b.indices())


def test_shards_partition_the_epoch():
    n, ws = 1003, 4
    seen = torch.cat([ShardedSampler(n, ws, r).indices() for r in range(ws)])
    assert seen.numel() == 1004  # padded by wrapping from the head
    assert set(seen.tolist()) == set(range(n))


def test_bad_rank():
    with pytest.raises(ValueError):
        ShardedSampler(10, 2, 2)


def test_load_mnist_names_its_source_and_warns_on_the_synthetic_fallback(tmp_path, capsys):
    """reference data.py:11-14 trains on MNIST (download=True).  Offline there is a synthetic stand-in,
    but never silently: the loader prints which source it used and warns loudly on the fallback."""
    import warnings

    import pytest
    import torch
    from ddp_b200.utils.data import load_mnist, synthetic_mnist, write_mnist_idx
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        ds = load_mnist(root=str(tmp_path), synthetic_size=64)
    assert ds.source.startswith("synthetic") and len(ds) == 64
    assert any("SYNTHETIC" in str(x.message) and "not on MNIST" in str(x.message) for x in w)
    assert "WARNING" in capsys.readouterr().out
    with pytest.raises(FileNotFoundError):
        load_mnist(root=str(tmp_path), synthetic_fallback=False)
    # real IDX files where torchvision keeps them: used, and announced
    write_mnist_idx(str(tmp_path), synthetic_mnist(n=32, seed=3))
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        ds = load_mnist(root=str(tmp_path))
    assert not w and len(ds) == 32 and "train-images-idx3-ubyte" in ds.source
    assert "32 images from" in capsys.readouterr().out
    assert torch.equal(ds.images, synthetic_mnist(n=32, seed=3).images)
