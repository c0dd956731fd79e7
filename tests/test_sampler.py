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
