"""Per-rank sharding equivalent to torch.utils.data.DistributedSampler as the reference uses it
(reference data.py:16-19: shuffle=True, seed=0, drop_last=False).

Index-for-index identical to DistributedSampler (tests/test_sampler.py) but works on a length,
not a Dataset, and can hand the whole epoch's shard out as one tensor so a device-resident
dataset can be gathered by a kernel instead of per-sample Python.
"""
from __future__ import annotations

import math

import torch


class ShardedSampler:
    def __init__(self, dataset_len: int, num_replicas: int, rank: int, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False):
        if not 0 <= rank < num_replicas:
            raise ValueError(f"invalid rank {rank} for {num_replicas} replicas")
        self.n = int(dataset_len)
        self.num_replicas = num_replicas
        self.rank = rank
        self.shuffle = shuffle
        self.seed = seed
        self.drop_last = drop_last
        self.epoch = 0
        if drop_last and self.n % num_replicas != 0:
            self.num_samples = math.ceil((self.n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(self.n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch: int) -> None:
        """Changes the permutation; called once per epoch (reference train_ddp.py:193)."""
        self.epoch = int(epoch)

    def indices(self) -> torch.Tensor:
        """int64 [num_samples]: this rank's sample indices for the current epoch."""
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g)
        else:
            idx = torch.arange(self.n)
        if not self.drop_last:
            pad = self.total_size - idx.numel()
            if pad > 0:
                if pad <= idx.numel():
                    idx = torch.cat([idx, idx[:pad]])
                else:
                    reps = math.ceil(pad / idx.numel())
                    idx = torch.cat([idx, idx.repeat(reps)[:pad]])
        else:
            idx = idx[: self.total_size]
        return idx[self.rank: self.total_size: self.num_replicas].contiguous()

    def __iter__(self):
        return iter(self.indices().tolist())

    def __len__(self) -> int:
        return self.num_samples
