"""Sampler index parity with torch, dataset shapes/dtypes (reference C5, C11, C17, N17)."""
import itertools

import pytest
import torch
from torch.utils.data import DataLoader
from torch.utils.data.distributed import DistributedSampler as TorchSampler

from pytorchdistributed_b200.data import (DistributedSampler, MyTrainDataset, SimpleDataset, SyntheticImageDataset,
                                          generate_random_data)


class _Len:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


@pytest.mark.parametrize("n,world,shuffle,drop_last", list(itertools.product([1, 7, 64, 2048, 2049], [1, 2, 3, 8],
                                                                               [True, False], [True, False])))
def test_sampler_matches_torch(n, world, shuffle, drop_last):
    ds = _Len(n)
    for rank in range(world):
        ours = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=3, drop_last=drop_last)
        ref = TorchSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=3, drop_last=drop_last)
        for epoch in (0, 1, 5):
            ours.set_epoch(epoch)
            ref.set_epoch(epoch)
            assert list(ours) == list(ref)
            assert len(ours) == len(ref)


def test_sampler_partitions_without_overlap():
    ds = _Len(2048)
    seen = []
    for rank in range(2):
        s = DistributedSampler(ds, num_replicas=2, rank=rank)
        idx = list(s)
        assert len(idx) == 1024
        seen += idx
    assert sorted(seen) == list(range(2048))


def test_sampler_needs_rank_or_group():
    with pytest.raises(RuntimeError):
        DistributedSampler(_Len(4))


def test_my_train_dataset():
    ds = MyTrainDataset(2048)
    assert len(ds) == 2048
    x, y = ds[5]
    assert x.shape == (20,) and y.shape == (1,) and x.dtype == torch.float32 and y.dtype == torch.float32
    assert 0 <= float(x.min()) and float(x.max()) < 1
    xb, yb = next(iter(DataLoader(ds, batch_size=32)))
    assert xb.shape == (32, 20) and yb.shape == (32, 1)


def test_simple_dataset():
    ds = SimpleDataset(1000)
    xb, yb = next(iter(DataLoader(ds, batch_size=32, shuffle=True)))
    assert xb.shape == (32, 10) and yb.shape == (32, 5)
    assert yb.dtype == torch.int64 and int(yb.abs().sum()) == 0      # randint(0, 1) is all zeros


def test_generate_random_data():
    x, y = generate_random_data()
    assert x.shape == (120, 3, 128, 128) and y.shape == (120, 1000)
    assert y.dtype == torch.float32 and torch.all(y.sum(1) == 1)


def test_synthetic_images():
    ds = SyntheticImageDataset(100, (32, 32), pool=8)
    x, y = ds[99]
    assert x.shape == (3, 32, 32) and 0 <= int(y) < 1000 and len(ds) == 100
