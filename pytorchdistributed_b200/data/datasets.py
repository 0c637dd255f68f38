"""Synthetic datasets with the shapes/dtypes of the reference's toy data.

* ``MyTrainDataset``  - ``R/02 DDP基本概念/ddp_gpus.py:57-66``: ``size`` pairs of
  ``(rand(20), rand(1))`` float32 (the "label" is a float of shape [1]).
* ``SimpleDataset``   - ``R/01 数据并行/01_multi_gpus_data_parallelism.ipynb#cell2``:
  ``randn(size, 10)`` inputs, ``randint(0, 1, (size, 5))`` int64 labels (all zero).
* ``generate_random_data`` - ``R/03 模型并行/03_model_parallel.ipynb#cell7``: a CPU batch
  ``randn(B,3,H,W)`` and float one-hot labels ``[B, num_classes]``.
* ``SyntheticImageDataset`` is an addition for the ResNet-50 runs of the CLI scripts (no dataset can be
  downloaded on the bench box).
* ``DeviceSyntheticBatches`` generates whole batches on the training device (no host memory, no copies): for kernel-side
  benchmarks where the input pipeline must not be part of the measurement.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch.utils.data import Dataset


class MyTrainDataset(Dataset):
    def __init__(self, size: int, in_features: int = 20, generator: Optional[torch.Generator] = None) -> None:
        self.size = size
        xs = torch.rand(size, in_features, generator=generator)
        ys = torch.rand(size, 1, generator=generator)
        self.data = [(xs[i], ys[i]) for i in range(size)]

    def __len__(self) -> int:
        return self.size

    def __getitem__(self, index: int):
        return self.data[index]


class SimpleDataset(Dataset):
    def __init__(self, size: int, input_size: int = 10, output_size: int = 5) -> None:
        self.size = size
        self.data = torch.randn(size, input_size)
        self.labels = torch.randint(0, 1, (size, output_size))

    def __len__(self) -> int:
        return self.size

    def __getitem__(self, index: int):
        return self.data[index], self.labels[index]


def generate_random_data(batch_size: int = 120, num_classes: int = 1000,
                         img_size: Tuple[int, int] = (128, 128),
                         generator: Optional[torch.Generator] = None):
    inputs = torch.randn(batch_size, 3, img_size[0], img_size[1], generator=generator)
    idx = torch.randint(0, num_classes, (batch_size, 1), generator=generator)
    one_hot = torch.zeros(batch_size, num_classes).scatter_(1, idx, 1.0)
    return inputs, one_hot


class SyntheticImageDataset(Dataset):
    """Fixed pool of random images + integer labels (host memory, optionally pinned)."""

    def __init__(self, size: int, img_size: Tuple[int, int] = (224, 224), num_classes: int = 1000,
                 seed: int = 0, pool: int = 512) -> None:
        g = torch.Generator().manual_seed(seed)
        self.size = size
        self.pool = min(pool, size)
        self.images = torch.randn(self.pool, 3, img_size[0], img_size[1], generator=g)
        self.labels = torch.randint(0, num_classes, (self.pool,), generator=g)

    def __len__(self) -> int:
        return self.size

    def __getitem__(self, index: int):
        j = index % self.pool
        return self.images[j], self.labels[j]


class DeviceSyntheticBatches:
    """``steps`` batches ``(images [B,3,H,W] fp32, labels [B] int64)`` living on ``device``.

    A small pool of batches is generated once with the device's own generator (seeded, so every run and every rank
    with the same seed sees the same data) and then cycled: iterating costs nothing on the host or on the device.
    ``len()`` is the number of steps, so it can stand in for a ``DataLoader`` in ``Trainer`` (``batch_size`` is there
    for the epoch banner).  Works on CPU too (tests)."""

    def __init__(self, steps: int, batch_size: int, img_size: Tuple[int, int] = (224, 224), num_classes: int = 1000,
                 device: Optional[torch.device] = None, seed: int = 0, pool: int = 2, one_hot: bool = False) -> None:
        self.steps, self.batch_size = int(steps), int(batch_size)
        self.device = torch.device(device) if device is not None else (
            torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu"))
        g = torch.Generator(device=self.device).manual_seed(seed)
        self._pool = []
        for _ in range(max(1, min(pool, steps))):
            x = torch.randn(batch_size, 3, img_size[0], img_size[1], device=self.device, generator=g)
            y = torch.randint(0, num_classes, (batch_size,), device=self.device, generator=g)
            if one_hot:
                y = torch.zeros(batch_size, num_classes, device=self.device).scatter_(1, y[:, None], 1.0)
            self._pool.append((x, y))

    def __len__(self) -> int:
        return self.steps

    def __iter__(self):
        for i in range(self.steps):
            yield self._pool[i % len(self._pool)]
