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


def test_device_synthetic_batches_cycle_a_seeded_pool():
    from pytorchdistributed_b200.data import DeviceSyntheticBatches
    a = DeviceSyntheticBatches(5, 4, img_size=(8, 8), num_classes=10, device="cpu", seed=7, pool=2)
    b = DeviceSyntheticBatches(5, 4, img_size=(8, 8), num_classes=10, device="cpu", seed=7, pool=2)
    assert len(a) == 5
    got = list(a)
    assert len(got) == 5
    x, y = got[0]
    assert x.shape == (4, 3, 8, 8) and x.dtype == torch.float32 and y.shape == (4,) and y.dtype == torch.int64
    assert got[2][0] is got[0][0] and got[1][0] is not got[0][0]          # pool of two, cycled
    assert torch.equal(list(b)[1][0], got[1][0])                          # same seed, same data
    oh = DeviceSyntheticBatches(1, 4, img_size=(8, 8), num_classes=10, device="cpu", one_hot=True)
    yy = next(iter(oh))[1]
    assert yy.shape == (4, 10) and torch.equal(yy.sum(1), torch.ones(4))


def test_data_parallel_demo_script_prints_the_notebook_lines():
    """R/01 数据并行/01_multi_gpus_data_parallelism.ipynb#cell3, #cell6 outputs (json L81-84, L139-145)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    out = subprocess.run([sys.executable, os.path.join(root, "scripts", "demo_data_parallel.py")], env=env,
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-500:]
    lines = out.stdout.strip().splitlines()
    assert lines[0] == "Data shape: torch.Size([32, 10]), Labels shape: torch.Size([32, 5])"
    assert lines[1] == "Outside: input size torch.Size([32, 10]) output_size torch.Size([32, 5])"
    assert len(lines) == 2


def test_svg_charts_are_well_formed_and_scaled():
    """The reference's two matplotlib figures (03_model_parallel.ipynb json L401-431, L599-624) as plain SVG."""
    import xml.etree.ElementTree as ET
    from pytorchdistributed_b200.utils.svgplot import bar_chart, line_chart
    ns = "{http://www.w3.org/2000/svg}"
    svg = bar_chart([2.0, 1.0], [0.5, 0.0], ["Model <Parallel>", "Single GPU"], title="t & t", ylabel="s")
    root = ET.fromstring(svg)
    bars = [r for r in root.iter(ns + "rect") if r.get("fill-opacity")]
    assert len(bars) == 2
    assert abs(float(bars[0].get("height")) - 2.0 * float(bars[1].get("height"))) < 0.2      # heights scale with the means
    assert "Model &lt;Parallel&gt;" in svg and "t &amp; t" in svg
    svg = line_chart([("a", [1, 3, 5], [3.0, 2.0, 1.0], [0.1, 0.1, 0.1]), ("b", [1, 5], [1.5, 1.5], [0.0, 0.0])],
                     xlabel="Pipeline Split Size")
    root = ET.fromstring(svg)
    assert len(list(root.iter(ns + "polyline"))) == 2
    assert len(list(root.iter(ns + "circle"))) == 5
    ys = [float(c.get("cy")) for c in root.iter(ns + "circle")][:3]
    assert ys[0] < ys[1] < ys[2]                  # larger value = higher on the page = smaller y
