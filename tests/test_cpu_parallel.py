"""Wrappers on CPU: split-model bookkeeping, micro-batch splitting, DataParallel passthrough (reference C13, C16, C19)."""
import pytest
import torch

from pytorchdistributed_b200.models import MLP, resnet50
from pytorchdistributed_b200.parallel import DataParallel, ModelParallelResNet50, PipelineParallelResNet50


@pytest.mark.parametrize("split", [1, 3, 4, 7, 20])
def test_pipeline_output_independent_of_split_size(split):
    torch.manual_seed(0)
    pp = PipelineParallelResNet50(split_size=split, num_classes=10, devices=("cpu", "cpu")).eval()
    plain = resnet50(num_classes=10).eval()
    plain.load_state_dict(pp.state_dict())
    x = torch.randn(7, 3, 64, 64)                 # 7 is not a multiple of most split sizes: uneven last micro-batch
    with torch.no_grad():
        assert torch.allclose(pp(x), plain(x), atol=1e-4)


def test_model_parallel_state_dict_layouts():
    mp = ModelParallelResNet50(num_classes=1000, devices=("cpu", "cpu"))
    sd = mp.state_dict()
    assert len(sd) == 320 and "seq1.0.weight" not in sd          # torchvision's keys only
    ref = mp.reference_state_dict()
    assert len(ref) == 638                                        # the reference class registers everything twice
    assert ref["seq1.0.weight"] is sd["conv1.weight"] or torch.equal(ref["seq1.0.weight"], sd["conv1.weight"])
    assert torch.equal(ref["seq2.1.2.bn3.weight"], sd["layer4.2.bn3.weight"])
    assert sum(p.numel() for p in mp.parameters()) == 25_557_032


def test_pipeline_trains_on_cpu():
    torch.manual_seed(0)
    pp = PipelineParallelResNet50(split_size=2, num_classes=10, devices=("cpu", "cpu")).train()
    x, y = torch.randn(4, 3, 32, 32), torch.randint(0, 10, (4,))
    loss = torch.nn.functional.cross_entropy(pp(x), y)
    loss.backward()
    assert all(p.grad is not None for p in pp.parameters())


def test_data_parallel_without_gpu_is_passthrough():
    m = MLP()
    dp = DataParallel(m)
    x = torch.randn(32, 10)
    assert dp(x).shape == (32, 5)
    assert torch.equal(dp(x), m(x))
    assert list(dp.state_dict().keys())[0].startswith("module.")


def _fd_exchange(rank, world):
    """Every rank opens a pipe, writes its rank into it and hands the READ end to everybody (unix-socket fd passing,
    the way the VMM handles of comm/p2p.py travel); reading from a received descriptor yields the sender's rank."""
    import os
    from pytorchdistributed_b200.comm.p2p import FdChannel
    chan = FdChannel(None)
    out = {}
    for round_ in range(2):                                  # two exchanges back to back: tags keep them apart
        r, w = os.pipe()
        os.write(w, bytes([10 * round_ + rank]) * (world - 1))
        got = chan.all_to_all(r)
        out[("a2a", round_)] = [None if fd is None else os.read(fd, 1)[0] for fd in got]
        for fd in got:
            if fd is not None:
                os.close(fd)
        os.close(r)
        os.close(w)
    r, w = os.pipe()
    os.write(w, bytes([77]) * world)
    fd = chan.from_root(r if rank == 0 else None, 0)
    out["root"] = None if fd is None else os.read(fd, 1)[0]
    chan.close()
    return out


def test_fd_channel_passes_descriptors_between_ranks():
    from tests._mp import run_ranks
    world = 3
    res = run_ranks(_fd_exchange, world, backend="gloo")
    for rank, out in enumerate(res):
        for round_ in range(2):
            assert out[("a2a", round_)] == [None if r == rank else 10 * round_ + r for r in range(world)]
        assert out["root"] == (None if rank == 0 else 77)


def test_numa_binding_reads_sysfs_and_is_conservative(tmp_path, monkeypatch):
    """utils/numa.py: CPUs of the GPU's NUMA node from a (fake) sysfs tree; no change when that is nothing new, too
    little, unknown, or switched off."""
    import os
    from pytorchdistributed_b200.utils import numa
    assert numa.parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert numa.parse_cpulist("") == set()
    assert numa.pci_address(0, 0x1b, 0) == "0000:1b:00.0"
    dev = tmp_path / "bus" / "pci" / "devices" / "0000:1b:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("1\n")
    node = tmp_path / "devices" / "system" / "node" / "node1"
    node.mkdir(parents=True)
    (node / "cpulist").write_text("56-111,168-223\n")
    cpus = numa.gpu_numa_cpus("0000:1b:00.0", str(tmp_path))
    assert cpus == set(range(56, 112)) | set(range(168, 224))
    assert numa.gpu_numa_cpus("0000:99:00.0", str(tmp_path)) is None            # unknown device
    (dev / "numa_node").write_text("-1\n")
    assert numa.gpu_numa_cpus("0000:1b:00.0", str(tmp_path)) is None            # platform without NUMA information
    allowed = set(range(224))
    assert numa.choose_affinity(cpus, allowed) == cpus
    assert numa.choose_affinity(cpus, cpus) is None                             # already there
    assert numa.choose_affinity(cpus, {0, 1, 2, 3, 56}) is None                 # one usable CPU: too little
    assert numa.choose_affinity(cpus, set(range(50, 60))) == {56, 57, 58, 59}   # a cpuset that overlaps the node
    assert numa.choose_affinity(None, allowed) is None
    before = os.sched_getaffinity(0)
    monkeypatch.setenv("PTDB200_NUMA_BIND", "0")
    assert numa.bind_to_gpu_node(0) is None
    monkeypatch.delenv("PTDB200_NUMA_BIND")
    assert numa.bind_to_gpu_node(0) is None                                     # no GPU here: nothing happens, nothing raised
    assert os.sched_getaffinity(0) == before
