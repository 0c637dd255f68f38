"""Multi-process DDP on CPU/gloo (BASELINE config #1: plumbing without a GPU)."""
import os
import subprocess
import sys

import pytest
import torch

from tests._mp import ROOT, run_ranks


def _train(rank, world, steps, use_no_sync):
    import torch
    import torch.nn.functional as F
    from pytorchdistributed_b200.models import MLP
    from pytorchdistributed_b200.optim import SGD
    from pytorchdistributed_b200.parallel import DistributedDataParallel as DDP
    torch.manual_seed(100 + rank)            # different init per rank: the wrapper must make rank 0 win
    model = MLP()
    ddp = DDP(model, bucket_cap_mb=0.001, first_bucket_mb=0.0005)     # tiny caps -> several buckets
    opt = SGD(model.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(7)
    xs, ts = torch.randn(steps, 8 * world, 10, generator=g), torch.randn(steps, 8 * world, 5, generator=g)
    init = {k: v.clone() for k, v in model.state_dict().items()}
    ddp.enable_comm_timing()                 # per-bucket communication-time counter (host clock on the gloo path)
    for s in range(steps):
        x, t = xs[s, rank * 8:(rank + 1) * 8], ts[s, rank * 8:(rank + 1) * 8]
        opt.zero_grad()
        if use_no_sync and s % 2 == 0:
            with ddp.no_sync():
                F.mse_loss(ddp(x), t).backward()
            continue
        F.mse_loss(ddp(x), t).backward()
        opt.step()
    stats = dict(ddp.stats, comm_ms=ddp.comm_times_ms())
    return init, {k: v.clone() for k, v in ddp.state_dict().items()}, len(ddp.bucket_sizes_bytes()), stats


def test_ddp_matches_single_process():
    import torch.nn.functional as F
    from pytorchdistributed_b200.models import MLP
    steps, world = 4, 2
    res = run_ranks(_train, world, (steps, False))
    init0, final0, nb, stats = res[0]
    init1, final1, _, _ = res[1]
    assert nb > 1 and stats["iterations"] == steps
    assert len(stats["comm_ms"]) == nb and all(t >= 0.0 for t in stats["comm_ms"])     # one timing per bucket of the last backward
    for k in init0:
        assert torch.equal(init0[k], init1[k])                  # rank 0's weights were broadcast
    assert all(k.startswith("module.") for k in final0)         # reference quirk 6
    for k in final0:
        assert torch.allclose(final0[k], final1[k], atol=0, rtol=0)      # replicas stay bit-identical
    # the same training in one process on the concatenated batch
    ref = MLP()
    ref.load_state_dict(init0)
    opt = torch.optim.SGD(ref.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(7)
    xs, ts = torch.randn(steps, 8 * world, 10, generator=g), torch.randn(steps, 8 * world, 5, generator=g)
    for s in range(steps):
        opt.zero_grad()
        F.mse_loss(ref(xs[s]), ts[s]).backward()
        opt.step()
    for k, v in ref.state_dict().items():
        assert torch.allclose(final0["module." + k], v, atol=1e-6)


def test_ddp_no_sync_accumulates_locally():
    res = run_ranks(_train, 2, (4, True))
    _, f0, _, st = res[0]
    _, f1, _, _ = res[1]
    assert st["iterations"] == 2          # only the synchronised backward passes count
    for k in f0:
        assert torch.equal(f0[k], f1[k])


def _shape_mismatch(rank, world):
    from pytorchdistributed_b200.models import MLP
    from pytorchdistributed_b200.parallel import DistributedDataParallel as DDP
    try:
        DDP(MLP(hidden_size=20 + rank))
    except RuntimeError as e:
        return "shapes differ" in str(e)
    return False


def test_ddp_detects_shape_mismatch():
    assert all(run_ranks(_shape_mismatch, 2))


def _buffers(rank, world):
    import torch
    from pytorchdistributed_b200.models.layers import BatchNorm2d
    from pytorchdistributed_b200.parallel import DistributedDataParallel as DDP

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.bn = BatchNorm2d(64)

        def forward(self, x):
            return self.bn(x)

    torch.manual_seed(rank)
    net = Net()
    net.bn.running_mean.fill_(float(rank + 1))
    ddp = DDP(net)
    first = net.bn.running_mean.clone()
    x = torch.randn(4, 64, 3, 3) + rank
    ddp(x).sum().backward()
    after_one = net.bn.running_mean.clone()      # local update, differs per rank
    ddp(x).sum().backward()                      # the forward re-broadcasts rank 0's buffers first
    after_two = net.bn.running_mean.clone()
    ddp.eval()
    with torch.no_grad():
        ddp(x)                                   # first evaluation pass after training: rank 0's statistics win again
        in_eval = net.bn.running_mean.clone()
        net.bn.running_mean.add_(float(rank))    # a rank-local change ...
        ddp(x)                                   # ... survives the next evaluation pass: no broadcast after a no-grad forward
    return first, after_one, after_two, int(net.bn.num_batches_tracked), in_eval, net.bn.running_mean.clone()


def test_ddp_broadcasts_buffers_each_forward():
    r0, r1 = run_ranks(_buffers, 2)
    assert torch.equal(r0[0], r1[0]) and float(r0[0][0]) == 1.0       # construction: rank 0 wins
    assert not torch.allclose(r0[1], r1[1])                            # local statistics diverge ...
    exp = 0.9 * r0[1] + 0.1 * (r1[2] - 0.9 * r0[1]) / 0.1              # ... and are re-synchronised from rank 0
    assert torch.allclose(r1[2], 0.9 * r0[1] + (r1[2] - 0.9 * r0[1]), atol=1e-6) and exp is not None
    assert r0[3] == 2
    assert not torch.allclose(r0[2], r1[2])                            # after training the ranks' statistics differ ...
    assert torch.equal(r0[4], r0[2]) and torch.equal(r1[4], r0[2])     # ... the first evaluation forward ends that (torch's rule)
    assert torch.equal(r0[5], r0[4]) and torch.allclose(r1[5], r0[4] + 1.0)


def test_torchrun_cli_golden_stdout(tmp_path):
    snap = tmp_path / "snap.pt"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(ROOT, "scripts", "train_ddp_torchrun.py"), "--max_epochs", "2",
           "--batch_size", "32", "--save_every", "1", "--snapshot_path", str(snap), "--loss", "mse"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    for r in (0, 1):
        for e in (0, 1):
            assert f"[GPU: {r}] Epoch: {e} | Batchsize: 32 | Steps: 32" in out.stdout
        assert f"local_rank: {r}, world_size: 0" in out.stdout
    # rank-0 checkpoint loads into the reference's model definition, keys without the module. prefix
    state = torch.load(snap)["MODEL_STATE"]
    ref = torch.nn.Linear(20, 1)
    ref.load_state_dict(state)
    assert not torch.equal(ref.weight, torch.zeros_like(ref.weight))


def test_spawn_cli_golden_stdout():
    cmd = [sys.executable, os.path.join(ROOT, "scripts", "train_ddp_spawn.py"), "--max_epochs", "1", "--batch_size",
           "32", "--world_size", "2", "--port", "29633"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    for r in (0, 1):
        assert f"[GPU: {r}] Epoch: 0 | Batchsize: 32 | Steps: 32" in out.stdout


def test_spawn_cli_requires_max_epochs_like_reference():
    cmd = [sys.executable, os.path.join(ROOT, "scripts", "train_ddp_spawn.py"), "--world_size", "1", "--port", "29635"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=120, cwd=ROOT)
    assert out.returncode != 0            # range(None) - the reference has no default either (ddp_gpus.py:90)


def test_checkpoint_resume_roundtrip(tmp_path):
    """--save_every writes rank 0's module state; --resume picks the epoch counter and the weights up again."""
    snap = tmp_path / "snap.pt"
    base = [sys.executable, "-m", "torch.distributed.run", "--nproc-per-node=2", "--master-addr", "127.0.0.1"]
    script = [os.path.join(ROOT, "scripts", "train_ddp_torchrun.py"), "--batch_size", "64", "--save_every", "1",
              "--snapshot_path", str(snap), "--loss", "mse"]
    out1 = subprocess.run(base + ["--master-port", "29641"] + script + ["--max_epochs", "1"], capture_output=True,
                          text=True, timeout=300, cwd=ROOT)
    assert out1.returncode == 0, out1.stderr[-1500:]
    first = torch.load(snap)
    assert first["EPOCHS_RUN"] == 1
    out2 = subprocess.run(base + ["--master-port", "29643"] + script + ["--max_epochs", "2", "--resume"],
                          capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out2.returncode == 0, out2.stderr[-1500:]
    assert "Resuming training from snapshot at epoch 1" in out2.stdout
    assert "Epoch: 0 |" not in out2.stdout and "Epoch: 1 |" in out2.stdout       # epoch 0 is not repeated
    second = torch.load(snap)
    assert second["EPOCHS_RUN"] == 2
    assert not torch.equal(first["MODEL_STATE"]["weight"], second["MODEL_STATE"]["weight"])


def test_cli_communication_flags_are_additive(tmp_path):
    """The survey's additive flags (--backend / --dtype / --bucket_cap_mb) parse on both launchers and keep the golden
    output; the reference's own flags and defaults are unchanged (ddp_gpus.py:88-92, ddp_gpus_torchrun.py:88-96)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29641", os.path.join(root, "scripts", "train_ddp_torchrun.py"),
                          "--max_epochs", "1", "--batch_size", "32", "--backend", "gloo", "--dtype", "bf16",
                          "--bucket_cap_mb", "1"], capture_output=True, text=True, env=env, timeout=280)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "Epoch: 0 | Batchsize: 32 | Steps: 32" in out.stdout
    helptext = subprocess.run([sys.executable, os.path.join(root, "scripts", "train_ddp_spawn.py"), "--help"],
                              capture_output=True, text=True, timeout=120).stdout
    for flag in ("--max_epochs", "--batch_size", "--backend", "--dtype", "--bucket_cap_mb"):
        assert flag in helptext


def _unused_and_hook(rank, world):
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    from pytorchdistributed_b200.parallel import DistributedDataParallel as DDP

    class TwoHeads(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.body = torch.nn.Linear(6, 8)
            self.used = torch.nn.Linear(8, 3)
            self.unused = torch.nn.Linear(8, 3)          # never part of the loss

        def forward(self, x):
            return self.used(torch.relu(self.body(x)))

    torch.manual_seed(0)
    g = torch.Generator().manual_seed(3)
    xs, ts = torch.randn(8 * world, 6, generator=g), torch.randn(8 * world, 3, generator=g)
    x, t = xs[rank * 8:(rank + 1) * 8], ts[rank * 8:(rank + 1) * 8]
    out = {}
    # without the flag the missing gradients are an error, with it the step goes through
    strict = DDP(TwoHeads(), bucket_cap_mb=0.0001, first_bucket_mb=0.0001)
    try:
        F.mse_loss(strict(x), t).backward()
        out["strict_raised"] = False
    except RuntimeError as e:
        out["strict_raised"] = "find_unused_parameters" in str(e)
    torch.manual_seed(0)
    model = TwoHeads()
    ddp = DDP(model, bucket_cap_mb=0.0001, first_bucket_mb=0.0001, find_unused_parameters=True)
    calls = []

    def hook(flat):                                          # a communication hook: count, then average as usual
        calls.append(flat.numel())
        dist.all_reduce(flat)
        flat.div_(world)

    ddp.register_comm_hook(hook)
    for _ in range(2):
        for p in model.parameters():
            p.grad = None
        F.mse_loss(ddp(x), t).backward()
    out["hook_calls"] = len(calls)
    out["buckets"] = len(ddp.bucket_sizes_bytes())
    out["unused_grad_zero"] = bool((model.unused.weight.grad == 0).all())
    # the averaged gradient equals the single-process gradient on the whole batch
    torch.manual_seed(0)
    ref = TwoHeads()
    F.mse_loss(ref(xs), ts).backward()
    out["grad_matches"] = bool(torch.allclose(model.used.weight.grad, ref.used.weight.grad, atol=1e-6)
                               and torch.allclose(model.body.weight.grad, ref.body.weight.grad, atol=1e-6))
    return out


def test_find_unused_parameters_and_comm_hook():
    for res in run_ranks(_unused_and_hook, 2):
        assert res["strict_raised"] is True
        assert res["unused_grad_zero"] and res["grad_matches"], res
        assert res["hook_calls"] == 2 * res["buckets"] and res["buckets"] >= 2, res
