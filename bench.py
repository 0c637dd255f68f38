#!/usr/bin/env python
"""Headline benchmark: ResNet-50 DDP training throughput (images/sec), bf16 compute, synthetic data.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Config (BASELINE.json #2): ResNet-50, 3x224x224, per-GPU batch 256, SGD(lr=1e-3) as in the reference's DDP
script, one process per GPU.  Weak scaling: per-GPU work is fixed, ``value`` is the whole-job images/sec.

Two numbers per arm:
  * ``value``      - K steps on device-resident inputs (a different pre-staged batch each step; one batch is
                     154 MB fp32 > L2), timed with CUDA events between barrier+synchronize, max over ranks;
  * ``e2e.value``  - the same K steps through the public API (``Trainer.run_batches`` for ours, the reference's own
                     ``.to(gpu)`` + ``Trainer._run_batch`` loop for the reference arm): every step's batch is copied
                     from pinned host memory and every step's result is read back to the host, inside the timed region.

``--impl reference`` runs the UNMODIFIED reference: ``Trainer``/``ddp_setup`` imported from the verbatim copy of
``ddp_gpus_torchrun.py`` under ``baseline/_ref`` (the reference is a pair of scripts, not an installable package -
see DESIGN.md), driving torchvision's ResNet-50 through torch DDP + NCCL + cuDNN/cuBLAS under bf16 autocast with
channels_last - its stock code path, none of our kernels.
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import socket
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

REF_DIR = os.path.join(ROOT, "baseline", "_ref")
REF_SCRIPT = os.path.join(REF_DIR, "02 DDP基本概念", "ddp_gpus_torchrun.py")
BASELINE_PUBLISHED = None      # the reference publishes no DDP ResNet-50 images/sec


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _env_setup(args) -> tuple[int, int, int]:
    if "RANK" not in os.environ:
        if args.gpus != 1:
            raise SystemExit("--gpus > 1 must be launched with torch.distributed.run (one process per GPU)")
        os.environ.update(RANK="0", LOCAL_RANK="0", WORLD_SIZE="1", MASTER_ADDR="127.0.0.1",
                          MASTER_PORT=str(_free_port()))
    return int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])


def _make_batches(n_batches: int, batch: int, img: int, pinned: bool, seed: int):
    g = torch.Generator().manual_seed(seed)
    xs, ys = [], []
    for _ in range(n_batches):
        x = torch.randn(batch, 3, img, img, generator=g)
        y = torch.randint(0, 1000, (batch,), generator=g)
        if pinned:
            x, y = x.pin_memory(), y.pin_memory()
        xs.append(x)
        ys.append(y)
    return xs, ys


def _timed(fn, steps: int, device) -> float:
    """Milliseconds for `steps` calls of fn(i), CUDA events, barrier + synchronize on both sides, max over ranks."""
    dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize(device)
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    every = [torch.zeros_like(ms) for _ in range(dist.get_world_size())]
    dist.all_gather(every, ms)
    _timed.by_rank = [round(float(t.item()) / steps, 3) for t in every]      # ms per step on every rank (spread = the
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)                                 # slowest chip sets the job's rate)
    return float(ms.item())


def _timed_block(fn, device) -> float:
    """Milliseconds for one call of fn(), CUDA events, barrier + synchronize on both sides, max over ranks."""
    dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize(device)
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


# ------------------------------------------------------------------------------------------------------
def run_ours(args, rank, local_rank, world):
    from pytorchdistributed_b200 import _lib
    from pytorchdistributed_b200.comm import ddp_setup, destroy_process_group
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.optim import SGD
    from pytorchdistributed_b200.train import Trainer
    from pytorchdistributed_b200.utils import ClockSampler

    if not _lib.available():
        raise SystemExit(f"native library missing: {_lib.load_error()}")
    ddp_setup()
    device = torch.device("cuda", local_rank)
    torch.manual_seed(0)
    model = resnet50()
    opt = SGD(model.parameters(), lr=1e-3)
    trainer = Trainer(model, train_dataloader=None, optimizer=opt, quiet=True, cuda_graph=args.graph,
                      fuse_optimizer=os.environ.get("PTDB200_BENCH_FUSE_SGD", "0") == "1")
    n_stage = 2
    hx, hy = _make_batches(n_stage, args.batch, args.img, pinned=True, seed=1234 + rank)
    dx = [x.to(device) for x in hx]
    dy = [y.to(device) for y in hy]

    def step_dev(i):
        trainer._run_batch(dx[i % n_stage], dy[i % n_stage])

    def run_e2e(steps):
        """`steps` steps through the public loop (Trainer.run_batches): every step's batch is copied from pinned host
        memory (prefetched on a copy stream while the previous step computes) and every step's loss is read back."""
        losses = trainer.run_batches(((hx[i % n_stage], hy[i % n_stage]) for i in range(steps)))
        return losses

    for i in range(args.warmup):
        step_dev(i)
    if args.graph:                       # capture happens on the first step after the trainer's eager warm-up
        for i in range(4):
            if trainer._graph is not None or trainer._graph_failed:
                break
            step_dev(i)
        step_dev(0)
    sampler = ClockSampler(local_rank, period_ms=50).start() if rank == 0 else None
    l0 = _lib.LAUNCHES
    ms = _timed(step_dev, args.steps, device)
    launches = _lib.LAUNCHES - l0
    if trainer._graph is not None:       # replayed steps re-issue the recorded entry-point launches on the device
        launches += args.steps * trainer._graph["launches"]
    clocks = sampler.stop() if sampler else {}
    run_e2e(max(2, args.warmup // 2))
    ms_e2e = _timed_block(lambda: run_e2e(args.steps), device)
    final_loss = float(trainer.last_loss.item())
    stats = dict(trainer.model.stats)
    comm = getattr(trainer.model.backend, "name", "?")
    if hasattr(trainer.model.backend, "mode"):          # nvls = reduced inside the NVSwitch, vmm / ipc = two-shot peer-to-peer
        comm += ":" + trainer.model.backend.mode + ":" + str(dict(trainer.model.backend.stats))
    by_rank = list(getattr(_timed, "by_rank", []) or [])
    destroy_process_group()
    return ms, ms_e2e, launches, clocks, {"final_loss": final_loss, "ddp": stats, "grad_comm": comm,
                                          "cuda_graph": bool(trainer._graph is not None),
                                          "fused_sgd_in_allreduce": bool(trainer.fused_optimizer),
                                          "native_reducer": bool(getattr(trainer.model, "_native", False)),
                                          "ms_per_step_by_rank": by_rank}


def run_reference(args, rank, local_rank, world):
    global REF_SCRIPT
    if not os.path.exists(REF_SCRIPT):
        # baseline/_ref is git-ignored: on a fresh clone fall back to the read-only reference tree itself
        alt = os.path.join("/root/reference", "02 DDP基本概念", "ddp_gpus_torchrun.py")
        if os.path.exists(alt):
            REF_SCRIPT = alt
    if not os.path.exists(REF_SCRIPT):
        return None, f"reference copy missing at {os.path.relpath(REF_SCRIPT, ROOT)}"
    try:
        import torchvision
    except Exception as e:  # noqa: BLE001
        return None, f"torchvision unavailable: {e}"
    spec = importlib.util.spec_from_file_location("ref_ddp_gpus_torchrun", REF_SCRIPT)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)          # defines ddp_setup / Trainer / MyTrainDataset; __main__ guard not taken
    from pytorchdistributed_b200.utils import ClockSampler

    ref.ddp_setup()                       # reference: init_process_group("nccl") + cuda.set_device(LOCAL_RANK)
    device = torch.device("cuda", local_rank)
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(0)

    class BF16(torch.nn.Module):
        """bf16 autocast + channels_last around torchvision's ResNet-50 (the reference passes any nn.Module)."""

        def __init__(self):
            super().__init__()
            self.net = torchvision.models.resnet50(num_classes=1000).to(memory_format=torch.channels_last)

        def forward(self, x):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return self.net(x.contiguous(memory_format=torch.channels_last)).float()

    model = BF16()
    opt = torch.optim.SGD(model.parameters(), lr=1e-3)
    trainer = ref.Trainer(model=model, train_dataloader=None, optimizer=opt)   # moves to GPU, wraps in torch DDP
    n_stage = 2
    hx, hy = _make_batches(n_stage, args.batch, args.img, pinned=True, seed=1234 + rank)
    dx = [x.to(device) for x in hx]
    dy = [y.to(device) for y in hy]
    probe = next(trainer.model.parameters())

    def step_dev(i):
        trainer._run_batch(dx[i % n_stage], dy[i % n_stage])

    host = torch.zeros(1).pin_memory()

    def step_e2e(i):
        x = hx[i % n_stage].to(local_rank, non_blocking=True)       # what the reference's _run_epoch does
        y = hy[i % n_stage].to(local_rank, non_blocking=True)
        trainer._run_batch(x, y)
        host.copy_(probe.detach().reshape(-1)[:1])                   # device -> host read after the step

    for i in range(args.warmup):
        step_dev(i)
    sampler = ClockSampler(local_rank, period_ms=50).start() if rank == 0 else None
    ms = _timed(step_dev, args.steps, device)
    clocks = sampler.stop() if sampler else {}
    for i in range(max(1, args.warmup // 2)):
        step_e2e(i)
    ms_e2e = _timed(step_e2e, args.steps, device)
    ref.destroy_process_group()
    return (ms, ms_e2e, clocks), None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--batch", type=int, default=256, help="per-GPU batch (BASELINE config: 256)")
    ap.add_argument("--img", type=int, default=224)
    ap.add_argument("--graph", type=int, default=int(os.environ.get("PTDB200_BENCH_GRAPH", "1")),
                    help="ours arm: Trainer(cuda_graph=True) - replay the whole step (fwd, bwd with the overlapped all-reduce, "
                         "optimizer) from one CUDA graph after an eager warm-up; falls back to eager if capture fails")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if not torch.cuda.is_available():
        if args.impl == "reference":
            print(json.dumps({"impl": "reference", "unavailable": "no CUDA device on this box"}))
            return
        raise SystemExit("bench.py needs a CUDA device")
    rank, local_rank, world = _env_setup(args)
    torch.cuda.set_device(local_rank)
    t0 = time.time()
    bytes_in = args.batch * 3 * args.img * args.img * 4 + args.batch * 8
    base = {
        "metric": "ResNet-50 DDP images/sec", "unit": "images/sec", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "dtype": "bf16", "data": "synthetic",
        "config": {"model": "resnet50", "global_batch": args.batch * world, "per_gpu_batch": args.batch,
                   "seq_len": None, "image": f"3x{args.img}x{args.img}", "optimizer": "SGD(lr=1e-3)",
                   "parallelism": f"dp{world}",
                   "l2": "inputs alternate between two 154 MB batches (> 126 MB L2); activations >> L2"},
    }
    if args.impl == "reference":
        res, why = run_reference(args, rank, local_rank, world)
        if res is None:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": why}))
            return
        ms, ms_e2e, clocks = res
        extra, launches = {}, None
    else:
        ms, ms_e2e, launches, clocks, extra = run_ours(args, rank, local_rank, world)
    if rank == 0:
        imgs = args.batch * world * args.steps
        value = imgs / (ms / 1e3)
        out = dict(base)
        out.update({
            "impl": args.impl, "value": round(value, 2), "ms_per_step": round(ms / args.steps, 3),
            "vs_baseline": (round(value / BASELINE_PUBLISHED, 4) if BASELINE_PUBLISHED else None),
            "clocks": {k: clocks.get(k) for k in ("sm_mhz", "sm_max_mhz", "reasons", "power_w_max", "samples")},
            "e2e": {"value": round(imgs / (ms_e2e / 1e3), 2), "unit": "images/sec",
                    "ms_per_step": round(ms_e2e / args.steps, 3),
                    "h2d_bytes_per_step": bytes_in, "d2h_bytes_per_step": 4},
            "wall_s": round(time.time() - t0, 1),
        })
        if launches is not None:
            out["gpu_launches"] = launches
        out.update(extra)
        print(json.dumps(out))


if __name__ == "__main__":
    main()
