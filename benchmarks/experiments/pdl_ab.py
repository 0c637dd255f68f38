"""A/B of programmatic dependent launch (PTDB200_PDL) on the flagship step, inside ONE process.

Two CUDA graphs of the same ResNet-50 training step (forward, loss, backward, SGD; batch 256 of 3x224x224, bf16) are
captured on the same tensors: one with the hot kernels launched normally, one with them launched as programmatic
dependents of their predecessor.  The graphs are then replayed in alternating blocks and timed with CUDA events, so box,
clocks and memory layout are identical for both arms.  Also checks that both graphs compute the same step.

    python benchmarks/experiments/pdl_ab.py [--rounds 6] [--block 20] [--out gpurun_out/pdl_ab.json]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from pytorchdistributed_b200 import _lib  # noqa: E402
from pytorchdistributed_b200.models import resnet50  # noqa: E402
from pytorchdistributed_b200.comm import ddp_setup, destroy_process_group  # noqa: E402
from pytorchdistributed_b200.optim import SGD  # noqa: E402
from pytorchdistributed_b200.train import Trainer  # noqa: E402
from pytorchdistributed_b200.utils import ClockSampler  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=6)
    ap.add_argument("--block", type=int, default=20)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--img", type=int, default=224)
    ap.add_argument("--mode", type=int, default=1, help="PTDB200_PDL value of the second arm (1 or 2)")
    ap.add_argument("--out", default="gpurun_out/pdl_ab.json")
    args = ap.parse_args()
    # the step bench.py times: Trainer -> DistributedDataParallel (gradients live in the flat buckets) -> fused SGD
    for k, v in (("MASTER_ADDR", "127.0.0.1"), ("MASTER_PORT", "29541"), ("RANK", "0"), ("WORLD_SIZE", "1"),
                 ("LOCAL_RANK", "0")):
        os.environ.setdefault(k, v)
    ddp_setup()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    model = resnet50()
    opt = SGD(model.parameters(), lr=1e-3)
    trainer = Trainer(model, train_dataloader=None, optimizer=opt, quiet=True, cuda_graph=False)
    model = trainer.model
    x = torch.randn(args.batch, 3, args.img, args.img, device=dev)
    y = torch.randint(0, 1000, (args.batch,), device=dev)

    def step():
        return trainer._step_eager(x, y)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    # everything a replay reads and writes: parameters, buffers, bf16 weight shadows, optimizer state (restored IN PLACE:
    # the graphs hold the addresses)
    live = list(model.parameters()) + list(model.buffers())
    for m in model.modules():
        sh = getattr(m, "_shadow", None)
        if torch.is_tensor(sh):
            live.append(sh)
        elif isinstance(sh, (tuple, list)):
            live += [t for t in sh if torch.is_tensor(t)]
    for st in opt.state.values():
        live += [v for v in st.values() if torch.is_tensor(v)]
    n_par = len(list(model.parameters()))
    state0 = [t.detach().clone() for t in live]

    graphs, losses = {}, {}
    for mode in (0, 1):
        _lib.call_ret("ptd_set_pdl_mode", args.mode if mode else 0)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            losses[mode] = step()
        torch.cuda.synchronize()
        graphs[mode] = g
    _lib.call_ret("ptd_set_pdl_mode", 0)

    def restore():
        with torch.no_grad():
            for t, s0 in zip(live, state0):
                t.copy_(s0)

    # -- same step?  three replays from the same state: normal twice (run-to-run noise of the atomics), then PDL ---------
    def run3(mode):
        restore()
        torch.cuda.synchronize()
        ls = []
        for _ in range(3):
            graphs[mode].replay()
            ls.append(float(losses[mode]))
        torch.cuda.synchronize()
        return ls, [p.detach().float().clone() for p in model.parameters()]

    def dist(a, b):
        return sum(float((u - v).pow(2).sum()) for u, v in zip(a, b)) ** 0.5

    la, pa = run3(0)
    lb, pb = run3(0)
    lp, pp = run3(1)
    upd = dist(pa, [s0.float() for s0 in state0[:n_par]])
    noise, diff = dist(pa, pb) / max(upd, 1e-30), dist(pa, pp) / max(upd, 1e-30)
    dl = max(abs(a - b) for a, b in zip(la, lp))
    print(f"losses normal {la}  normal again {lb}  pdl {lp}")
    print(f"update distance / update size: normal vs normal {noise:.3e}, normal vs pdl {diff:.3e}")

    # -- timing: alternating blocks ---------------------------------------------------------------------------------
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    times = {0: [], 1: []}
    clocks = ClockSampler(0, period_ms=50)
    clocks.start()
    for r in range(args.rounds + 1):
        for mode in ((0, 1) if r % 2 == 0 else (1, 0)):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(args.block):
                graphs[mode].replay()
            e1.record()
            torch.cuda.synchronize()
            if r > 0:                       # round 0 = warm-up of both graphs
                times[mode].append(e0.elapsed_time(e1) / args.block)
    clocks.stop()
    med = {m: sorted(v)[len(v) // 2] for m, v in times.items()}
    out = {"experiment": "programmatic dependent launch on the ResNet-50 step (1 GPU, batch 256, bf16, whole-step graph)",
           "pdl_mode": args.mode,
           "ms_per_step_normal": [round(t, 4) for t in times[0]], "ms_per_step_pdl": [round(t, 4) for t in times[1]],
           "median_normal_ms": round(med[0], 4), "median_pdl_ms": round(med[1], 4),
           "pdl_over_normal": round(med[1] / med[0], 5), "block": args.block, "rounds": args.rounds,
           "loss_max_abs_diff_normal_vs_pdl": dl, "update_distance_normal_vs_normal": noise,
           "update_distance_normal_vs_pdl": diff, "losses_normal": la, "losses_normal_again": lb, "losses_pdl": lp,
           "clocks": clocks.summary()}
    print(json.dumps(out))
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w") as f:
        f.write(json.dumps(out) + "\n")
    destroy_process_group()


if __name__ == "__main__":
    main()
