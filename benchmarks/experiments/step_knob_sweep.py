"""In-step A/B of the library's environment knobs: the ResNet-50 training step of bench.py (1 GPU, batch 256, bf16,
whole-step CUDA graph) timed under each setting, one fresh process per setting on the same box, back to back.

Per-layer timings (benchmarks/op_bench.py) cannot see what a knob does to the overlap of the weight-gradient side
stream with the data-gradient chain; this can.  The first and the last run use the defaults: their difference is the
drift of the box over the sweep.

    python benchmarks/experiments/step_knob_sweep.py "PTD_WGRAD_WAVES=0.5" "PTD_WGRAD_WAVES=2" "PTDB200_PDL=1" ...
    python benchmarks/experiments/step_knob_sweep.py --child        (one measurement, prints one JSON line)
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def child(args):
    import torch

    from pytorchdistributed_b200.comm import ddp_setup, destroy_process_group
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.optim import SGD
    from pytorchdistributed_b200.train import Trainer
    from pytorchdistributed_b200.utils import ClockSampler

    for k, v in (("MASTER_ADDR", "127.0.0.1"), ("MASTER_PORT", "29543"), ("RANK", "0"), ("WORLD_SIZE", "1"),
                 ("LOCAL_RANK", "0")):
        os.environ.setdefault(k, v)
    ddp_setup()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    model = resnet50()
    opt = SGD(model.parameters(), lr=1e-3)
    trainer = Trainer(model, train_dataloader=None, optimizer=opt, quiet=True, cuda_graph=False)
    x = torch.randn(args.batch, 3, args.img, args.img, device=dev)
    y = torch.randint(0, 1000, (args.batch,), device=dev)
    for _ in range(3):
        trainer._step_eager(x, y)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        loss = trainer._step_eager(x, y)
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    times = []
    clocks = ClockSampler(0, period_ms=50)
    clocks.start()
    for r in range(args.rounds + 1):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.block):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        if r > 0:
            times.append(e0.elapsed_time(e1) / args.block)
    clocks.stop()
    times.sort()
    print("RESULT " + json.dumps({"median_ms": round(times[len(times) // 2], 4), "min_ms": round(times[0], 4),
                                  "max_ms": round(times[-1], 4), "loss": float(loss), "clocks": clocks.summary()}),
          flush=True)
    destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("settings", nargs="*", help='"NAME=VALUE[,NAME=VALUE...]" per run; defaults run first and last')
    ap.add_argument("--child", action="store_true")
    ap.add_argument("--rounds", type=int, default=5)
    ap.add_argument("--block", type=int, default=20)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--img", type=int, default=224)
    ap.add_argument("--out", default="gpurun_out/step_knob_sweep.json")
    args = ap.parse_args()
    if args.child:
        child(args)
        return
    rows = []
    for setting in ["defaults"] + list(args.settings) + ["defaults"]:
        env = dict(os.environ)
        env["PTDB200_QUIET"] = "1"
        if setting != "defaults":
            for kv in setting.split(","):
                k, v = kv.split("=", 1)
                env[k] = v
        cmd = [sys.executable, os.path.abspath(__file__), "--child", "--rounds", str(args.rounds), "--block",
               str(args.block), "--batch", str(args.batch), "--img", str(args.img)]
        try:
            res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=120)
            line = next((ln for ln in res.stdout.splitlines() if ln.startswith("RESULT ")), None)
            row = json.loads(line[7:]) if line else {"error": (res.stderr or res.stdout)[-400:]}
        except subprocess.TimeoutExpired:
            row = {"error": "timeout"}
        row["setting"] = setting
        rows.append(row)
        print(json.dumps(row), flush=True)
    base = [r["median_ms"] for r in rows if r["setting"] == "defaults" and "median_ms" in r]
    ref = sum(base) / len(base) if base else None
    for r in rows:
        if ref and "median_ms" in r:
            r["vs_defaults"] = round(r["median_ms"] / ref, 5)
    out = {"experiment": "ResNet-50 step (1 GPU, batch 256, bf16, whole-step graph) under each knob setting; "
                         "one process per row, same box, back to back", "block": args.block, "rounds": args.rounds,
           "defaults_ms": base, "rows": rows}
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w") as f:
        f.write(json.dumps(out, indent=1) + "\n")
    for r in rows:
        print(f"{r['setting']:40s} {r.get('median_ms', 'ERR')}  x{r.get('vs_defaults', '-')}")


if __name__ == "__main__":
    main()
