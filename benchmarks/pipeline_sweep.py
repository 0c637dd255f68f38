"""Model-parallel / pipeline benchmark (BASELINE config 4; reference notebook 03 cells 7-9, 12-13).

ResNet-50 split over 2 GPUs, batch 120 of 3x128x128, 1000 classes, one-hot float labels, Adam(lr=1e-3),
`train(model)` = 3 iterations.  Arms: single GPU, naive model parallel, pipeline for every split_size of the
reference sweep.  Two timings per arm, reported next to each other:

  * the notebook's procedure: host wall clock of train() (timeit style, no explicit sync) AND the same with both
    devices synchronised - CPU `randn` of the batch (~25 ms per iteration) is INSIDE the timed region, as published;
  * ``device_*``: the same train() on inputs generated ONCE outside the timed region, CUDA events on both devices'
    streams, synchronised on both sides - what the GPUs need, without the host's data generation.

`--impl reference` runs the reference's own classes (lifted verbatim from the notebook JSON under baseline/_ref) on
torchvision + cuDNN: in fp32 as published, or with `--autocast` under bf16 autocast (the precision of our arm).
SM clocks / throttle reasons are sampled while each arm runs.

    python benchmarks/pipeline_sweep.py [--impl ours|reference] [--autocast] [--graph] [--prewarm] [--repeat 10]
"""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch import nn  # noqa: E402

NUM_CLASSES, NUM_EPOCH, BATCH, IMG = 1000, 3, 120, (128, 128)
SPLITS = [1, 3, 5, 8, 10, 12, 20, 40, 60]
NB = os.path.join(ROOT, "baseline", "_ref", "03 模型并行", "03_model_parallel.ipynb")


_CACHED = []          # device-timed mode: batches generated once, outside the timed region
_USE_CACHE = False
_NEXT = 0


def _fresh_batch():
    inputs = torch.randn(BATCH, 3, IMG[0], IMG[1])
    idx = torch.LongTensor(BATCH).random_(0, NUM_CLASSES).view(BATCH, 1)
    one_hot = torch.zeros(BATCH, NUM_CLASSES).scatter_(1, idx, 1)
    return inputs, one_hot


def generate_random_data():
    global _NEXT
    if _USE_CACHE:
        _NEXT += 1
        return _CACHED[_NEXT % len(_CACHED)]
    return _fresh_batch()


def reference_namespace():
    """Execute the reference's own class/func definitions (cells 5, 7, 12 of notebook 03), unmodified."""
    nb = json.load(open(NB))
    ns = {"torch": torch, "nn": nn}
    exec("from torchvision.models.resnet import ResNet, Bottleneck", ns)
    code5 = "".join(nb["cells"][5]["source"])
    code7 = "".join(nb["cells"][7]["source"])
    code12 = "".join(nb["cells"][12]["source"]).split("setup =")[0]      # class only, not the timing lines
    for code in (code5, code7, code12):
        exec(code, ns)
    return ns


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--repeat", type=int, default=10)
    ap.add_argument("--autocast", action="store_true", help="reference arm under bf16 autocast (precision-matched)")
    ap.add_argument("--splits", default=",".join(map(str, SPLITS)))
    ap.add_argument("--out", default=None)
    ap.add_argument("--graph", action="store_true", help="ours: pipeline stages replayed from CUDA graphs")
    ap.add_argument("--transfer", default="nccl", choices=["nccl", "p2p"],
                    help="ours: stage boundary by NCCL send/recv on side streams, or by our peer-copy kernel (with "
                         "--graph the copy is captured inside stage 2's graphs)")
    ap.add_argument("--prewarm", action="store_true",
                    help="build each model once and run one untimed train() before the timed repetitions (steady "
                         "state: allocator warm, graphs captured) instead of the notebook's fresh model per repetition")
    args = ap.parse_args()
    assert torch.cuda.device_count() >= 2, "needs 2 GPUs"
    splits = [int(s) for s in args.splits.split(",") if s]
    if args.impl == "reference":
        ns = reference_namespace()
        ns["generate_random_data"] = generate_random_data      # same function; lets the device-timed pass reuse batches
        import torchvision.models as models
        ref_train = ns["train"]

        def train(model):
            if args.autocast:
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    ref_train(model)
            else:
                ref_train(model)
        builders = {"single": lambda: models.resnet50(num_classes=NUM_CLASSES).to("cuda:0"),
                    "model_parallel": lambda: ns["ModelParallelResNet50"](),
                    **{f"pipeline_{k}": (lambda k=k: ns["PipelineParallelResNet50"](split_size=k)) for k in splits}}
    else:
        from pytorchdistributed_b200.models import resnet50
        from pytorchdistributed_b200.ops import functional as K
        from pytorchdistributed_b200.optim import Adam
        from pytorchdistributed_b200.parallel import ModelParallelResNet50, PipelineParallelResNet50

        def train(model):
            model.train()
            optimizer = Adam(model.parameters(), lr=1e-3, model=model)
            for _ in range(NUM_EPOCH):
                inputs, labels = generate_random_data()
                optimizer.zero_grad()
                outputs = model(inputs.to("cuda:0"))
                loss = K.cross_entropy(outputs, labels.to(outputs.device))
                loss.backward()
                optimizer.step()

        builders = {"single": lambda: resnet50(num_classes=NUM_CLASSES).to("cuda:0"),
                    "model_parallel": lambda: ModelParallelResNet50(transfer=args.transfer),
                    **{f"pipeline_{k}": (lambda k=k: PipelineParallelResNet50(split_size=k, cuda_graph=args.graph,
                                                                              transfer=args.transfer))
                       for k in splits}}

    from pytorchdistributed_b200.utils import ClockSampler
    global _USE_CACHE
    _CACHED.extend(tuple(t.pin_memory() for t in _fresh_batch()) for _ in range(NUM_EPOCH))
    results = {}
    sync = lambda: (torch.cuda.synchronize("cuda:0"), torch.cuda.synchronize("cuda:1"))  # noqa: E731
    for name, build in builders.items():
        wall, synced, dev_ms = [], [], []
        kept = None
        sampler = ClockSampler(0, period_ms=100).start()
        for rep in range(args.repeat + 1):
            if args.prewarm:
                if kept is None:
                    kept = build()
                    train(kept)              # untimed: allocator, kernel attributes, graph capture
                model = kept
            else:
                model = build()
            sync()
            t0 = time.perf_counter()
            train(model)
            t1 = time.perf_counter()
            sync()
            t2 = time.perf_counter()
            # device-timed pass: same train(), batches from the cache, CUDA events on both devices
            _USE_CACHE = True
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(2)]
            for d in (0, 1):
                with torch.cuda.device(d):
                    ev[d][0].record()
            train(model)
            for d in (0, 1):
                with torch.cuda.device(d):
                    ev[d][1].record()
            sync()
            _USE_CACHE = False
            if rep > 0:                      # first repetition warms up allocator / cuDNN heuristics
                wall.append(t1 - t0)
                synced.append(t2 - t0)
                dev_ms.append(max(e0.elapsed_time(e1) for e0, e1 in ev))
            del model
        kept = None
        clocks = sampler.stop()
        results[name] = {"wall_s_mean": statistics.mean(wall), "wall_s_std": statistics.pstdev(wall),
                         "synced_s_mean": statistics.mean(synced), "synced_s_std": statistics.pstdev(synced),
                         "device_s_mean": statistics.mean(dev_ms) / 1e3, "device_s_std": statistics.pstdev(dev_ms) / 1e3,
                         "clocks": clocks}
        print(f"{args.impl:9s} {name:16s} train() = {results[name]['synced_s_mean']:.4f} s synced "
              f"(+-{results[name]['synced_s_std']:.4f}), {results[name]['wall_s_mean']:.4f} s host wall, "
              f"{results[name]['device_s_mean']:.4f} s device-timed on cached inputs "
              f"[{results[name]['device_s_mean'] / NUM_EPOCH * 1e3:.1f} ms/iter]  clocks {clocks.get('sm_mhz')} {clocks.get('reasons')}",
              flush=True)
    if args.out:
        json.dump({"impl": args.impl, "graph": args.graph, "transfer": args.transfer, "prewarm": args.prewarm,
                   "autocast": args.autocast,
                   "precision": ("bf16 kernels" if args.impl == "ours" else ("bf16 autocast" if args.autocast else "fp32")),
                   "repeat": args.repeat, "batch": BATCH, "img": IMG, "iters_per_train": NUM_EPOCH, "results": results},
                  open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
