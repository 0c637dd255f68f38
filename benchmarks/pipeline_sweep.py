"""Model-parallel / pipeline benchmark (BASELINE config 4; reference notebook 03 cells 7-9, 12-13).

ResNet-50 split over 2 GPUs, batch 120 of 3x128x128, 1000 classes, one-hot float labels, Adam(lr=1e-3),
`train(model)` = 3 iterations - the reference's own procedure, timed two ways: host wall clock of train()
like the notebook (timeit, no explicit sync) AND device-synchronised.  Arms: single GPU, naive model parallel,
pipeline for every split_size of the reference sweep.  `--impl reference` runs the reference's own classes
(lifted verbatim from the notebook JSON under baseline/_ref) on torchvision + cuDNN in fp32, as published.

    python benchmarks/pipeline_sweep.py [--impl ours|reference] [--repeat 5] [--out file.json]
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


def generate_random_data():
    inputs = torch.randn(BATCH, 3, IMG[0], IMG[1])
    idx = torch.LongTensor(BATCH).random_(0, NUM_CLASSES).view(BATCH, 1)
    one_hot = torch.zeros(BATCH, NUM_CLASSES).scatter_(1, idx, 1)
    return inputs, one_hot


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
    ap.add_argument("--repeat", type=int, default=5)
    ap.add_argument("--splits", default=",".join(map(str, SPLITS)))
    ap.add_argument("--out", default=None)
    ap.add_argument("--graph", action="store_true", help="ours: pipeline stages replayed from CUDA graphs")
    ap.add_argument("--prewarm", action="store_true",
                    help="build each model once and run one untimed train() before the timed repetitions (steady "
                         "state: allocator warm, graphs captured) instead of the notebook's fresh model per repetition")
    args = ap.parse_args()
    assert torch.cuda.device_count() >= 2, "needs 2 GPUs"
    splits = [int(s) for s in args.splits.split(",") if s]
    if args.impl == "reference":
        ns = reference_namespace()
        import torchvision.models as models
        train = ns["train"]
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
                    "model_parallel": lambda: ModelParallelResNet50(),
                    **{f"pipeline_{k}": (lambda k=k: PipelineParallelResNet50(split_size=k, cuda_graph=args.graph))
                       for k in splits}}

    results = {}
    for name, build in builders.items():
        wall, synced = [], []
        kept = None
        for rep in range(args.repeat + 1):
            if args.prewarm:
                if kept is None:
                    kept = build()
                    train(kept)              # untimed: allocator, kernel attributes, graph capture
                model = kept
            else:
                model = build()
            torch.cuda.synchronize("cuda:0"); torch.cuda.synchronize("cuda:1")
            t0 = time.perf_counter()
            train(model)
            t1 = time.perf_counter()
            torch.cuda.synchronize("cuda:0"); torch.cuda.synchronize("cuda:1")
            t2 = time.perf_counter()
            if rep > 0:                      # first repetition warms up allocator / cuDNN heuristics
                wall.append(t1 - t0)
                synced.append(t2 - t0)
            del model
        kept = None
        results[name] = {"wall_s_mean": statistics.mean(wall), "wall_s_std": statistics.pstdev(wall),
                         "synced_s_mean": statistics.mean(synced), "synced_s_std": statistics.pstdev(synced)}
        print(f"{args.impl:9s} {name:16s} train() = {results[name]['synced_s_mean']:.4f} s synced "
              f"(+-{results[name]['synced_s_std']:.4f}), {results[name]['wall_s_mean']:.4f} s host wall  "
              f"[{results[name]['synced_s_mean'] / NUM_EPOCH * 1e3:.1f} ms/iter]", flush=True)
    if args.out:
        json.dump({"impl": args.impl, "graph": args.graph, "prewarm": args.prewarm, "batch": BATCH, "img": IMG,
                   "iters_per_train": NUM_EPOCH, "results": results},
                  open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
