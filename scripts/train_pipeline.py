"""Model-parallel / pipeline training of ResNet-50 over two GPUs - notebook 03's ``train(model)`` as a command line
(``R/03 模型并行/03_model_parallel.ipynb`` cells 7 and 12: batch 120 of 3x128x128, 1000 classes, one-hot labels, Adam).

    python scripts/train_pipeline.py --split_size 20 [--mode pipeline|model_parallel|single] [--cuda_graph]
        [--transfer nccl|p2p] [--iters 3] [--batch_size 120] [--img_size 128]

``--split_size`` is the reference's micro-batch size (``PipelineParallelResNet50(split_size=...)``).
"""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from pytorchdistributed_b200.data import generate_random_data  # noqa: E402
from pytorchdistributed_b200.models import resnet50  # noqa: E402
from pytorchdistributed_b200.ops import functional as K  # noqa: E402
from pytorchdistributed_b200.optim import Adam  # noqa: E402
from pytorchdistributed_b200.parallel import ModelParallelResNet50, PipelineParallelResNet50  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser(description="split-model pipeline training")
    ap.add_argument("--mode", choices=["pipeline", "model_parallel", "single"], default="pipeline")
    ap.add_argument("--split_size", type=int, default=20)
    ap.add_argument("--cuda_graph", action="store_true")
    ap.add_argument("--transfer", choices=["nccl", "p2p"], default="nccl")
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--batch_size", type=int, default=120)
    ap.add_argument("--img_size", type=int, default=128)
    ap.add_argument("--num_classes", type=int, default=1000)
    args = ap.parse_args()
    gpu = torch.cuda.is_available()
    if args.mode == "single" or not gpu or torch.cuda.device_count() < 2:
        model = resnet50(num_classes=args.num_classes)
        model = model.to("cuda:0") if gpu else model
    elif args.mode == "model_parallel":
        model = ModelParallelResNet50(num_classes=args.num_classes, transfer=args.transfer)
    else:
        model = PipelineParallelResNet50(split_size=args.split_size, num_classes=args.num_classes,
                                         cuda_graph=args.cuda_graph, transfer=args.transfer)
    model.train()
    optimizer = Adam(model.parameters(), lr=1e-3, model=model)
    t0 = time.perf_counter()
    for it in range(args.iters):
        inputs, labels = generate_random_data(args.batch_size, args.num_classes, (args.img_size, args.img_size))
        optimizer.zero_grad()
        outputs = model(inputs.to("cuda:0") if gpu else inputs)
        loss = K.cross_entropy(outputs, labels.to(outputs.device))
        loss.backward()
        optimizer.step()
        print(f"iter {it} | loss {float(loss):.4f} | split_size {args.split_size} | mode {args.mode}", flush=True)
    if gpu:
        for d in range(torch.cuda.device_count()):
            torch.cuda.synchronize(d)
    print(f"{args.iters} iterations in {time.perf_counter() - t0:.3f} s")


if __name__ == "__main__":
    main()
