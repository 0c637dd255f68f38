"""Single-process DataParallel benchmark (BASELINE config 3): ResNet-50 replicate/scatter/parallel-apply/gather
across all visible GPUs, forward+backward+SGD, per-GPU batch 64 by default (single-process DP is host-bound).

    python benchmarks/dp_bench.py [--impl ours|reference] [--per-gpu-batch 64] [--steps 10]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--per-gpu-batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--img", type=int, default=224)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    n = torch.cuda.device_count()
    B = args.per_gpu_batch * n
    x = torch.randn(B, 3, args.img, args.img, device="cuda:0")
    y = torch.randint(0, 1000, (B,), device="cuda:0")
    if args.impl == "ours":
        from pytorchdistributed_b200.models import resnet50
        from pytorchdistributed_b200.ops import functional as K
        from pytorchdistributed_b200.optim import SGD
        from pytorchdistributed_b200.parallel import DataParallel
        model = resnet50().to("cuda:0").train()
        dp = DataParallel(model, criterion=K.cross_entropy)       # gather fused into the loss
        opt = SGD(model.parameters(), lr=1e-3)

        def step():
            opt.zero_grad()
            loss = dp(x, y)
            loss.backward()
            opt.step()
            return loss
    else:
        import torchvision
        model = torchvision.models.resnet50().to("cuda:0").train()
        dp = torch.nn.DataParallel(model)                          # the reference's call (notebook 01 cell 5)
        opt = torch.optim.SGD(model.parameters(), lr=1e-3)

        def step():
            opt.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = dp(x)
            loss = torch.nn.functional.cross_entropy(out.float(), y)
            loss.backward()
            opt.step()
            return loss

    for _ in range(args.warmup):
        step()
    for d in range(n):
        torch.cuda.synchronize(d)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        loss = step()
    for d in range(n):
        torch.cuda.synchronize(d)
    dt = time.perf_counter() - t0
    res = {"impl": args.impl, "n_gpus": n, "global_batch": B, "images_per_sec": round(B * args.steps / dt, 1),
           "ms_per_step": round(dt / args.steps * 1e3, 2), "loss": float(loss)}
    if args.impl == "ours":
        res["replicate"] = dict(dp.stats)
    print(json.dumps(res))
    if args.out:
        json.dump(res, open(args.out, "w"))


if __name__ == "__main__":
    main()
