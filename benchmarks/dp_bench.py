"""Single-process DataParallel benchmark (BASELINE config 3): ResNet-50 replicate / scatter / parallel-apply / gather
across all visible GPUs, forward + backward + SGD, synthetic 3x224x224 batches that live on GPU 0.

    python benchmarks/dp_bench.py [--impl ours|reference] [--per-gpu-batch 256] [--steps 10] [--graph 1]

``ours``       ``pytorchdistributed_b200.parallel.DataParallel`` with the loss fused into the gather; ``--graph 1``
               (default) replays one captured (forward, backward) graph pair per replica from one thread.
``reference``  ``torch.nn.DataParallel`` around torchvision's ResNet-50 under bf16 autocast - the notebook's call
               (R/01 数据并行/01_multi_gpus_data_parallelism.ipynb cell 5), same precision as our arm.

Timing: CUDA events on GPU 0's stream (the root gathers the loss and reduces the gradients, so it finishes last)
around ``steps`` steps after ``warmup`` ones, every device synchronised on both sides; the two alternating input
batches are 154 MB x n_gpus each (> L2).  SM clocks / throttle reasons are sampled during the timed region.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--per-gpu-batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--img", type=int, default=224)
    ap.add_argument("--graph", type=int, default=1)
    ap.add_argument("--gpus", type=int, default=0, help="use the first N visible GPUs (default: all)")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    from pytorchdistributed_b200.utils import ClockSampler
    n = args.gpus or torch.cuda.device_count()
    B = args.per_gpu_batch * n
    xs = [torch.randn(B, 3, args.img, args.img, device="cuda:0") for _ in range(2)]
    ys = [torch.randint(0, 1000, (B,), device="cuda:0") for _ in range(2)]
    if args.impl == "ours":
        from pytorchdistributed_b200.models import resnet50
        from pytorchdistributed_b200.ops import functional as K
        from pytorchdistributed_b200.optim import SGD
        from pytorchdistributed_b200.parallel import DataParallel
        model = resnet50().to("cuda:0").train()
        dp = DataParallel(model, device_ids=list(range(n)), criterion=K.cross_entropy,      # gather fused into the loss
                          cuda_graph=bool(args.graph))
        opt = SGD(model.parameters(), lr=1e-3, model=model)

        def step(i):
            opt.zero_grad()
            loss = dp(xs[i % 2], ys[i % 2])
            loss.backward()
            opt.step()
            return loss
    else:
        import torchvision
        model = torchvision.models.resnet50().to("cuda:0").train()
        dp = torch.nn.DataParallel(model, device_ids=list(range(n)))      # the reference's call (notebook 01 cell 5)
        opt = torch.optim.SGD(model.parameters(), lr=1e-3)

        def step(i):
            opt.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = dp(xs[i % 2])
            loss = torch.nn.functional.cross_entropy(out.float(), ys[i % 2])
            loss.backward()
            opt.step()
            return loss

    for i in range(max(args.warmup, 3)):
        step(i)
    for d in range(n):
        torch.cuda.synchronize(d)
    sampler = ClockSampler(0).start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.device(0):
        e0.record()
        for i in range(args.steps):
            loss = step(i)
        e1.record()
    for d in range(n):
        torch.cuda.synchronize(d)
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    res = {"impl": args.impl, "n_gpus": n, "global_batch": B, "per_gpu_batch": args.per_gpu_batch, "dtype": "bf16",
           "images_per_sec": round(B * args.steps / (ms / 1e3), 1), "ms_per_step": round(ms / args.steps, 2),
           "steps": args.steps, "warmup": max(args.warmup, 3), "loss": float(loss), "clocks": clocks,
           "timing": "CUDA events on GPU 0, all devices synchronised on both sides; inputs alternate between two batches > L2"}
    if args.impl == "ours":
        res["cuda_graph"] = bool(args.graph)
        res["replicate"] = dict(dp.stats)
    print(json.dumps(res))
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
