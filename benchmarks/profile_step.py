"""One training step of the flagship config inside a cudaProfilerStart/Stop window.

    python benchmarks/profile_step.py [--batch 256] [--img 224] [--impl ours|reference]
    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python benchmarks/profile_step.py

Without ncu it prints the host-side time to issue one step (CPU-bound check) and the device time.
"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--img", type=int, default=224)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    x = torch.randn(args.batch, 3, args.img, args.img, device=dev)
    y = torch.randint(0, 1000, (args.batch,), device=dev)
    if args.impl == "ours":
        from pytorchdistributed_b200.models import resnet50
        from pytorchdistributed_b200.ops import functional as K
        from pytorchdistributed_b200.optim import SGD
        model = resnet50().to(dev).train()
        opt = SGD(model.parameters(), lr=1e-3, model=model)
        loss_fn = K.cross_entropy
        fwd = model
    else:
        import torchvision
        model = torchvision.models.resnet50().to(dev).to(memory_format=torch.channels_last).train()
        opt = torch.optim.SGD(model.parameters(), lr=1e-3)
        loss_fn = torch.nn.functional.cross_entropy
        torch.backends.cudnn.benchmark = True

        def fwd(inp):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return model(inp.contiguous(memory_format=torch.channels_last)).float()

    def step():
        opt.zero_grad()
        loss = loss_fn(fwd(x), y)
        loss.backward()
        opt.step()
        return loss

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()
    t0 = time.perf_counter()
    e0.record()
    step()
    e1.record()
    t_issue = time.perf_counter() - t0
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print(f"impl={args.impl} host issue time {t_issue * 1e3:.2f} ms, device time {e0.elapsed_time(e1):.2f} ms")


if __name__ == "__main__":
    main()
