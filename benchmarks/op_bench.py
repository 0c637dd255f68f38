"""Per-layer microbenchmark of the ResNet-50 ops at the flagship batch (256 x 3x224x224).

Each distinct convolution shape is timed for fprop / dgrad / wgrad with CUDA events (warm-up, L2 flushed by a
256 MB write between iterations), plus the BN passes.  Prints achieved TFLOP/s and the fraction of the measured
bf16 peak / HBM bandwidth from MEASURED_PEAKS.json.  `--only NAME --once` runs a single op once (for ncu).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from pytorchdistributed_b200.ops import functional as K  # noqa: E402

CL = torch.channels_last
# (name, cin, cout, hin, k, stride, pad)
SHAPES = [
    ("l1.conv1_64-64_1x1", 64, 64, 56, 1, 1, 0), ("l1.conv2_64-64_3x3", 64, 64, 56, 3, 1, 1),
    ("l1.conv3_64-256_1x1", 64, 256, 56, 1, 1, 0), ("l1.conv1b_256-64_1x1", 256, 64, 56, 1, 1, 0),
    ("l2.conv1_256-128_1x1", 256, 128, 56, 1, 1, 0), ("l2.conv2s_128-128_3x3s2", 128, 128, 56, 3, 2, 1),
    ("l2.conv3_128-512_1x1", 128, 512, 28, 1, 1, 0), ("l2.ds_256-512_1x1s2", 256, 512, 56, 1, 2, 0),
    ("l2.conv1b_512-128_1x1", 512, 128, 28, 1, 1, 0), ("l2.conv2_128-128_3x3", 128, 128, 28, 3, 1, 1),
    ("l3.conv1_512-256_1x1", 512, 256, 28, 1, 1, 0), ("l3.conv2s_256-256_3x3s2", 256, 256, 28, 3, 2, 1),
    ("l3.conv3_256-1024_1x1", 256, 1024, 14, 1, 1, 0), ("l3.ds_512-1024_1x1s2", 512, 1024, 28, 1, 2, 0),
    ("l3.conv1b_1024-256_1x1", 1024, 256, 14, 1, 1, 0), ("l3.conv2_256-256_3x3", 256, 256, 14, 3, 1, 1),
    ("l4.conv1_1024-512_1x1", 1024, 512, 14, 1, 1, 0), ("l4.conv2s_512-512_3x3s2", 512, 512, 14, 3, 2, 1),
    ("l4.conv3_512-2048_1x1", 512, 2048, 7, 1, 1, 0), ("l4.ds_1024-2048_1x1s2", 1024, 2048, 14, 1, 2, 0),
    ("l4.conv1b_2048-512_1x1", 2048, 512, 7, 1, 1, 0), ("l4.conv2_512-512_3x3", 512, 512, 7, 3, 1, 1),
]


def x_like(t):
    """A distinct tensor with the same values (so that dz and y of a BN backward are not the same memory)."""
    return t.clone(memory_format=torch.preserve_format)


def timeit(fn, flush, iters):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--only", default=None)
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--out", default=None)
    ap.add_argument("--bn", action="store_true", help="with --once: also run the BN passes on the layer's output")
    args = ap.parse_args()
    dev = "cuda"
    peaks = {"bf16_tflops": 1643.2, "hbm_gbs": 6555.8}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks.update(json.load(open(pk)))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    from pytorchdistributed_b200.utils import ClockSampler
    sampler = ClockSampler(0, period_ms=100).start()       # SM clocks / throttle reasons over the whole table
    rows = []
    n = args.batch
    for name, cin, cout, h, k, s, p in SHAPES:
        if args.only and args.only not in name:
            continue
        oh = (h + 2 * p - k) // s + 1
        x = torch.randn(n, cin, h, h, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)
        w = (torch.randn(cout, cin, k, k, device=dev) * 0.05).to(torch.bfloat16).contiguous(memory_format=CL)
        dy = torch.randn(n, cout, oh, oh, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)
        flops = 2.0 * n * oh * oh * cout * cin * k * k
        act_bytes = (x.numel() + dy.numel()) * 2
        ops = {
            "fprop": lambda: K.conv_fprop_raw(x, w, s, p, True),
            "dgrad": lambda: K.conv_dgrad_raw(dy, w, x.shape, s, p),
            "wgrad": lambda: K.conv_wgrad_raw(x, dy, w.shape, s, p),
        }
        for opn, fn in ops.items():
            if args.once:
                fn()
                torch.cuda.synchronize()
                continue
            us = timeit(fn, flush, args.iters)
            tf = flops / us / 1e6
            gbs = act_bytes / us / 1e3
            rows.append({"layer": name, "op": opn, "us": round(us, 1), "tflops": round(tf, 1),
                         "frac_bf16_peak_of_measured": round(tf / peaks["bf16_tflops"], 3),
                         "min_gbs": round(gbs, 0), "frac_hbm_of_measured": round(gbs / peaks["hbm_gbs"], 3)})
            print(f"{name:28s} {opn:6s} {us:9.1f} us  {tf:7.1f} TFLOP/s ({tf / peaks['bf16_tflops']:.2f} of measured peak)"
                  f"  {gbs:6.0f} GB/s min-traffic ({gbs / peaks['hbm_gbs']:.2f} of measured HBM)", flush=True)
        # BN passes on this layer's output
        if args.once and args.bn:
            c = cout
            gamma, beta = torch.ones(c, device=dev), torch.zeros(c, device=dev)
            rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
            stats = K.bn_stats_raw(dy)
            z, saved = K.bn_apply_raw(dy, stats, gamma, beta, rm, rv, 1e-5, 0.1, True)
            K.bn_backward_raw(dy, z, x_like(dy), saved, gamma, True, False)
            z, saved, mask = K.bn_apply_raw(dy, stats, gamma, beta, rm, rv, 1e-5, 0.1, True, want_mask=True)
            K.bn_backward_raw(dy, None, x_like(dy), saved, gamma, True, False, mask=mask)
            torch.cuda.synchronize()
        if not args.once and not args.only:
            c = cout
            y = dy
            gamma, beta = torch.ones(c, device=dev), torch.zeros(c, device=dev)
            rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
            stats = K.bn_stats_raw(y)
            z, saved, mask = K.bn_apply_raw(y, stats, gamma, beta, rm, rv, 1e-5, 0.1, True, want_mask=True)
            y2 = x_like(y)
            # minimum traffic in units of one bf16 pass over the tensor: apply 2 (+1/16 mask); backward reads
            # dz, y, z twice and writes dy = 7, or 5 + 1/8 when the bit mask stands in for z
            for opn, fn, nb in (("bn_apply", lambda: K.bn_apply_raw(y, stats, gamma, beta, rm, rv, 1e-5, 0.1, True), 2),
                                ("bn_apply_m", lambda: K.bn_apply_raw(y, stats, gamma, beta, rm, rv, 1e-5, 0.1, True,
                                                                     want_mask=True), 2.0625),
                                ("bn_bwd", lambda: K.bn_backward_raw(y2, z, y, saved, gamma, True, False), 7),
                                ("bn_bwd_m", lambda: K.bn_backward_raw(y2, None, y, saved, gamma, True, False, mask=mask),
                                 5.125)):
                us = timeit(fn, flush, args.iters)
                gbs = nb * y.numel() * 2 / us / 1e3
                rows.append({"layer": name, "op": opn, "us": round(us, 1), "gbs": round(gbs, 0),
                             "frac_hbm_of_measured": round(gbs / peaks["hbm_gbs"], 3)})
                print(f"{name:28s} {opn:10s} {us:9.1f} us  {gbs:6.0f} GB/s ({gbs / peaks['hbm_gbs']:.2f} of measured HBM)", flush=True)
    clocks = sampler.stop()
    print(f"clocks over the run: {clocks}")
    if args.out:
        json.dump({"batch": n, "iters": args.iters, "clocks": clocks, "peaks": peaks, "rows": rows}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
