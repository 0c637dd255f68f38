"""Per layer shape (batch 256): data gradient alone, + separate bn_bwd_reduce, and the fused kernel.  CUDA events, L2
flushed between iterations."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from pytorchdistributed_b200 import _lib  # noqa: E402

CL = torch.channels_last
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
# (name, cin (= channels of dx / the BN), cout (channels of dy), k, stride, h_in)   dx: [n, cin, h, h]
SHAPES = [("l1 conv3 dgrad 256->64 @56", 64, 256, 1, 1, 56), ("l1 conv2 dgrad 3x3 64 @56", 64, 64, 3, 1, 56),
          ("l2 conv3 dgrad 512->128 @28", 128, 512, 1, 1, 28), ("l2 conv2 dgrad 3x3 128 @28", 128, 128, 3, 1, 28),
          ("l2.0 conv2 dgrad 3x3s2 128 @56", 128, 128, 3, 2, 56),
          ("l3 conv3 dgrad 1024->256 @14", 256, 1024, 1, 1, 14), ("l3 conv2 dgrad 3x3 256 @14", 256, 256, 3, 1, 14),
          ("l4 conv3 dgrad 2048->512 @7", 512, 2048, 1, 1, 7), ("l4 conv2 dgrad 3x3 512 @7", 512, 512, 3, 1, 7),
          ("l1.1 conv1 dgrad 64->256 @56 (+addend)", 256, 64, 1, 1, 56)]
n = 256


def timed(fn, iters=5):
    fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3


for name, cin, cout, k, stride, h in SHAPES:
    pad = k // 2
    oh = (h + 2 * pad - k) // stride + 1
    w = (torch.randn(cout, cin, k, k, device=dev) * 0.05).to(torch.bfloat16).contiguous(memory_format=CL)
    dy = torch.randn(n, cout, oh, oh, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)
    y = torch.randn(n, cin, h, h, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)
    m = n * h * h
    mask = torch.randint(0, 256, (m, cin // 8), dtype=torch.uint8, device=dev)
    mean, invstd = torch.zeros(cin, device=dev), torch.ones(cin, device=dev)
    dx = torch.empty_like(y)
    sums = torch.zeros(2, cin, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    gated = "addend" in name
    add = torch.randn_like(y) if gated else None
    t_d = timed(lambda: _lib.call("ptd_conv_dgrad_fused", dy, w, dx, add, mask if gated else None, None, None, None, None, None, None,
                                  n, h, h, cin, cout, k, k, stride, pad, st))
    t_r = timed(lambda: _lib.call("ptd_bn_bwd_reduce", dx, mask, y, mean, invstd, sums[0], sums[1], m, cin, 2, st))
    t_f = timed(lambda: _lib.call("ptd_conv_dgrad_fused", dy, w, dx, add, mask if gated else None, y, mask, mean, invstd, sums[0], sums[1],
                                  n, h, h, cin, cout, k, k, stride, pad, st))
    print(f"{name:42s} dgrad {t_d:7.1f} us + reduce {t_r:6.1f} us = {t_d + t_r:7.1f} us   fused {t_f:7.1f} us   "
          f"({'saves' if t_f < t_d + t_r else 'LOSES'} {abs(t_d + t_r - t_f):6.1f} us)", flush=True)
