"""One data-gradient shape for ncu: layer1 conv3 (256 -> 64 channels, 56 x 56, batch 256), plain or with the fused
BatchNorm-backward reduction.   python benchmarks/experiments/ncu_dgrad_case.py [--fused 1]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from pytorchdistributed_b200 import _lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--fused", type=int, default=0)
args = ap.parse_args()
CL = torch.channels_last
dev = "cuda"
n, cin, cout, h = 256, 64, 256, 56
w = (torch.randn(cout, cin, 1, 1, device=dev) * 0.05).to(torch.bfloat16).contiguous(memory_format=CL)
dy = torch.randn(n, cout, h, h, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)
y = torch.randn(n, cin, h, h, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)
m = n * h * h
mask = torch.randint(0, 256, (m, cin // 8), dtype=torch.uint8, device=dev)
mean, invstd = torch.zeros(cin, device=dev), torch.ones(cin, device=dev)
dx = torch.empty_like(y)
sums = torch.zeros(2, cin, device=dev)
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    if args.fused:
        _lib.call("ptd_conv_dgrad_fused", dy, w, dx, None, None, y, mask, mean, invstd, sums[0], sums[1], n, h, h, cin, cout, 1, 1, 1, 0, st)
    else:
        _lib.call("ptd_conv_dgrad", dy, w, dx, None, n, h, h, cin, cout, 1, 1, 1, 0, st)
torch.cuda.synchronize()
print("ok", float(dx.float().abs().mean()))
