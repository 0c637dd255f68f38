"""The smallest run that touches every kernel family once - what scripts/sanitize.sh puts under compute-sanitizer
(ONE tool per gpurun call).  Exits non-zero if a result is off, so the sanitizer log belongs to a correct program."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from pytorchdistributed_b200.models import resnet50  # noqa: E402
from pytorchdistributed_b200.ops import functional as K  # noqa: E402
from pytorchdistributed_b200.optim import SGD  # noqa: E402

CL = torch.channels_last
dev = "cuda"
torch.manual_seed(0)


def rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


bad = []
# convolution: 1x1 and 3x3 (im2col TMA), forward / data gradient / weight gradient
for cin, cout, r, stride, pad in [(64, 64, 1, 1, 0), (64, 128, 3, 1, 1), (128, 128, 3, 2, 1)]:
    x = torch.randn(2, cin, 16, 16, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)
    w = (torch.randn(cout, cin, r, r, device=dev) * 0.05).to(torch.bfloat16).contiguous(memory_format=CL)
    y, stats = K.conv_fprop_raw(x, w, stride, pad, True)
    ref = F.conv2d(x.float(), w.float(), None, stride, pad)
    dy = torch.randn_like(ref).to(torch.bfloat16).contiguous(memory_format=CL)
    dx = K.conv_dgrad_raw(dy, w, x.shape, stride, pad)
    dw = K.conv_wgrad_raw(x, dy, w.shape, stride, pad)
    rdx = torch.nn.grad.conv2d_input(x.shape, w.float(), dy.float(), stride, pad)
    rdw = torch.nn.grad.conv2d_weight(x.float(), w.shape, dy.float(), stride, pad)
    for name, a, b in (("fprop", y, ref), ("dgrad", dx, rdx), ("wgrad", dw, rdw)):
        if rel(a, b) > 2e-2:
            bad.append((cin, cout, r, stride, name, rel(a, b)))
# one tiny ResNet-50 step: stem, fused bottlenecks (BN, masks, dual BN), pools, classifier, loss, optimizer
model = resnet50(num_classes=16).to(dev).train()
opt = SGD(model.parameters(), lr=0.01, momentum=0.9, model=model)
x = torch.randn(2, 3, 32, 32, device=dev)
y = torch.randint(0, 16, (2,), device=dev)
for _ in range(2):
    opt.zero_grad()
    loss = K.cross_entropy(model(x), y)
    loss.backward()
    opt.step()
torch.cuda.synchronize()
if not torch.isfinite(loss):
    bad.append(("resnet loss", float(loss)))
print("loss", float(loss), "bad", bad)
sys.exit(1 if bad else 0)
