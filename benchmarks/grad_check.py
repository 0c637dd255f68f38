"""End-to-end gradient check: our ResNet-50 (bf16 kernels) vs torchvision fp32 on the same weights/batch.
Prints the relative error of the loss, logits and every parameter gradient (worst first)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
import torchvision  # noqa: E402

from pytorchdistributed_b200.models import resnet50  # noqa: E402
from pytorchdistributed_b200.ops import functional as K  # noqa: E402


def rel(a, b):
    a, b = a.float(), b.float()
    return float((a - b).norm() / (b.norm() + 1e-12))


def main(batch=32, img=128):
    torch.manual_seed(0)
    dev = "cuda"
    m = resnet50().to(dev).train()
    with torch.no_grad():
        for p in m.parameters():
            p.copy_(p.to(torch.bfloat16).float())
    tv = torchvision.models.resnet50().to(dev).train()
    tv.load_state_dict(m.state_dict())
    x = torch.randn(batch, 3, img, img, device=dev).to(torch.bfloat16).float()
    y = torch.randint(0, 1000, (batch,), device=dev)
    lo, lr_ = m(x), tv(x)
    print("logits rel err", rel(lo, lr_))
    loss, lref = K.cross_entropy(lo, y), F.cross_entropy(lr_, y)
    print("loss", float(loss), float(lref))
    loss.backward(); lref.backward()
    errs = []
    for (n, p), (_, q) in zip(m.named_parameters(), tv.named_parameters()):
        errs.append((rel(p.grad, q.grad), n, float(q.grad.norm())))
    errs.sort(reverse=True)
    for e, n, nr in errs[:25]:
        print(f"{e:8.4f}  {n:40s} |ref grad| {nr:.3e}")
    print("median", sorted(e for e, _, _ in errs)[len(errs) // 2])


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 32, int(sys.argv[2]) if len(sys.argv) > 2 else 128)
