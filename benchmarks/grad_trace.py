"""Where does the backward diverge?  Activation-gradient relative error at every block boundary, ours vs torchvision."""
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


def tap(model, names, store):
    mods = dict(model.named_modules())
    for n in names:
        def hook(mod, inp, out, n=n):
            out.retain_grad()
            store[n] = out
        mods[n].register_forward_hook(hook)


def main(batch=32, img=128):
    torch.manual_seed(0)
    dev = "cuda"
    m = resnet50().to(dev).train()
    with torch.no_grad():
        for p in m.parameters():
            p.copy_(p.to(torch.bfloat16).float())
    tv = torchvision.models.resnet50().to(dev).train()
    tv.load_state_dict(m.state_dict())
    names = ["maxpool"] + [f"layer{l}.{b}" for l, nb in ((1, 3), (2, 4), (3, 6), (4, 3)) for b in range(nb)] + ["avgpool"]
    a, b = {}, {}
    tap(m, names, a)
    tap(tv, names, b)
    x = torch.randn(batch, 3, img, img, device=dev).to(torch.bfloat16).float()
    y = torch.randint(0, 1000, (batch,), device=dev)
    K.cross_entropy(m(x), y).backward()
    F.cross_entropy(tv(x), y).backward()
    print(f"{'boundary':12s} {'act err':>9s} {'grad err':>9s} {'|grad ref|':>11s} {'|grad ours|':>11s}")
    for n in reversed(names):
        ga, gb = a[n].grad, b[n].grad
        print(f"{n:12s} {rel(a[n], b[n].reshape(a[n].shape)):9.4f} {rel(ga, gb.reshape(ga.shape)):9.4f} "
              f"{float(gb.float().norm()):11.3e} {float(ga.float().norm()):11.3e}")
    print("fc.weight grad err", rel(m.fc.weight.grad, tv.fc.weight.grad))
    # control: torch's own bf16 autocast against the same fp32 run - how much divergence is inherent to bf16?
    tv2 = torchvision.models.resnet50().to(dev).train()
    tv2.load_state_dict(m.state_dict())
    c = {}
    tap(tv2, names, c)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out2 = tv2(x)
    F.cross_entropy(out2.float(), y).backward()
    print("control (torch bf16 autocast vs torch fp32):")
    for n in reversed(names):
        print(f"{n:12s} {rel(c[n], b[n]):9.4f} {rel(c[n].grad, b[n].grad):9.4f}")
    print("ours vs torch bf16 autocast:")
    for n in reversed(names):
        print(f"{n:12s} {rel(a[n], c[n].reshape(a[n].shape)):9.4f} {rel(a[n].grad, c[n].grad.reshape(a[n].grad.shape)):9.4f}")


if __name__ == "__main__":
    main()
