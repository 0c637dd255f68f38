"""Where does a DataParallel(cuda_graph=True) step spend its time?  Times each replica's forward / backward graph alone
and both replicas together (CUDA events per device)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from pytorchdistributed_b200.models import resnet50  # noqa: E402
from pytorchdistributed_b200.ops import functional as K  # noqa: E402
from pytorchdistributed_b200.optim import SGD  # noqa: E402
from pytorchdistributed_b200.parallel import DataParallel  # noqa: E402

n = torch.cuda.device_count()
B = 256 * n
x = torch.randn(B, 3, 224, 224, device="cuda:0")
y = torch.randint(0, 1000, (B,), device="cuda:0")
model = resnet50().to("cuda:0").train()
dp = DataParallel(model, criterion=K.cross_entropy, cuda_graph=True)
opt = SGD(model.parameters(), lr=1e-3, model=model)
for _ in range(3):
    opt.zero_grad(); dp(x, y).backward(); opt.step()
for d in range(n):
    torch.cuda.synchronize(d)
gs = dp._graphed[1]


def timed(fns, reps=5):
    """fns: list of (device index, callable); all launched back to back, each timed on its own device."""
    evs = []
    for d, _ in fns:
        with torch.cuda.device(d):
            evs.append((torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)))
    for (d, _), (e0, _) in zip(fns, evs):
        with torch.cuda.device(d):
            e0.record()
    for _ in range(reps):
        for d, f in fns:
            with torch.cuda.device(d):
                f()
    for (d, _), (_, e1) in zip(fns, evs):
        with torch.cuda.device(d):
            e1.record()
    for d in range(n):
        torch.cuda.synchronize(d)
    return [round(e0.elapsed_time(e1) / reps, 2) for e0, e1 in evs]


for i, g in enumerate(gs):
    print(f"replica {i}: fwd alone {timed([(i, g.fwd.replay)])} ms, bwd alone {timed([(i, g.bwd.replay)])} ms", flush=True)
print("all fwd together", timed([(i, g.fwd.replay) for i, g in enumerate(gs)]))
print("all bwd together", timed([(i, g.bwd.replay) for i, g in enumerate(gs)]))
import time
for d in range(n):
    torch.cuda.synchronize(d)
t0 = time.perf_counter()
for _ in range(5):
    opt.zero_grad(); loss = dp(x, y); loss.backward(); opt.step()
t1 = time.perf_counter()
for d in range(n):
    torch.cuda.synchronize(d)
t2 = time.perf_counter()
print(f"5 steps: host issue {1e3 * (t1 - t0) / 5:.2f} ms/step, until done {1e3 * (t2 - t0) / 5:.2f} ms/step")
t0 = time.perf_counter(); dp.replicate(force=True)
for d in range(n):
    torch.cuda.synchronize(d)
print(f"replicate(force) {1e3 * (time.perf_counter() - t0):.2f} ms")
