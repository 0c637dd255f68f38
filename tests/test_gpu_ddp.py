"""DDP / Trainer on ONE GPU (world size 1, process group inside the test process): the gradient sink that lets
weight-gradient kernels write into the bucket views, and whole-step CUDA-graph replay against eager execution.

Deep in a randomly initialised ResNet the bf16 gradients are chaotic (fp32 atomics land in a different order on every
run), so every comparison is calibrated against the distance between two runs of the SAME configuration."""
import os
import socket

import pytest
import torch
import torch.distributed as dist

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.float(), b.float()
    return float((a - b).norm() / (b.norm() + 1e-12))


@pytest.fixture()
def pg():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    yield
    from pytorchdistributed_b200.comm import destroy_process_group
    destroy_process_group()


def _grads(direct: bool, seed_model=0, steps=2):
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.ops import functional as K
    from pytorchdistributed_b200.parallel import DistributedDataParallel as DDP
    torch.manual_seed(seed_model)
    model = resnet50(num_classes=32).cuda().train()
    ddp = DDP(model, device_ids=[0], direct_grads=direct)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(8, 3, 64, 64, generator=g).cuda()
    y = torch.randint(0, 32, (8,), generator=g).cuda()
    out = []
    for p in model.parameters():
        p.grad = None
    # two backward passes over the SAME forward (retain_graph): the second must ADD into the bucket views, not
    # overwrite them - and with identical saved activations the two passes differ only by atomics noise
    loss = K.cross_entropy(ddp(x), y)
    for s in range(steps):
        loss.backward(retain_graph=s + 1 < steps)
        torch.cuda.synchronize()
        out.append({k: p.grad.detach().clone() for k, p in model.named_parameters()})
    stats = dict(ddp.stats)
    del ddp
    return out, stats


def test_direct_bucket_gradients_match_accumulate_grad_path(pg):
    a, sa = _grads(True)
    b, sb = _grads(False)
    c, _ = _grads(False)                 # noise floor of the AccumulateGrad path against itself
    assert sa["direct_grads"] > 100 and sb["direct_grads"] == 0
    for step in range(2):
        for k in a[step]:
            assert torch.isfinite(a[step][k]).all(), k
            noise = _rel(b[step][k], c[step][k])
            assert _rel(a[step][k], b[step][k]) < 3 * noise + 2e-2, (step, k, noise, _rel(a[step][k], b[step][k]))
    # accumulation: the second backward over the same graph doubles every gradient - on the direct path (weight-gradient
    # kernels reducing into the views, BatchNorm sums added by the block's scatter kernel) and on the autograd path
    for k in ("layer4.2.conv3.weight", "layer4.2.bn3.weight", "layer1.0.bn1.weight", "conv1.weight", "fc.weight"):
        ratio = float(a[1][k].float().norm() / a[0][k].float().norm())
        assert 1.7 < ratio < 2.3, (k, ratio)              # added: 2; overwritten: 1
        assert _rel(a[1][k], 2 * a[0][k]) < 0.2, (k, _rel(a[1][k], 2 * a[0][k]))


def _train(cuda_graph: bool, steps: int = 8):
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.optim import SGD
    from pytorchdistributed_b200.train import Trainer
    torch.manual_seed(0)
    model = resnet50(num_classes=32)
    opt = SGD(model.parameters(), lr=1e-3)
    tr = Trainer(model, None, opt, gpu_id=0, quiet=True, cuda_graph=cuda_graph, graph_warmup=2)
    g = torch.Generator().manual_seed(7)
    xs = [torch.randn(8, 3, 64, 64, generator=g).cuda() for _ in range(2)]
    ys = [torch.randint(0, 32, (8,), generator=g).cuda() for _ in range(2)]
    losses = []
    for s in range(steps):
        losses.append(float(tr._run_batch(xs[s % 2], ys[s % 2])))
    torch.cuda.synchronize()
    sd = {k: v.detach().float().clone() for k, v in tr.model.module.state_dict().items()}
    return losses, sd, tr._graph is not None


def test_trainer_whole_step_graph_matches_eager(pg):
    le, sde, ge = _train(False)
    le2, sde2, _ = _train(False)
    lg, sdg, gg = _train(True)
    assert gg and not ge                                  # the graph really was captured and replayed
    noise = 0.0
    for i, (a, b, c) in enumerate(zip(le, le2, lg)):
        noise = max(noise, abs(a - b))                    # run-to-run distance of the eager loop so far
        assert abs(a - c) < 3 * noise + 5e-2 * max(1.0, abs(a)), (i, le, le2, lg)
    for k in sde:
        if "num_batches" in k:
            assert torch.equal(sde[k], sdg[k]), k        # BatchNorm step counters advance inside the replayed graph
            continue
        noise = _rel(sde[k], sde2[k])
        assert _rel(sdg[k], sde[k]) < 3 * noise + 2e-2, (k, noise, _rel(sdg[k], sde[k]))


def test_adam_step_count_advances_under_graph_replay(pg):
    """The bias corrections of Adam depend on the step count; a captured step must not freeze it."""
    from pytorchdistributed_b200.models import MLP
    from pytorchdistributed_b200.optim import Adam
    results = []
    for graphed in (False, True):
        torch.manual_seed(0)
        model = MLP().cuda()
        opt = Adam(model.parameters(), lr=1e-2)
        g = torch.Generator().manual_seed(1)
        x, t = torch.randn(32, 10, generator=g).cuda(), torch.randn(32, 5, generator=g).cuda()

        def step():
            opt.zero_grad(set_to_none=False)
            torch.nn.functional.mse_loss(model(x), t).backward()
            opt.step()

        for _ in range(3):
            step()
        if graphed:
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                step()
            for _ in range(8):                            # capture records the step without running it
                graph.replay()
        else:
            for _ in range(8):
                step()
        torch.cuda.synchronize()
        results.append(([p.detach().clone() for p in model.parameters()], opt.step_count()))
    (pe, se), (pg_, sg) = results
    assert se == sg == 11
    for a, b in zip(pe, pg_):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-5)
