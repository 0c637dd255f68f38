"""Multi-GPU tests (need >= 2 devices): peer-to-peer collectives against NCCL, DDP over the p2p backend."""
import pytest
import torch

from tests._mp import run_ranks

pytestmark = pytest.mark.gpu


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


def _collectives(rank, world, mode):
    import os
    os.environ["PTDB200_P2P_MODE"] = mode
    os.environ["PTDB200_AR_ALGO"] = "nvls" if mode == "nvls" else "auto"     # at 2 ranks `auto` prefers two-shot
    os.environ["PTDB200_QUIET"] = "1"
    import torch
    import torch.distributed as dist
    from pytorchdistributed_b200.comm.p2p import P2PCollectives
    dev = torch.device("cuda", rank)
    be = P2PCollectives(None, dev)
    out = {"mode": be.mode, "mode_note": be.mode_note}
    for n in (4, 21, 1000, 262144 + 12, 4 * 1024 * 1024 + 4):
        for dtype in (torch.float32, torch.bfloat16):
            flat = be.alloc(n, dtype, dev)
            g = torch.Generator(device="cpu").manual_seed(100 * rank + n)
            # integers in float: the sum is exact in any order, so the comparison with NCCL is bit-exact
            vals = torch.randint(-8, 9, (n,), generator=g).to(dtype).to(dev)
            flat.copy_(vals)
            ref = vals.clone().float()
            dist.all_reduce(ref)
            be.all_reduce_avg(flat, scale=1.0)
            torch.cuda.synchronize()
            out[(n, str(dtype), "sum_exact")] = bool(torch.equal(flat.float(), ref))
            flat.copy_(vals)
            be.all_reduce_avg(flat)
            torch.cuda.synchronize()
            out[(n, str(dtype), "avg")] = bool(torch.allclose(flat.float(), ref / world, rtol=1e-2, atol=1e-2))
            if dtype == torch.float32:
                # bf16 on the wire: small integers (and their sums, and the power-of-two average) are exact in bf16
                flat.copy_(vals)
                be.all_reduce_avg(flat, scale=1.0, wire=True)
                torch.cuda.synchronize()
                out[(n, "wire_bf16", "sum_exact")] = bool(torch.equal(flat, ref))
                # random data: every rank ends with the same values, close to the fp32 sum
                r = torch.randn(n, generator=g).to(dev)
                flat.copy_(r)
                exact = r.clone()
                dist.all_reduce(exact)
                be.all_reduce_avg(flat, wire=True)
                torch.cuda.synchronize()
                out[(n, "wire_bf16", "close")] = bool(torch.allclose(flat, exact / world, rtol=2e-2, atol=2e-2))
                lo, hi = flat.clone(), flat.clone()
                dist.all_reduce(lo, op=dist.ReduceOp.MIN)
                dist.all_reduce(hi, op=dist.ReduceOp.MAX)
                out[(n, "wire_bf16", "identical_on_all_ranks")] = bool(torch.equal(lo, hi))
    # back-to-back calls of every kind with no host sync in between
    flat = be.alloc(1 << 20, torch.float32, dev)
    flat.fill_(float(rank + 1))
    for i in range(20):
        be.all_reduce_avg(flat, wire=(i % 5 == 4))
    torch.cuda.synchronize()
    exp = sum(range(1, world + 1)) / world
    out["repeat"] = bool(torch.allclose(flat, torch.full_like(flat, exp), rtol=1e-2))
    b = be.alloc(100003, torch.float32, dev)
    b.fill_(float(rank))
    be.broadcast(b, src=world - 1)
    torch.cuda.synchronize()
    out["bcast"] = bool((b == world - 1).all())
    tmp = torch.full((777,), float(rank), device=dev)          # not symmetric memory: staged
    be.broadcast(tmp, src=0)
    torch.cuda.synchronize()
    out["bcast_staged"] = bool((tmp == 0).all())
    out["stats"] = dict(be.stats)
    be.close()
    return out


@pytest.mark.parametrize("mode", ["nvls", "vmm", "ipc"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_p2p_collectives_match_nccl(world, mode):
    _need(world)
    if mode == "vmm" and world != 2:
        pytest.skip("the unicast-VMM provider runs the same kernels as ipc: covered once, at 2 ranks")
    results = run_ranks(_collectives, world, (mode,), backend="nccl")
    for res in results:
        bad = [k for k, v in res.items() if isinstance(k, tuple) and not v] + [k for k in ("repeat", "bcast", "bcast_staged") if not res[k]]
        assert not bad, (res["mode"], bad)
    got = results[0]["mode"]
    if got != mode:
        pytest.skip(f"asked for {mode}, the box granted {got} ({results[0]['mode_note']}); that path passed")
    if mode == "nvls":
        assert results[0]["stats"]["nvls"] > 0 and results[0]["stats"]["two_shot"] == 0


def _sharded(rank, world):
    import torch
    import torch.distributed as dist
    from pytorchdistributed_b200.comm.p2p import P2PCollectives
    dev = torch.device("cuda", rank)
    be = P2PCollectives(None, dev)
    out = {}
    for n in (8, 1001, 65536 + 20, 3 * 1024 * 1024 + 4):
        def fresh(seed_rank=rank):
            g = torch.Generator(device="cpu").manual_seed(1000 * seed_rank + n)
            return torch.randint(-8, 9, (n,), generator=g).float().to(dev)       # exact sums in any order
        every = [fresh(r) for r in range(world)]
        total = torch.stack(every).sum(0)
        flat = be.alloc(n, torch.float32, dev)
        sl = [be.shard(flat, r) for r in range(world)]
        lo = [s_.storage_offset() - flat.storage_offset() for s_ in sl]
        out[(n, "shards_cover")] = sum(s_.numel() for s_ in sl) == n
        # all_gather: every rank contributes its own shard of rank-specific data
        flat.fill_(-1.0)
        be.shard(flat).copy_(every[rank][lo[rank]:lo[rank] + sl[rank].numel()])
        be.all_gather(flat)
        exp = torch.cat([every[r][lo[r]:lo[r] + sl[r].numel()] for r in range(world)])
        torch.cuda.synchronize()
        out[(n, "all_gather")] = bool(torch.equal(flat, exp))
        # gather to the last rank
        flat.fill_(-1.0)
        be.shard(flat).copy_(every[rank][lo[rank]:lo[rank] + sl[rank].numel()])
        be.gather(flat, dst=world - 1)
        torch.cuda.synchronize()
        if rank == world - 1:
            out[(n, "gather")] = bool(torch.equal(flat, exp))
        else:
            out[(n, "gather")] = bool((flat[:lo[rank]] == -1).all())
        # scatter from rank 0
        flat.copy_(every[rank])
        be.scatter(flat, src=0)
        torch.cuda.synchronize()
        out[(n, "scatter")] = bool(torch.equal(be.shard(flat), every[0][lo[rank]:lo[rank] + sl[rank].numel()]))
        # reduce_scatter
        flat.copy_(every[rank])
        be.reduce_scatter(flat)
        torch.cuda.synchronize()
        out[(n, "reduce_scatter")] = bool(torch.equal(be.shard(flat), total[lo[rank]:lo[rank] + sl[rank].numel()]))
        # reduce to rank 0
        flat.copy_(every[rank])
        be.reduce(flat, dst=0)
        torch.cuda.synchronize()
        out[(n, "reduce")] = bool(torch.equal(flat, total if rank == 0 else every[rank]))
        # one-shot all-reduce against NCCL
        flat.copy_(every[rank])
        if n * 4 <= (4 << 20):
            ref = every[rank].clone()
            dist.all_reduce(ref)
            be.all_reduce_oneshot(flat, scale=1.0)
            torch.cuda.synchronize()
            out[(n, "oneshot")] = bool(torch.equal(flat, ref))
        else:                                   # larger than the grid's registers: refused, not truncated
            try:
                be.all_reduce_oneshot(flat, scale=1.0)
                out[(n, "oneshot_refused")] = False
            except RuntimeError:
                out[(n, "oneshot_refused")] = True
    # reduce_scatter followed by all_gather == all-reduce, back to back without host syncs
    flat = be.alloc(1 << 18, torch.float32, dev)
    for it in range(10):
        flat.fill_(float(rank + it))
        be.reduce_scatter(flat, scale=1.0 / world)
        be.all_gather(flat)
    torch.cuda.synchronize()
    out["rs_ag"] = bool(torch.allclose(flat, torch.full_like(flat, sum(r + 9 for r in range(world)) / world)))
    be.close()
    return out


@pytest.mark.parametrize("world", [2, 4])
def test_p2p_sharded_collectives(world):
    _need(world)
    for res in run_ranks(_sharded, world, backend="nccl"):
        bad = [k for k, v in res.items() if not v]
        assert not bad, bad


def _ddp(rank, world, comm):
    import torch
    import torch.nn.functional as F
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.ops import functional as K
    from pytorchdistributed_b200.optim import SGD
    from pytorchdistributed_b200.parallel import DistributedDataParallel as DDP
    dev = torch.device("cuda", rank)
    torch.manual_seed(rank)
    model = resnet50(num_classes=64).to(dev).train()
    ddp = DDP(model, device_ids=[rank], comm=comm)
    opt = SGD(model.parameters(), lr=0.05, model=model)
    g = torch.Generator().manual_seed(5)
    xs = torch.randn(3, 4 * world, 3, 64, 64, generator=g)
    ys = torch.randint(0, 64, (3, 4 * world), generator=g)
    losses = []
    for s in range(3):
        opt.zero_grad()
        x = xs[s, rank * 4:(rank + 1) * 4].to(dev)
        y = ys[s, rank * 4:(rank + 1) * 4].to(dev)
        loss = K.cross_entropy(ddp(x), y)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    torch.cuda.synchronize()
    sd = {k: v.detach().float().cpu() for k, v in model.state_dict().items()}
    return losses, sd, dict(ddp.stats), ddp.backend.name


@pytest.mark.parametrize("world", [2])
def test_ddp_p2p_replicas_identical_and_match_nccl_path(world):
    _need(world)
    p2p = run_ranks(_ddp, world, ("p2p",), backend="nccl")
    nccl = run_ranks(_ddp, world, ("torch",), backend="nccl")
    assert p2p[0][3] == "p2p" and nccl[0][3] == "torch"
    for k in p2p[0][1]:
        if "running_" in k or "num_batches" in k:
            continue              # BN statistics are per-rank until the next forward re-broadcasts rank 0's
        assert torch.equal(p2p[0][1][k], p2p[1][1][k]), k            # parameter replicas stay bit-identical
    # same training through NCCL: split-K atomics make conv grads non-deterministic in the last bits,
    # so compare loosely
    num = sum(float((p2p[0][1][k] - nccl[0][1][k]).norm()) for k in p2p[0][1] if "running" not in k and "num_b" not in k)
    den = sum(float(nccl[0][1][k].norm()) for k in nccl[0][1] if "running" not in k and "num_b" not in k)
    assert num / den < 0.3        # three chaotic bf16 steps at lr 0.05: same ballpark, not bitwise
    assert p2p[0][2]["iterations"] == 3 and p2p[0][2]["allreduce_launches"] >= 3


def _ddp_fused_sgd(rank, world, model_name, fuse, native):
    """DDP over the p2p backend with SGD(momentum, weight decay): the step applied by the all-reduce kernels' epilogue
    (fuse) against the separate optimizer kernel, and the native launch path against the Python one."""
    import os
    os.environ["PTDB200_QUIET"] = "1"
    os.environ["PTDB200_DDP_NATIVE"] = "1" if native else "0"
    import torch
    from pytorchdistributed_b200.models import MLP, resnet50
    from pytorchdistributed_b200.ops import functional as K
    from pytorchdistributed_b200.optim import SGD
    from pytorchdistributed_b200.parallel import DistributedDataParallel as DDP
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    if model_name == "mlp":
        model = MLP().to(dev).train()
        g = torch.Generator().manual_seed(5)
        xs = torch.randn(6, 16 * world, 10, generator=g)
        ts = torch.randn(6, 16 * world, 5, generator=g)
        loss_fn = torch.nn.functional.mse_loss
        per = 16
    else:
        model = resnet50(num_classes=32).to(dev).train()
        g = torch.Generator().manual_seed(5)
        xs = torch.randn(4, 4 * world, 3, 64, 64, generator=g)
        ts = torch.randint(0, 32, (4, 4 * world), generator=g)
        loss_fn = K.cross_entropy
        per = 4
    ddp = DDP(model, device_ids=[rank], comm="p2p")
    opt = SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-3, model=model)
    fused = ddp.fuse_optimizer(opt) if fuse else False
    losses = []
    for s in range(xs.size(0)):
        opt.zero_grad()
        x = xs[s, rank * per:(rank + 1) * per].to(dev)
        t = ts[s, rank * per:(rank + 1) * per].to(dev)
        loss = loss_fn(ddp(x), t)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    # the bf16 shadows the next forward would read must be the cast of the final parameters
    shadow_ok = True
    for m in model.modules():
        if hasattr(m, "shadow") and not getattr(m, "is_stem", False) and not getattr(m, "small", False):
            shadow_ok = shadow_ok and bool(torch.equal(m.shadow().float(), m.weight.detach().to(torch.bfloat16).float()))
    torch.cuda.synchronize()
    mom = [opt.state[p]["momentum_buffer"].detach().cpu() for p in model.parameters()]
    return losses, [p.detach().cpu() for p in model.parameters()], mom, fused, ddp._native, shadow_ok, dict(ddp.stats)


def test_ddp_fused_sgd_epilogue_and_native_reducer_match_unfused():
    _need(2)
    rel = lambda a, b: float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))  # noqa: E731
    # deterministic model: the fused epilogue, the native launch path and the Python launch path agree tightly
    base = run_ranks(_ddp_fused_sgd, 2, ("mlp", False, False), backend="nccl")
    nat = run_ranks(_ddp_fused_sgd, 2, ("mlp", False, True), backend="nccl")
    fus = run_ranks(_ddp_fused_sgd, 2, ("mlp", True, True), backend="nccl")
    assert not base[0][4] and nat[0][4] and fus[0][4] and fus[0][3] and not nat[0][3]
    for other in (nat, fus):
        for a, b in zip(other[0][1], base[0][1]):
            assert torch.allclose(a, b, rtol=1e-5, atol=1e-6), float((a - b).abs().max())
        for a, b in zip(other[0][2], base[0][2]):
            assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)                      # momentum buffers too
        for a, b in zip(other[0][1], other[1][1]):
            assert torch.equal(a, b)                                                # replicas bit-identical
    # ResNet-50: within the unfused path's own run-to-run noise; shadows consistent with the parameters
    r0 = run_ranks(_ddp_fused_sgd, 2, ("resnet", False, True), backend="nccl")
    r1 = run_ranks(_ddp_fused_sgd, 2, ("resnet", False, True), backend="nccl")
    rf = run_ranks(_ddp_fused_sgd, 2, ("resnet", True, True), backend="nccl")
    assert rf[0][3] and rf[0][5] and r0[0][5]
    for a, b in zip(rf[0][1], rf[1][1]):
        assert torch.equal(a, b)
    for a, b, c in zip(rf[0][1], r0[0][1], r1[0][1]):
        assert rel(a, b) < 3 * rel(c, b) + 2e-2, (rel(a, b), rel(c, b))
    assert abs(rf[0][0][0] - r0[0][0][0]) < 5e-2           # same first loss (same weights, same batch)


def _ddp_mlp(rank, world, comm, wire):
    """Deterministic little model (fp32 CUDA-core kernels, MSE): p2p and NCCL gradient averaging must agree tightly."""
    import os
    os.environ["PTDB200_QUIET"] = "1"
    import torch
    from pytorchdistributed_b200.models import MLP
    from pytorchdistributed_b200.optim import SGD
    from pytorchdistributed_b200.parallel import DistributedDataParallel as DDP
    dev = torch.device("cuda", rank)
    torch.manual_seed(100 + rank)                     # different init per rank: rank 0's must win
    model = MLP().to(dev).train()
    ddp = DDP(model, device_ids=[rank], comm=comm, wire_dtype=torch.bfloat16 if wire else None)
    opt = SGD(model.parameters(), lr=0.05, momentum=0.9)
    g = torch.Generator().manual_seed(5)
    xs = torch.randn(10, 16 * world, 10, generator=g)
    ts = torch.randn(10, 16 * world, 5, generator=g)
    losses = []
    for s in range(10):
        opt.zero_grad()
        x = xs[s, rank * 16:(rank + 1) * 16].to(dev)
        t = ts[s, rank * 16:(rank + 1) * 16].to(dev)
        loss = torch.nn.functional.mse_loss(ddp(x), t)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    torch.cuda.synchronize()
    return losses, [p.detach().cpu() for p in model.parameters()], getattr(ddp.backend, "mode", "torch"), \
        dict(getattr(ddp.backend, "stats", {}))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_ddp_p2p_matches_nccl_on_deterministic_model(world):
    _need(world)
    p2p = run_ranks(_ddp_mlp, world, ("p2p", False), backend="nccl")
    nccl = run_ranks(_ddp_mlp, world, ("torch", False), backend="nccl")
    wire = run_ranks(_ddp_mlp, world, ("p2p", True), backend="nccl")
    for r in range(1, world):
        for a, b in zip(p2p[0][1], p2p[r][1]):
            assert torch.equal(a, b)                         # replicas bit-identical
        for a, b in zip(wire[0][1], wire[r][1]):
            assert torch.equal(a, b)                         # also with bf16 on the wire (the slice owner rounds once)
    for a, b in zip(p2p[0][1], nccl[0][1]):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-5), float((a - b).abs().max())
    for a, b in zip(wire[0][1], nccl[0][1]):
        assert torch.allclose(a, b, rtol=5e-2, atol=5e-3), float((a - b).abs().max())     # bf16 gradients: ~3 digits
    assert wire[0][3].get("wire_bf16", 0) > 0 and p2p[0][3].get("wire_bf16", 0) == 0
    assert max(abs(x - y) for x, y in zip(p2p[0][0], nccl[0][0])) < 1e-4


def _loss_curve(rank, world, impl):
    """20 steps of ResNet-50 on a fixed batch (so the loss must fall), ours vs torchvision + torch DDP + bf16 autocast."""
    import os
    os.environ["PTDB200_QUIET"] = "1"
    import torch
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    from pytorchdistributed_b200.models import resnet50
    ours = resnet50(num_classes=16)
    init = {k: v.clone() for k, v in ours.state_dict().items()}
    g = torch.Generator().manual_seed(11)
    x = torch.randn(8 * world, 3, 64, 64, generator=g)[rank * 8:(rank + 1) * 8].to(dev)
    y = torch.randint(0, 16, (8 * world,), generator=g)[rank * 8:(rank + 1) * 8].to(dev)
    if impl == "ours":
        from pytorchdistributed_b200.ops import functional as K
        from pytorchdistributed_b200.optim import SGD
        from pytorchdistributed_b200.parallel import DistributedDataParallel as DDP
        model = ours.to(dev).train()
        ddp = DDP(model, device_ids=[rank])
        opt = SGD(model.parameters(), lr=0.003, momentum=0.9)
        fwd, loss_fn = ddp, K.cross_entropy
    else:
        import torchvision
        model = torchvision.models.resnet50(num_classes=16)
        model.load_state_dict(init)                                   # same initial weights: the layouts are compatible
        model = model.to(dev).to(memory_format=torch.channels_last).train()
        ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[rank])
        opt = torch.optim.SGD(model.parameters(), lr=0.003, momentum=0.9)

        def fwd(inp):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return ddp(inp.contiguous(memory_format=torch.channels_last)).float()
        loss_fn = torch.nn.functional.cross_entropy
    losses = []
    for _ in range(20):
        opt.zero_grad()
        loss = loss_fn(fwd(x), y)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    torch.cuda.synchronize()
    return losses


def test_ddp_resnet_loss_curve_tracks_torch_ddp():
    _need(2)
    ours = run_ranks(_loss_curve, 2, ("ours",), backend="nccl")[0]
    ref = run_ranks(_loss_curve, 2, ("torch",), backend="nccl")[0]
    print("ours :", [round(v, 3) for v in ours])
    print("torch:", [round(v, 3) for v in ref])
    assert abs(ours[0] - ref[0]) < 0.15                               # same weights, same batch: same first loss (bf16, BN over 8 rows)
    assert ours[-1] < ours[0] - 0.3 and ref[-1] < ref[0] - 0.3        # both learn the fixed batch
    diff = [abs(a - b) for a, b in zip(ours, ref)]
    assert sum(diff) / len(diff) < 0.25, diff                         # two bf16 runs of a BN network drift, not diverge


def test_pipeline_cuda_graph_matches_eager_pipeline():
    """Graph-replayed pipeline stages (both stages on one GPU here: the graph logic does not care) behave like the
    eager pipeline, step after step, uneven last split included.

    Deep in a randomly initialised ResNet with 4-image micro-batches the gradients are chaotic: two EAGER runs of
    the same model on the same batch differ by O(1) there (fp32 atomics land in a different order and 50
    BatchNorms over 16-element statistics amplify it).  So a second eager model measures that noise per parameter
    and the graphed model has to stay within a small multiple of it; on top: every gradient present, finite and of
    the same magnitude - what a missing or doubled accumulation would break."""
    from pytorchdistributed_b200.ops import functional as K
    from pytorchdistributed_b200.parallel import PipelineParallelResNet50
    torch.manual_seed(0)
    dev = ("cuda:0", "cuda:0")
    mk = lambda **kw: PipelineParallelResNet50(split_size=4, num_classes=32, devices=dev, transfer="p2p", **kw).train()
    eager, eager2, graph = mk(), mk(), mk(cuda_graph=True)
    eager2.load_state_dict(eager.state_dict())
    graph.load_state_dict(eager.state_dict())
    models = (eager, eager2, graph)
    opts = [torch.optim.SGD(m.parameters(), lr=1e-3) for m in models]
    g = torch.Generator().manual_seed(1)
    rel = lambda a, b: float((a - b).norm() / (a.norm() + 1e-6))
    for step in range(3):
        x = torch.randn(10, 3, 64, 64, generator=g).cuda()            # 4 + 4 + 2: the tail runs eagerly
        y = torch.randint(0, 32, (10,), generator=g).cuda()
        losses = []
        for model, opt in zip(models, opts):
            opt.zero_grad()
            loss = K.cross_entropy(model(x), y)
            loss.backward()
            losses.append(float(loss.detach()))
        # one |eager - eager2| sample is itself noisy (it can come out near zero): allow 5 % of the loss on top
        assert abs(losses[0] - losses[2]) < 3 * abs(losses[0] - losses[1]) + 5e-2 * max(1.0, abs(losses[0])), (step, losses)
        for (k, pe), p2, pg in zip(eager.named_parameters(), eager2.parameters(), graph.parameters()):
            assert pg.grad is not None and torch.isfinite(pg.grad).all(), k
            ne, ng = float(pe.grad.norm()), float(pg.grad.norm())
            assert 0.4 * ne <= ng <= 2.5 * ne + 1e-6, (step, k, ne, ng)
            noise = rel(pe.grad, p2.grad)
            assert rel(pe.grad, pg.grad) < 3 * noise + 0.05, (step, k, noise, rel(pe.grad, pg.grad))
        for opt in opts:
            opt.step()
    assert graph._graphs is not None and graph._graphs[0] == 2
    sd_e, sd_2, sd_g = eager.state_dict(), eager2.state_dict(), graph.state_dict()
    for k in sd_e:
        if "running" in k or "num_batches" in k:        # BatchNorm buffers are updated by the replayed graphs too
            noise = rel(sd_e[k].float(), sd_2[k].float())
            assert rel(sd_e[k].float(), sd_g[k].float()) < 3 * noise + 0.05, (k, noise)


def test_data_parallel_forward_backward_matches_single_device():
    _need(2)
    from pytorchdistributed_b200.models import MLP
    from pytorchdistributed_b200.parallel import DataParallel
    torch.manual_seed(0)
    n = min(torch.cuda.device_count(), 4)
    ref = MLP().cuda(0)
    model = MLP().cuda(0)
    model.load_state_dict(ref.state_dict())
    dp = DataParallel(model, device_ids=list(range(n)))
    x = torch.randn(32, 10, device="cuda:0")
    t = torch.randn(32, 5, device="cuda:0")
    out = dp(x)
    assert out.shape == (32, 5) and out.device.index == 0          # notebook 01: "Outside: ... output_size [32, 5]"
    assert torch.allclose(out, ref(x), atol=1e-5)
    torch.nn.functional.mse_loss(out, t).backward()
    torch.nn.functional.mse_loss(ref(x), t).backward()
    for p, q in zip(model.parameters(), ref.parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-5)
    # replicas are refreshed only when the root parameters change
    c0 = dp.stats["replicate_copies"]
    dp(x)
    assert dp.stats["replicate_copies"] == c0 and dp.stats["replicate_skipped"] >= 1
    with torch.no_grad():
        for p in model.parameters():
            p.add_(1.0)
        for p in ref.parameters():
            p.add_(1.0)
    assert torch.allclose(dp(x), ref(x), atol=1e-4)
    assert dp.stats["replicate_copies"] > c0


def test_data_parallel_resnet_gradients_and_second_step_match_single_device():
    """Gradient parity (not just "finite") and the step after an optimizer update: replicas must run on the UPDATED
    weights (their bf16 shadows are rebuilt after the raw peer copy), else the second loss comes out different."""
    _need(2)
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.ops import functional as K
    from pytorchdistributed_b200.optim import SGD
    from pytorchdistributed_b200.parallel import DataParallel
    rel = lambda a, b: float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))  # noqa: E731
    torch.manual_seed(0)
    models = [resnet50(num_classes=32).cuda(0).train() for _ in range(3)]       # dp, single, single again (noise)
    for m in models[1:]:
        m.load_state_dict(models[0].state_dict())
    dp = DataParallel(models[0], device_ids=[0, 1], criterion=K.cross_entropy)
    opts = [SGD(m.parameters(), lr=0.1) for m in models]
    g = torch.Generator().manual_seed(2)
    x = torch.randn(8, 3, 64, 64, generator=g).cuda(0)
    y = torch.randint(0, 32, (8,), generator=g).cuda(0)

    def single(m):
        # what two replicas compute: each half on its own (its own BatchNorm statistics), losses weighted by share
        total = 0.0
        for lo in (0, 4):
            l = K.cross_entropy(m(x[lo:lo + 4]), y[lo:lo + 4]) * 0.5
            l.backward()
            total += float(l)
        return total

    losses = []
    for step in range(2):
        for o in opts:
            o.zero_grad()
        l_dp = dp(x, y)
        l_dp.backward()
        l_s, l_n = single(models[1]), single(models[2])
        losses.append((float(l_dp), l_s, l_n))
        if step == 0:
            for (k, p), q, r in zip(models[0].named_parameters(), models[1].parameters(), models[2].parameters()):
                noise = rel(q.grad, r.grad)
                assert rel(p.grad, q.grad) < 3 * noise + 2e-2, (k, noise, rel(p.grad, q.grad))
        for o in opts:
            o.step()
    assert abs(losses[0][0] - losses[0][1]) < 3 * abs(losses[0][1] - losses[0][2]) + 2e-2 * max(1.0, abs(losses[0][1])), losses
    # the regression itself: after the update the replica's bf16 weight copy is the cast of the NEW root weights
    root_w = models[0].layer1[0].conv1.weight.detach()
    rep_mod = dp.replicas[1].module.layer1[0].conv1
    dp(x, y)                                            # replicate + forward on the updated weights
    torch.cuda.synchronize(1)
    assert rep_mod._shadow is not None                  # (the replica's fp32 master is not refreshed: nothing reads it)
    assert torch.equal(rep_mod._shadow.float().cpu(), root_w.to(torch.bfloat16).float().cpu())
    # and the loss of the second step follows the single-device run
    assert abs(losses[1][0] - losses[1][1]) < 3 * abs(losses[1][1] - losses[1][2]) + 0.05 * max(1.0, abs(losses[1][1])), losses
    assert dp.stats["replicate_copies"] >= 2


def test_data_parallel_cuda_graph_matches_eager():
    """Graph-replayed replicas (one forward and one backward graph per GPU, launched from one thread) train like the
    thread-pool eager path: losses and root gradients within the eager path's own run-to-run noise, three steps."""
    _need(2)
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.ops import functional as K
    from pytorchdistributed_b200.optim import SGD
    from pytorchdistributed_b200.parallel import DataParallel
    rel = lambda a, b: float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))  # noqa: E731
    torch.manual_seed(0)
    models = [resnet50(num_classes=32).cuda(0).train() for _ in range(3)]
    for m in models[1:]:
        m.load_state_dict(models[0].state_dict())
    dps = [DataParallel(models[0], device_ids=[0, 1], criterion=K.cross_entropy),
           DataParallel(models[1], device_ids=[0, 1], criterion=K.cross_entropy),
           DataParallel(models[2], device_ids=[0, 1], criterion=K.cross_entropy, cuda_graph=True)]
    opts = [SGD(m.parameters(), lr=1e-3, model=m) for m in models]
    g = torch.Generator().manual_seed(4)
    noise_l = 0.0
    for step in range(3):
        x = torch.randn(8, 3, 64, 64, generator=g).cuda(0)
        y = torch.randint(0, 32, (8,), generator=g).cuda(0)
        losses = []
        for dp, opt in zip(dps, opts):
            opt.zero_grad()
            loss = dp(x, y)
            loss.backward()
            losses.append(float(loss.detach()))
        noise_l = max(noise_l, abs(losses[0] - losses[1]))
        assert abs(losses[0] - losses[2]) < 3 * noise_l + 5e-2 * max(1.0, abs(losses[0])), (step, losses)
        for (k, pe), p2, pg in zip(models[0].named_parameters(), models[1].parameters(), models[2].parameters()):
            assert pg.grad is not None and torch.isfinite(pg.grad).all(), k
            noise = rel(pe.grad, p2.grad)
            assert rel(pg.grad, pe.grad) < 3 * noise + 5e-2, (step, k, noise, rel(pg.grad, pe.grad))
        for opt in opts:
            opt.step()
    assert dps[2].stats["graph_captures"] == 1 and dps[2].stats["graph_replays"] == 3
    for k, v in models[0].state_dict().items():
        if "num_batches" in k:
            assert torch.equal(v, models[2].state_dict()[k]), k


def test_data_parallel_fused_loss_resnet():
    _need(2)
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.ops import functional as K
    from pytorchdistributed_b200.parallel import DataParallel
    torch.manual_seed(0)
    model = resnet50(num_classes=64).cuda(0).train()
    dp = DataParallel(model, device_ids=[0, 1], criterion=K.cross_entropy)
    x = torch.randn(8, 3, 64, 64, device="cuda:0")
    y = torch.randint(0, 64, (8,), device="cuda:0")
    loss = dp(x, y)
    assert loss.dim() == 0 and 3.0 < float(loss) < 6.0             # ~ln(64) at random init
    loss.backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.parameters())
    assert float(model.fc.weight.grad.abs().sum()) > 0


@pytest.mark.parametrize("transfer", ["nccl", "p2p"])
def test_pipeline_matches_unsplit_model(transfer):
    _need(2)
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.ops import functional as K
    from pytorchdistributed_b200.parallel import ModelParallelResNet50, PipelineParallelResNet50
    torch.manual_seed(0)
    pp = PipelineParallelResNet50(split_size=3, num_classes=32, transfer=transfer).eval()
    single = resnet50(num_classes=32).cuda(0).eval()
    single.load_state_dict(pp.state_dict())
    x = torch.randn(8, 3, 64, 64, device="cuda:0")
    with torch.no_grad():
        a, b = pp(x), single(x)
    assert a.device.index == 1 and a.shape == (8, 32)
    assert torch.allclose(a.cpu(), b.cpu(), atol=2e-2, rtol=2e-2)   # eval mode: running stats, deterministic
    # training step: finite loss, every parameter (on both devices) gets a gradient
    pp.train()
    y = torch.randint(0, 32, (8,), device="cuda:1")
    loss = K.cross_entropy(pp(x), y)
    loss.backward()
    assert torch.isfinite(loss)
    devs = {p.device.index for p in pp.parameters()}
    assert devs == {0, 1}
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in pp.parameters())
    mp_ = ModelParallelResNet50(num_classes=32, transfer=transfer).eval()
    mp_.load_state_dict(pp.state_dict())
    with torch.no_grad():
        assert torch.allclose(mp_(x).cpu(), pp.eval()(x).cpu(), atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("split,graph", [(2, False), (3, False), (8, False), (3, True)])
def test_pipeline_gradients_match_unsplit_model_per_split_size(split, graph):
    """Training-mode gradient parity per split size (3 leaves an uneven tail of 2): the pipeline must equal the
    UNSPLIT model run micro-batch by micro-batch (BatchNorm sees micro-batch statistics in both), within the
    distance between two runs of that reference; output rows land in one preallocated tensor (no concatenation)."""
    _need(2)
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.ops import functional as K
    from pytorchdistributed_b200.parallel import PipelineParallelResNet50
    rel = lambda a, b: float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))  # noqa: E731
    torch.manual_seed(0)
    pp = PipelineParallelResNet50(split_size=split, num_classes=32, transfer="p2p", cuda_graph=graph).train()
    refs = [resnet50(num_classes=32).cuda(0).train() for _ in range(2)]
    for r in refs:
        r.load_state_dict(pp.state_dict())
    g = torch.Generator().manual_seed(9)
    x = torch.randn(8, 3, 64, 64, generator=g).cuda(0)
    y = torch.randint(0, 32, (8,), generator=g)
    for it in range(2 if graph else 1):              # graph mode: the second pass replays what the first captured
        for m in [pp] + refs:
            for p in m.parameters():
                p.grad = None
        out = pp(x)
        assert out.shape == (8, 32) and out.device.index == 1
        loss = K.cross_entropy(out, y.cuda(1))
        loss.backward()
        ref_losses = []
        for r in refs:
            # cross_entropy is the mean over rows, so micro-batch losses are weighted by their share of the batch
            total = 0.0
            for lo in range(0, 8, split):
                hi = min(8, lo + split)
                l = K.cross_entropy(r(x[lo:hi]), y[lo:hi].cuda(0)) * ((hi - lo) / 8)
                l.backward()
                total += float(l.detach())
            ref_losses.append(total)
        noise_l = abs(ref_losses[0] - ref_losses[1])
        assert abs(float(loss.detach()) - ref_losses[0]) < 3 * noise_l + 3e-2 * max(1.0, abs(ref_losses[0])), (it, float(loss), ref_losses)
        for (k, p), q, q2 in zip(pp.named_parameters(), refs[0].parameters(), refs[1].parameters()):
            assert p.grad is not None and torch.isfinite(p.grad).all(), k
            noise = rel(q.grad, q2.grad)
            assert rel(p.grad.cpu(), q.grad.cpu()) < 3 * noise + 5e-2, (it, k, noise, rel(p.grad.cpu(), q.grad.cpu()))


def _stage_run_gpu(rank, world, schedule, recompute):
    """Process-per-stage pipeline on GPUs: rank r owns one stage of ResNet-50, activations / gradients travel with
    NCCL send/recv; every rank also runs the FULL model on its own GPU as the reference for its stage's gradients."""
    import torch
    from pytorchdistributed_b200.models import resnet50
    from pytorchdistributed_b200.ops import functional as K
    from pytorchdistributed_b200.parallel import PipelineStage
    from pytorchdistributed_b200.parallel.pipeline import _Stage2
    dev = torch.device("cuda", rank)
    rel = lambda a, b: float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))  # noqa: E731
    torch.manual_seed(0)
    models = [resnet50(num_classes=32).to(dev).train() for _ in range(3)]          # staged, reference, reference again
    for m in models[1:]:
        m.load_state_dict(models[0].state_dict())
    seq1, seq2, fc = models[0].stage_modules()
    part = seq1 if rank == 0 else _Stage2(seq2, fc)
    stage = PipelineStage(part, rank, world, device=dev, loss_fn=K.cross_entropy, schedule=schedule, recompute=recompute)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(10, 3, 64, 64, generator=g)
    y = torch.randint(0, 32, (10,), generator=g)
    loss = stage.train_step(x if rank == 0 else None, y if rank == world - 1 else None, split_size=4, batch_rows=10)
    ref_losses = []
    for m in models[1:]:
        total = 0.0
        for lo in range(0, 10, 4):
            hi = min(10, lo + 4)
            l = K.cross_entropy(m(x[lo:hi].to(dev)), y[lo:hi].to(dev)) * ((hi - lo) / 10)
            l.backward()
            total += float(l.detach())
        ref_losses.append(total)
    torch.cuda.synchronize(dev)
    mine = {id(p) for p in part.parameters()}
    worst = 0.0
    for (k, p), q, q2 in zip(models[0].named_parameters(), models[1].parameters(), models[2].parameters()):
        if id(p) not in mine:
            continue
        assert p.grad is not None and torch.isfinite(p.grad).all(), k
        worst = max(worst, rel(p.grad, q.grad) - 3 * rel(q.grad, q2.grad))
    return (float(loss) if loss is not None else None), ref_losses, worst


@pytest.mark.parametrize("schedule,recompute", [("1f1b", False), ("gpipe", True)])
def test_process_per_stage_pipeline_on_gpus(schedule, recompute):
    _need(2)
    res = run_ranks(_stage_run_gpu, 2, (schedule, recompute), backend="nccl")
    loss, ref_losses, _ = res[1]
    assert abs(loss - ref_losses[0]) < 3 * abs(ref_losses[0] - ref_losses[1]) + 3e-2 * max(1.0, abs(ref_losses[0])), (loss, ref_losses)
    for rank, (_, _, worst) in enumerate(res):
        assert worst < 5e-2, (rank, worst)            # beyond 3x the reference's own run-to-run noise by at most 5 %


def _stress(rank, world, mode):
    """>= 1000 collectives of mixed kinds and sizes back to back, on both signal pads, with NO host synchronisation in
    between; every buffer is poisoned with a rank-specific value right before its collective and checked on the device
    right after it (a mismatch bumps a device-side counter).  What a lost flag, a stale sequence number, a reused
    staging buffer or a missing fence would turn into a wrong value or a time-out."""
    import os
    os.environ["PTDB200_P2P_MODE"] = mode
    os.environ["PTDB200_AR_ALGO"] = "nvls" if mode == "nvls" else "auto"
    os.environ["PTDB200_QUIET"] = "1"
    os.environ["PTDB200_P2P_TIMEOUT_S"] = "30"
    import torch
    from pytorchdistributed_b200.comm.p2p import P2PCollectives
    dev = torch.device("cuda", rank)
    be = P2PCollectives(None, dev)
    sizes = [8, 1000, 65536 + 8, (1 << 20) + 24, 3 * (1 << 20)]
    bufs = [be.alloc(n, torch.float32, dev) for n in sizes]
    errors = torch.zeros((), dtype=torch.int64, device=dev)
    total = float(sum(range(1, world + 1)))
    count = 0
    for it in range(260):
        buf = bufs[it % len(bufs)]
        base = float(it % 7 + 1)                       # small integers: exact in fp32 and in bf16, in any order
        kind = it % 5
        buf.fill_(base * (rank + 1))
        if kind == 0:
            be.all_reduce_avg(buf, scale=1.0)
            exp_val = base * total
        elif kind == 1:
            be.all_reduce_avg(buf, scale=1.0, wire=True)
            exp_val = base * total
        elif kind == 2:
            src = it % world
            be.broadcast(buf, src=src)
            exp_val = base * (src + 1)
        elif kind == 3:
            be.reduce_scatter(buf, scale=1.0)
            be.all_gather(buf)
            exp_val = base * total
            count += 1
        else:
            if buf.numel() * 4 <= (4 << 20):
                be.all_reduce_oneshot(buf, scale=1.0)
                exp_val = base * total
            else:
                be.barrier()
                exp_val = base * (rank + 1)
        count += 1
        exp = torch.full_like(buf, exp_val)               # (expected values are known on the host: no read-back here)
        errors += (buf != exp).any().to(torch.int64)
        # a second, different collective on the same buffer right behind it (write-after-read across ranks)
        summed = base * total if (kind == 4 and buf.numel() * 4 > (4 << 20)) else exp_val * world
        be.all_reduce_avg(buf, scale=1.0 / total)
        errors += ((buf - summed / total).abs() > 1e-3 * abs(summed)).any().to(torch.int64)
        count += 1
        for _ in range(2):                             # and two tiny ones: latency-bound, flags flip fastest
            bufs[0].fill_(float(rank + 1))
            be.all_reduce_avg(bufs[0], scale=1.0)
            errors += (bufs[0] != total).any().to(torch.int64)
            count += 1
    torch.cuda.synchronize()
    out = {"errors": int(errors.item()), "collectives": count, "mode": be.mode, "stats": dict(be.stats)}
    be.close()
    return out


@pytest.mark.parametrize("mode", ["nvls", "ipc"])
def test_p2p_stress_mixed_collectives(mode):
    _need(2)
    for res in run_ranks(_stress, 2, (mode,), backend="nccl", timeout=400):
        assert res["collectives"] >= 1000
        assert res["errors"] == 0, res


def _dead_peer(rank, world):
    """Rank 1 never joins the collective: rank 0's kernel must give up after PTDB200_P2P_TIMEOUT_S, name the missing
    peer and fail the stream - not spin on the GPU forever."""
    import os
    import time
    os.environ["PTDB200_P2P_TIMEOUT_S"] = "3"
    os.environ["PTDB200_QUIET"] = "1"
    import torch
    from pytorchdistributed_b200.comm.p2p import P2PCollectives
    dev = torch.device("cuda", rank)
    be = P2PCollectives(None, dev)
    buf = be.alloc(1 << 16, torch.float32, dev)
    buf.fill_(1.0)
    be.all_reduce_avg(buf)                             # one healthy round first
    torch.cuda.synchronize()
    if rank == 1:
        time.sleep(12)                                 # alive, but never launches the second collective
        return {"role": "absent"}
    t0 = time.time()
    be.all_reduce_avg(buf)
    try:
        torch.cuda.synchronize()
        return {"role": "waiter", "failed": False, "seconds": time.time() - t0}
    except Exception as e:  # noqa: BLE001 - the trap surfaces as a CUDA error on the next synchronisation
        return {"role": "waiter", "failed": True, "seconds": time.time() - t0, "error": str(e)[:200]}


def test_dead_peer_is_reported_instead_of_hanging():
    _need(2)
    res = run_ranks(_dead_peer, 2, backend="nccl", timeout=120, hard_exit=True)
    waiter = res[0]
    assert waiter["failed"], waiter
    assert 2.0 < waiter["seconds"] < 11.0, waiter      # the 3 s budget, not the 12 s the peer stays away
