"""Process-per-stage pipeline: schedule shapes and gradient equivalence with single-process training (gloo)."""
import pytest
import torch

from pytorchdistributed_b200.parallel import max_live_activations, schedule_ops
from tests._mp import run_ranks


@pytest.mark.parametrize("n_stages,n_micro", [(2, 6), (4, 8), (4, 3), (3, 1)])
def test_schedules_are_complete_and_ordered(n_stages, n_micro):
    for sched in ("gpipe", "1f1b"):
        for st in range(n_stages):
            ops = schedule_ops(sched, st, n_stages, n_micro)
            assert sorted(ops) == sorted([("F", i) for i in range(n_micro)] + [("B", i) for i in range(n_micro)])
            for i in range(n_micro):                       # a micro-batch is never backpropagated before its forward
                assert ops.index(("F", i)) < ops.index(("B", i))
            fs = [i for k, i in ops if k == "F"]
            bs = [i for k, i in ops if k == "B"]
            assert fs == sorted(fs) and bs == sorted(bs)
    # 1F1B bounds the activations alive per stage by the pipeline depth, GPipe by the number of micro-batches
    assert max_live_activations(schedule_ops("gpipe", 0, 4, 8)) == 8
    assert max_live_activations(schedule_ops("1f1b", 0, 4, 8)) == 4
    assert max_live_activations(schedule_ops("1f1b", 3, 4, 8)) == 1


def _stage_run(rank, world, schedule, recompute):
    import torch
    import torch.nn.functional as F
    from pytorchdistributed_b200.parallel import PipelineStage
    torch.manual_seed(0)
    full = torch.nn.Sequential(torch.nn.Linear(10, 16), torch.nn.ReLU(), torch.nn.Linear(16, 12), torch.nn.ReLU(),
                               torch.nn.Linear(12, 4))
    parts = [full[:2], full[2:4], full[4:]][:world] if world == 3 else [full[:2], full[2:]]
    stage = PipelineStage(parts[rank], rank, world, device=torch.device("cpu"), loss_fn=F.mse_loss, schedule=schedule,
                          recompute=recompute)
    g = torch.Generator().manual_seed(1)
    x, t = torch.randn(10, 10, generator=g), torch.randn(10, 4, generator=g)
    loss = stage.train_step(x if rank == 0 else None, t if rank == world - 1 else None, split_size=3, batch_rows=10)
    grads = [p.grad.clone() for p in parts[rank].parameters()]
    return (float(loss) if loss is not None else None), grads


@pytest.mark.parametrize("world,schedule,recompute", [(2, "gpipe", False), (2, "1f1b", False), (3, "1f1b", True),
                                                      (2, "gpipe", True)])
def test_pipeline_stages_match_single_process(world, schedule, recompute):
    import torch.nn.functional as F
    res = run_ranks(_stage_run, world, (schedule, recompute))
    torch.manual_seed(0)
    full = torch.nn.Sequential(torch.nn.Linear(10, 16), torch.nn.ReLU(), torch.nn.Linear(16, 12), torch.nn.ReLU(),
                               torch.nn.Linear(12, 4))
    g = torch.Generator().manual_seed(1)
    x, t = torch.randn(10, 10, generator=g), torch.randn(10, 4, generator=g)
    # micro-batches of 3,3,3,1 rows weighted by their share of the batch = mean over elements per micro-batch, weighted
    sizes = [3, 3, 3, 1]
    loss = sum(F.mse_loss(full(x[s:s + n]), t[s:s + n]) * (n / 10) for s, n in zip([0, 3, 6, 9], sizes))
    loss.backward()
    assert abs(res[-1][0] - float(loss)) < 1e-5
    got = [g_ for r in res for g_ in r[1]]
    for a, p in zip(got, full.parameters()):
        assert torch.allclose(a, p.grad, atol=1e-5)
