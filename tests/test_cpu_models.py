"""Model definitions: torchvision state_dict parity, parameter count, CPU forward equivalence."""
import pytest
import torch

from pytorchdistributed_b200.models import MLP, ToyLinear, resnet50
from pytorchdistributed_b200.optim import SGD, Adam
from pytorchdistributed_b200.ops import functional as K


def test_resnet50_matches_torchvision_layout():
    torchvision = pytest.importorskip("torchvision")
    m = resnet50()
    tv = torchvision.models.resnet50()
    assert sum(p.numel() for p in m.parameters()) == 25_557_032
    sd, tsd = m.state_dict(), tv.state_dict()
    assert list(sd.keys()) == list(tsd.keys())
    assert all(sd[k].shape == tsd[k].shape and sd[k].dtype == tsd[k].dtype for k in sd)
    tv.load_state_dict(sd)          # our checkpoint loads into the reference's model class
    m2 = resnet50()
    m2.load_state_dict(tv.state_dict())   # and back
    x = torch.randn(2, 3, 64, 64)
    m.train(); tv.train()
    assert torch.allclose(m(x), tv(x), atol=1e-3)
    m.eval(); tv.eval()
    assert torch.allclose(m(x), tv(x), atol=1e-3)


def test_resnet_stage_split_shares_parameters():
    m = resnet50()
    seq1, seq2, fc = m.stage_modules()
    ids = {id(p) for p in m.parameters()}
    assert {id(p) for mod in (seq1, seq2, fc) for p in mod.parameters()} == ids
    n1 = sum(p.numel() for p in seq1.parameters())
    assert n1 == 1_444_928            # seq1 parameter count of the reference split
    x = torch.randn(2, 3, 64, 64)
    m.eval()
    y = fc(seq2(seq1(x)).view(2, -1))
    assert torch.allclose(y, m(x), atol=1e-4)


def test_toy_models_match_reference_shapes():
    lin = ToyLinear()
    ref = torch.nn.Linear(20, 1)
    ref.load_state_dict(lin.state_dict())
    x = torch.rand(32, 20)
    assert torch.allclose(lin(x), ref(x), atol=1e-6)
    mlp = MLP()
    assert list(mlp.state_dict().keys()) == [f"net.{i}.{k}" for i in (0, 2, 4, 6) for k in ("weight", "bias")]
    assert mlp(torch.randn(32, 10)).shape == (32, 5)


def test_degenerate_reference_loss_is_zero():
    out = torch.randn(32, 1, requires_grad=True)
    loss = K.cross_entropy(out, torch.rand(32, 1))
    loss.backward()
    assert float(loss.abs()) == 0.0 and float(out.grad.abs().max()) == 0.0


@pytest.mark.parametrize("ours,theirs,kw", [
    (SGD, torch.optim.SGD, dict(lr=1e-2)),
    (SGD, torch.optim.SGD, dict(lr=1e-2, momentum=0.9, weight_decay=1e-3)),
    (Adam, torch.optim.Adam, dict(lr=1e-3)),
    (Adam, torch.optim.Adam, dict(lr=1e-3, weight_decay=1e-2)),
])
def test_optimizers_match_torch_on_cpu(ours, theirs, kw):
    torch.manual_seed(0)
    a, b = MLP(), MLP()
    b.load_state_dict(a.state_dict())
    oa, ob = ours(a.parameters(), **kw), theirs(b.parameters(), **kw)
    for step in range(4):
        x, t = torch.randn(16, 10), torch.randn(16, 5)
        for m, o in ((a, oa), (b, ob)):
            o.zero_grad()
            torch.nn.functional.mse_loss(m(x), t).backward()
            o.step()
    for pa, pb in zip(a.parameters(), b.parameters()):
        assert torch.allclose(pa, pb, atol=1e-6)
