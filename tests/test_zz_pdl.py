"""Programmatic dependent launch (csrc/common.cuh: ptd_launch / pdl_wait): the convolution and BatchNorm kernels of a
bottleneck block chained as programmatic dependents must compute what the plainly launched kernels compute."""
import pytest
import torch

from pytorchdistributed_b200 import _lib

pytestmark = pytest.mark.gpu
CL = torch.channels_last
DEV = "cuda"


def _rel(a, b):
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


@pytest.mark.parametrize("mode", [1, 2])
def test_pdl_launch_modes_compute_the_same_bottleneck(mode):
    from pytorchdistributed_b200.models.layers import BatchNorm2d, Conv2d
    from pytorchdistributed_b200.models.resnet import Bottleneck
    torch.manual_seed(0)
    down = torch.nn.Sequential(Conv2d(64, 256, 1, 1), BatchNorm2d(256))
    blk = Bottleneck(64, 64, 1, down).to(DEV).train()
    g = torch.Generator(device="cpu").manual_seed(3)
    x = torch.randn(8, 64, 32, 32, generator=g).to(DEV).to(torch.bfloat16).contiguous(memory_format=CL)
    dz = torch.randn(8, 256, 32, 32, generator=g).to(DEV).to(torch.bfloat16).contiguous(memory_format=CL)

    def run():
        blk.zero_grad()
        x1 = x.clone().requires_grad_()
        out = blk(x1)
        out.backward(dz)
        torch.cuda.synchronize()
        return [out.detach().float().clone(), x1.grad.float().clone()] + [p.grad.float().clone() for p in blk.parameters()]

    old = _lib.call_ret("ptd_set_pdl_mode", 0)
    try:
        base = run()
        assert all(torch.isfinite(t).all() for t in base)
        _lib.call_ret("ptd_set_pdl_mode", mode)
        for _ in range(3):
            got = run()
            for i, (a, b) in enumerate(zip(got, base)):
                assert _rel(a, b) < 2e-2, (mode, i)          # atomics order is the only freedom between two runs; a race is O(1)
    finally:
        _lib.call_ret("ptd_set_pdl_mode", old)
