"""The allocation-free BatchNorm-backward entry point, driven through the C ABI like a foreign host would.

Who needs it: code that is handed its inputs and outputs and may not grow the process's device memory - a plug-in
living inside another runtime's pre-planned arena, a service with a fixed footprint.  Such a host cannot supply the
2*C-float scratch that ``ptd_bn_bwd_reduce`` / ``ptd_bn_bwd_apply`` expect, so ``ptd_bn_backward_static`` keeps it in an
array that ships inside the library (at most 2048 channels).

The test calls the symbol with raw pointers via ctypes and compares with the regular two-call path, twice in a row on
different data: the second call only comes out right if the first call's leftovers in the shared array were cleared.
This file sorts last on purpose, so a run that stops at the first failure has already been through everything else.
"""
import ctypes

import pytest
import torch

from pytorchdistributed_b200 import _lib

pytestmark = pytest.mark.gpu
CL = torch.channels_last
P, L, I = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _case(c, relu, seed, dev):
    g = torch.Generator().manual_seed(seed)
    n, h = 6, 10
    mk = lambda s: (torch.randn(n, c, h, h, generator=g) * s).to(dev).to(torch.bfloat16).contiguous(memory_format=CL)  # noqa: E731
    y, dz = mk(2.0), mk(1.0)
    z = torch.relu(mk(1.0)) if relu else mk(1.0)
    mean = (torch.randn(c, generator=g) * 0.1).to(dev)
    invstd = (torch.rand(c, generator=g) + 0.5).to(dev)
    gamma = (torch.rand(c, generator=g) + 0.5).to(dev)
    return y, z, dz, mean, invstd, gamma, n * h * h


def _reference_path(lib, y, z, dz, mean, invstd, gamma, m, c, relu, stream):
    sums = torch.zeros(2, c, device=y.device)                    # the scratch a normal caller allocates and zeroes
    dy, gout = torch.empty_like(y), torch.empty_like(y)
    assert lib.ptd_bn_bwd_reduce(dz.data_ptr(), z.data_ptr(), y.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                 sums[0].data_ptr(), sums[1].data_ptr(), m, c, relu, stream) == 0
    assert lib.ptd_bn_bwd_apply(dz.data_ptr(), z.data_ptr(), y.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                gamma.data_ptr(), sums[0].data_ptr(), sums[1].data_ptr(), dy.data_ptr(),
                                gout.data_ptr(), m, c, relu, stream) == 0
    return dy, gout, sums[1].clone(), sums[0].clone()            # dy, masked grad, dgamma, dbeta


def test_static_scratch_entry_point_matches_regular_path_twice():
    lib = _lib.require()
    fn = lib.ptd_bn_backward_static
    fn.restype = I
    fn.argtypes = [P, P, P, P, P, P, P, P, P, P, L, I, I, P]
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream(dev).cuda_stream
    # outputs are allocated up front, like a host that owns only its inputs and outputs would
    for call, (c, relu, seed) in enumerate([(256, 1, 11), (2048, 0, 12), (64, 1, 13)]):
        y, z, dz, mean, invstd, gamma, m = _case(c, relu, seed, dev)
        dy, gout = torch.empty_like(y), torch.empty_like(y)
        dgamma, dbeta = torch.empty(c, device=dev), torch.empty(c, device=dev)
        rc = fn(dz.data_ptr(), z.data_ptr(), y.data_ptr(), mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(),
                dy.data_ptr(), gout.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), m, c, relu, stream)
        assert rc == 0, lib.ptd_last_error()
        rdy, rgout, rdgamma, rdbeta = _reference_path(lib, y, z, dz, mean, invstd, gamma, m, c, relu, stream)
        torch.cuda.synchronize()
        assert _rel(dy, rdy) < 1e-2, f"call {call}"
        assert _rel(gout, rgout) < 1e-2, f"call {call}"
        assert _rel(dgamma, rdgamma) < 1e-2 and _rel(dbeta, rdbeta) < 1e-2, f"call {call}"


def _torch_fp32_bn_backward(y, z, dz, mean, invstd, gamma, relu):
    """The same backward written in plain fp32 PyTorch: the oracle that does not share a line with the library."""
    yf, dzf = y.float(), dz.float()
    g = dzf * (z.float() > 0) if relu else dzf
    cshape = (1, -1, 1, 1)
    xhat = (yf - mean.view(cshape)) * invstd.view(cshape)
    dbeta = g.sum((0, 2, 3))
    dgamma = (g * xhat).sum((0, 2, 3))
    m = yf.numel() // yf.shape[1]
    dy = (gamma * invstd).view(cshape) * (g - dbeta.view(cshape) / m - xhat * dgamma.view(cshape) / m)
    return dy, g, dgamma, dbeta


@pytest.mark.parametrize("dev_index", [0, 1])
def test_static_scratch_entry_point_matches_fp32_pytorch(dev_index):
    """Against plain PyTorch, and once on a second GPU: the array's address is looked up per device, so the first
    call there takes the branch that the calls on device 0 no longer do."""
    if torch.cuda.device_count() <= dev_index:
        pytest.skip(f"needs {dev_index + 1} GPUs")
    lib = _lib.require()
    fn = lib.ptd_bn_backward_static
    fn.restype = I
    fn.argtypes = [P, P, P, P, P, P, P, P, P, P, L, I, I, P]
    dev = torch.device("cuda", dev_index)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        for c, relu, seed in [(128, 1, 21), (512, 0, 22)]:           # two calls in a row on this device
            y, z, dz, mean, invstd, gamma, m = _case(c, relu, seed, dev)
            dy, gout = torch.empty_like(y), torch.empty_like(y)
            dgamma, dbeta = torch.empty(c, device=dev), torch.empty(c, device=dev)
            rc = fn(dz.data_ptr(), z.data_ptr(), y.data_ptr(), mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(),
                    dy.data_ptr(), gout.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), m, c, relu, stream)
            assert rc == 0, lib.ptd_last_error()
            tdy, tg, tdgamma, tdbeta = _torch_fp32_bn_backward(y, z, dz, mean, invstd, gamma, bool(relu))
            torch.cuda.synchronize(dev)
            assert _rel(dy, tdy) < 1e-2 and _rel(gout, tg) < 1e-2, (c, relu)
            assert _rel(dgamma, tdgamma) < 1e-2 and _rel(dbeta, tdbeta) < 1e-2, (c, relu)


def test_static_scratch_entry_point_rejects_bad_arguments():
    """Nothing may launch on arguments the kernels cannot honour: the bit-mask mode of the two-call path, a missing z
    with the ReLU on, an empty tensor, a channel count that is not a power of two."""
    lib = _lib.require()
    fn = lib.ptd_bn_backward_static
    fn.restype = I
    fn.argtypes = [P, P, P, P, P, P, P, P, P, P, L, I, I, P]
    dev = torch.device("cuda:0")
    c = 64
    t = torch.zeros(1, c, 2, 2, device=dev, dtype=torch.bfloat16).contiguous(memory_format=CL)
    v = torch.ones(c, device=dev)
    out = torch.empty(c, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    call = lambda z, m, ch, relu: fn(t.data_ptr(), z, t.data_ptr(), v.data_ptr(), v.data_ptr(), v.data_ptr(),  # noqa: E731
                                     t.data_ptr(), None, out.data_ptr(), out.data_ptr(), m, ch, relu, st)
    assert call(t.data_ptr(), 4, c, 2) != 0 and b"relu" in lib.ptd_last_error()
    assert call(None, 4, c, 1) != 0 and b"z" in lib.ptd_last_error()
    assert call(t.data_ptr(), 0, c, 0) != 0
    assert call(t.data_ptr(), 4, 96, 0) != 0 and b"power of two" in lib.ptd_last_error()
    assert call(t.data_ptr(), 4, c, 1) == 0, lib.ptd_last_error()      # and the same arguments, valid, go through
    torch.cuda.synchronize()


def test_static_scratch_entry_point_rejects_oversize():
    lib = _lib.require()
    fn = lib.ptd_bn_backward_static
    fn.restype = I
    fn.argtypes = [P, P, P, P, P, P, P, P, P, P, L, I, I, P]
    dev = torch.device("cuda:0")
    c = 4096                                                       # one step past the built-in 2048-channel scratch
    t = torch.zeros(1, c, 2, 2, device=dev, dtype=torch.bfloat16).contiguous(memory_format=CL)
    v = torch.ones(c, device=dev)
    out = torch.empty(c, device=dev)
    rc = fn(t.data_ptr(), t.data_ptr(), t.data_ptr(), v.data_ptr(), v.data_ptr(), v.data_ptr(), t.data_ptr(), None,
            out.data_ptr(), out.data_ptr(), 4, c, 0, torch.cuda.current_stream(dev).cuda_stream)
    assert rc != 0
    assert b"2048" in lib.ptd_last_error()
