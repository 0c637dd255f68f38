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
