"""NVTX ranges (no-ops without CUDA): step / bucket / micro-batch markers for nsys / ncu timelines.
The reference has no tracing at all; torch itself only marks ``DataParallel.forward`` /
``DistributedDataParallel.forward``."""
from __future__ import annotations

import contextlib

import torch


@contextlib.contextmanager
def nvtx_range(name: str):
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()
