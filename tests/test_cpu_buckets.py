"""Native bucket planner / reducer bookkeeping versus torch's DDP internals (reference N3)."""
import ctypes

import pytest
import torch
import torch.distributed as dist

from pytorchdistributed_b200 import _lib
from pytorchdistributed_b200.parallel import plan_buckets


def _torch_assign(sizes_bytes, first, cap):
    tensors = [torch.empty(s // 4) for s in sizes_bytes]
    res = dist._compute_bucket_assignment_by_size(tensors, [first, cap])
    buckets = res[0] if isinstance(res, tuple) else res
    out = [0] * len(tensors)
    for b, idxs in enumerate(buckets):
        for i in idxs:
            out[i] = b
    return out


@pytest.mark.parametrize("sizes", [
    [84], [4] * 10, [400_000, 700_000, 100, 30_000_000, 100], [2 << 20, 2 << 20, 30 << 20, 4, 4, 4],
    [p.numel() * 4 for p in reversed(list(__import__("pytorchdistributed_b200.models", fromlist=["resnet50"]).resnet50().parameters()))],
])
def test_plan_matches_torch(sizes):
    first, cap = 1 << 20, 25 << 20
    assert plan_buckets(sizes, first, cap) == _torch_assign(sizes, first, cap)


def test_reducer_tracks_buckets():
    bucket_of = [0, 0, 1, 1, 1, 2]
    arr = (ctypes.c_int * 6)(*bucket_of)
    rid = _lib.call_ret("ptd_reducer_create", ctypes.cast(arr, ctypes.c_void_p), 6, 3)
    assert rid > 0
    mark = lambda i: _lib.call_ret("ptd_reducer_mark_ready", rid, i)  # noqa: E731
    assert mark(5) == 2            # single-parameter bucket completes at once
    assert mark(2) == -1 and mark(3) == -1 and mark(2) == -2 and mark(4) == 1
    assert _lib.call_ret("ptd_reducer_all_done", rid) == 0
    assert mark(0) == -1 and mark(1) == 0
    assert _lib.call_ret("ptd_reducer_all_done", rid) == 1
    out = (ctypes.c_int * 6)()
    assert _lib.call_ret("ptd_reducer_ready_order", rid, ctypes.cast(out, ctypes.c_void_p), 6) == 6
    assert list(out) == [5, 2, 3, 4, 0, 1]
    _lib.call_ret("ptd_reducer_reset", rid)
    assert _lib.call_ret("ptd_reducer_all_done", rid) == 0 and mark(5) == 2
    assert _lib.call_ret("ptd_reducer_destroy", rid) == 0
    assert _lib.call_ret("ptd_reducer_mark_ready", rid, 0) == -3
