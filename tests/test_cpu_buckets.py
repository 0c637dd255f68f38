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


def test_native_reducer_rejects_bad_bucket_recipes():
    """ptd_reducer_bind_bucket validates its recipe before it touches CUDA (this box has no GPU)."""
    import ctypes
    from pytorchdistributed_b200 import _lib
    from pytorchdistributed_b200.parallel import ddp  # noqa: F401  (registers the reducer entry points)
    arr = (ctypes.c_int * 3)(0, 0, 1)
    rid = _lib.call_ret("ptd_reducer_create", ctypes.cast(arr, ctypes.c_void_p), 3, 2)
    assert rid > 0
    c = _lib.BucketComm()
    c.algo, c.rank, c.world = 7, 0, 2                      # unknown algorithm
    assert _lib.call_ret("ptd_reducer_bind_bucket", rid, 0, ctypes.addressof(c)) != 0
    c.algo, c.world = 2, 1                                 # a one-rank group has nothing to reduce
    assert _lib.call_ret("ptd_reducer_bind_bucket", rid, 0, ctypes.addressof(c)) != 0
    c.algo, c.world, c.fuse_sgd = 3, 2, 1                  # the fused SGD step needs fp32 on the wire
    assert _lib.call_ret("ptd_reducer_bind_bucket", rid, 0, ctypes.addressof(c)) != 0
    assert b"fp32" in _lib.require().ptd_last_error()
    # an unbound reducer still counts: marking through the launching entry point behaves like the plain one
    assert _lib.call_ret("ptd_reducer_mark_ready_launch", rid, 2, None, None) == 1      # bucket 1 complete (one parameter)
    assert _lib.call_ret("ptd_reducer_mark_ready_launch", rid, 0, None, None) == -1
    assert _lib.call_ret("ptd_reducer_mark_ready_launch", rid, 0, None, None) == -2     # marked twice
    assert _lib.call_ret("ptd_reducer_mark_ready_launch", rid, 1, None, None) == 0
    assert _lib.call_ret("ptd_reducer_finalize", rid, None) == 1
    _lib.call_ret("ptd_reducer_destroy", rid)


def test_baseline_tables_match_the_committed_profiles():
    """BASELINE.md section B is generated from profiles/*.json: the text between the markers must be what the generator
    prints today (nobody typed a number by hand, and nobody forgot to regenerate)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "benchmarks", "make_baseline_tables.py")],
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    text = open(os.path.join(root, "BASELINE.md")).read()
    begin, end = "<!-- BEGIN GENERATED TABLES -->", "<!-- END GENERATED TABLES -->"
    inside = text[text.index(begin) + len(begin):text.index(end)]
    assert inside.strip() == out.stdout.strip()
    assert "REJECTED" not in inside                          # no quoted row ran under a thermal / hardware slowdown
