"""Process-group bootstrap (control plane).

Two entry points mirror the reference's two launch styles:

* ``ddp_setup(rank, world_size)`` - spawn style, ``R/02 DDP基本概念/ddp_gpus.py:11-23``:
  sets ``MASTER_ADDR``/``MASTER_PORT`` itself, then initialises the group and
  selects ``cuda:rank``.
* ``ddp_setup()`` - torchrun style, ``R/02 DDP基本概念/ddp_gpus_torchrun.py:11-19``:
  everything comes from the environment (``RANK``, ``WORLD_SIZE``, ``MASTER_*``,
  ``LOCAL_RANK``).

The torch process group is only the *control plane* here (rendezvous, handle
exchange for the peer-to-peer runtime, pipeline send/recv).  Gradient traffic goes
through ``pytorchdistributed_b200.comm.p2p`` on a GPU box.  On a box without CUDA the
backend falls back to gloo so the same scripts run as CPU plumbing tests.
"""
from __future__ import annotations

import datetime
import os
from typing import Optional

import torch
import torch.distributed as dist

from ..utils.numa import bind_to_gpu_node

DEFAULT_ADDR = "127.0.0.1"      # the reference hard-codes "localhost" (ddp_gpus.py:18); same host, resolvable everywhere
DEFAULT_PORT = "12355"          # ddp_gpus.py:19


def pick_backend(requested: Optional[str] = None) -> str:
    if requested in (None, "auto"):
        requested = os.environ.get("PTDB200_BACKEND", "auto")
    if requested == "auto":
        return "nccl" if torch.cuda.is_available() else "gloo"
    if requested == "nccl" and not torch.cuda.is_available():
        return "gloo"
    return requested


def local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0")))


def ddp_setup(rank: Optional[int] = None, world_size: Optional[int] = None, backend: Optional[str] = None,
              timeout_s: int = 600) -> int:
    """Initialise the default process group; returns the device index to use.

    With ``rank``/``world_size`` given it behaves like the spawn variant, otherwise
    like the torchrun variant.
    """
    be = pick_backend(backend)
    kw = dict(backend=be, timeout=datetime.timedelta(seconds=timeout_s))
    if rank is not None:
        os.environ.setdefault("MASTER_ADDR", DEFAULT_ADDR)
        os.environ.setdefault("MASTER_PORT", DEFAULT_PORT)
        dev = rank
        if torch.cuda.is_available():
            torch.cuda.set_device(dev)
            kw["device_id"] = torch.device("cuda", dev)
            bind_to_gpu_node(dev)          # host threads and pinned staging memory next to this rank's GPU
        dist.init_process_group(rank=rank, world_size=world_size, **kw)
    else:
        dev = local_rank()
        if torch.cuda.is_available():
            torch.cuda.set_device(dev)
            kw["device_id"] = torch.device("cuda", dev)
            bind_to_gpu_node(dev)
        dist.init_process_group(**kw)
    return dev


def destroy_process_group() -> None:
    from . import p2p
    p2p.shutdown()
    if dist.is_initialized():
        dist.destroy_process_group()


def is_main() -> bool:
    return (not dist.is_initialized()) or dist.get_rank() == 0


def world() -> tuple[int, int]:
    if dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1
