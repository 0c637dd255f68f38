"""Run a function in `world` processes with an initialised gloo/nccl process group (test helper)."""
import os
import socket
import sys
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fn, args, ret, backend, hard_exit=False):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank),
                      WORLD_SIZE=str(world))
    try:
        if backend == "nccl":
            torch.cuda.set_device(rank)
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        else:
            dist.init_process_group("gloo", rank=rank, world_size=world)
        out = fn(rank, world, *args)
        ret[rank] = ("ok", out)
        if hard_exit:              # the function killed its CUDA context on purpose: no orderly teardown is possible
            os._exit(0)
    except Exception:  # noqa: BLE001
        ret[rank] = ("err", traceback.format_exc())
    finally:
        try:
            from pytorchdistributed_b200.comm import destroy_process_group
            destroy_process_group()
        except Exception:  # noqa: BLE001
            pass


def run_ranks(fn, world=2, args=(), backend="gloo", timeout=240, hard_exit=False):
    ctx = mp.get_context("spawn")
    mgr = ctx.Manager()
    ret = mgr.dict()
    port = _free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, ret, backend, hard_exit)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
    for p in procs:
        if p.is_alive():
            p.kill()
            raise RuntimeError("rank process timed out")
    out = []
    for r in range(world):
        status, val = ret.get(r, ("err", "no result (process died)"))
        if status != "ok":
            raise RuntimeError(f"rank {r} failed:\n{val}")
        out.append(val)
    return out
