"""Probe what the box exposes for NVLS: driver multicast attribute, fabric handles, torch symmetric memory's
multicast pointer, and whether NCCL itself picks NVLS.  Run under torchrun with >= 2 ranks."""
import os
import sys

import torch
import torch.distributed as dist


def main() -> None:
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from cuda.bindings import driver as cu
    cu.cuInit(0)
    err, dev = cu.cuDeviceGet(local)
    for name in ("CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED", "CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED",
                 "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED",
                 "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED"):
        attr = getattr(cu.CUdevice_attribute, name, None)
        if attr is None:
            print(f"[{rank}] {name}: not in bindings")
            continue
        err, val = cu.cuDeviceGetAttribute(attr, dev)
        print(f"[{rank}] {name} = {val} ({err})", flush=True)
    x = torch.ones(1 << 20, device="cuda")
    dist.all_reduce(x)
    torch.cuda.synchronize()
    try:
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(1 << 20, dtype=torch.float32, device=torch.device("cuda", local))
        h = symm.rendezvous(t, dist.group.WORLD.group_name)
        print(f"[{rank}] symm multicast_ptr = {h.multicast_ptr:#x} buffer_ptrs = {[hex(p) for p in h.buffer_ptrs]}", flush=True)
    except Exception as e:  # noqa: BLE001
        print(f"[{rank}] torch symmetric memory failed: {type(e).__name__}: {e}", flush=True)
    # pidfd_getfd availability (fd passing between ranks without a unix socket)
    import ctypes
    libc = ctypes.CDLL(None, use_errno=True)
    pidfd = libc.syscall(434, os.getppid(), 0)
    print(f"[{rank}] pidfd_open(parent) -> {pidfd} errno {ctypes.get_errno()}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
