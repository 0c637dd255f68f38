"""Keep a rank's host threads - and with them its pinned staging memory - on the NUMA node its GPU hangs off.

Why: every step's batch crosses PCIe from pinned host memory (``Trainer.run_batches``).  Pinned pages are placed on the
node of the thread that allocates them; with eight ranks scheduled wherever the kernel likes, about half of them end up
reading their batch across the socket interconnect.  The builder's 8-GPU runs show it in the one copy per run that is
not hidden behind compute, the first: 154 MB took ~13 ms with 8 ranks against ~3 ms with one
(``profiles/bench_n{1,8}_r2*.json``: e2e minus device-timed step, times the step count).

The reference leaves placement to the launcher (``torchrun`` sets ``OMP_NUM_THREADS`` only,
``R/02 DDP基本概念/ddp_gpus_torchrun.py:101-102``); here ``ddp_setup`` calls ``bind_to_gpu_node`` (``PTDB200_NUMA_BIND=0``
switches it off).  Everything is read from sysfs; any surprise (no sysfs, no NUMA information, a cpuset that excludes
the node) leaves the affinity as it was.
"""
from __future__ import annotations

import os
from typing import Optional, Set


def parse_cpulist(text: str) -> Set[int]:
    """``"0-3,8,10-11"`` -> {0, 1, 2, 3, 8, 10, 11}."""
    out: Set[int] = set()
    for part in text.strip().split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            lo, hi = part.split("-", 1)
            out.update(range(int(lo), int(hi) + 1))
        else:
            out.add(int(part))
    return out


def pci_address(domain: int, bus: int, device: int) -> str:
    return f"{domain:04x}:{bus:02x}:{device:02x}.0"


def gpu_numa_cpus(pci_addr: str, sysfs: str = "/sys") -> Optional[Set[int]]:
    """CPUs of the NUMA node the PCI device ``pci_addr`` is attached to, or ``None`` when sysfs does not say."""
    try:
        with open(os.path.join(sysfs, "bus", "pci", "devices", pci_addr, "numa_node")) as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open(os.path.join(sysfs, "devices", "system", "node", f"node{node}", "cpulist")) as f:
            cpus = parse_cpulist(f.read())
        return cpus or None
    except (OSError, ValueError):
        return None


def choose_affinity(node_cpus: Optional[Set[int]], allowed: Set[int], min_cpus: int = 4) -> Optional[Set[int]]:
    """The CPUs to bind to: the node's CPUs this process may use - or ``None`` (leave things alone) when that is
    nothing new (already inside the node) or too little to run a rank's threads on."""
    if not node_cpus:
        return None
    inter = node_cpus & allowed
    if len(inter) < min(min_cpus, len(allowed)) or inter == allowed:
        return None
    return inter


def bind_to_gpu_node(device_index: int, sysfs: str = "/sys") -> Optional[Set[int]]:
    """Bind the calling thread (threads started later inherit it) to the CPUs next to CUDA device ``device_index``.
    Returns the CPU set that was applied, or ``None`` when nothing was changed."""
    if os.environ.get("PTDB200_NUMA_BIND", "1") == "0" or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        import torch
        props = torch.cuda.get_device_properties(device_index)
        addr = pci_address(int(props.pci_domain_id), int(props.pci_bus_id), int(props.pci_device_id))
        target = choose_affinity(gpu_numa_cpus(addr, sysfs), set(os.sched_getaffinity(0)))
        if target is None:
            return None
        os.sched_setaffinity(0, target)
        return target
    except Exception:  # noqa: BLE001 - placement is an optimisation, never a reason to fail
        return None
