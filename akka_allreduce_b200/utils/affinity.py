"""CPU/NUMA placement for a rank process: run on the cores next to the GPU it drives.

The end-to-end path copies every round's input from pinned host memory to the device.  On a two-socket host
half of the GPUs hang off each socket; a rank whose pinned buffers were first touched on the far socket pays the
inter-socket link on every H2D copy (and the coordinator/worker threads pay it on every flag they poll).  Binding
the process to the GPU's NUMA node before it allocates anything keeps buffers, threads and PCIe root on one side.
The reference leaves placement to the JVM and the OS; NCCL does the same thing for its own threads.

Everything here is best effort: on a single-node host, in a container that hides sysfs, or with
``AAR_NO_AFFINITY`` set, nothing happens.
"""
from __future__ import annotations

import os


MIN_LOCAL_CPUS = 4


def parse_cpulist(text: str) -> set[int]:
    """'0-3,8,10-11' -> {0,1,2,3,8,10,11} (the format of /sys/devices/system/node/nodeN/cpulist)."""
    cpus: set[int] = set()
    for part in text.strip().split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            lo, hi = part.split("-", 1)
            cpus.update(range(int(lo), int(hi) + 1))
        else:
            cpus.add(int(part))
    return cpus


def numa_node_of_pci(address: str, sysfs: str = "/sys") -> int | None:
    """NUMA node of a PCI function ('0000:19:00.0'), None when the platform does not say (-1 or no file)."""
    try:
        node = int(open(os.path.join(sysfs, "bus/pci/devices", address.lower(), "numa_node")).read().strip())
    except (OSError, ValueError):
        return None
    return node if node >= 0 else None


def cpus_of_node(node: int, sysfs: str = "/sys") -> set[int]:
    try:
        return parse_cpulist(open(os.path.join(sysfs, f"devices/system/node/node{node}/cpulist")).read())
    except (OSError, ValueError):
        return set()


def pci_address(device_index: int) -> str | None:
    import torch

    try:
        p = torch.cuda.get_device_properties(device_index)
        return f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except Exception:
        return None


def bind_to_gpu(device_index: int, sysfs: str = "/sys") -> dict | None:
    """Restricts the calling thread (and every thread it starts later) to the cores of the GPU's NUMA node.
    Returns what was done ({'numa_node': n, 'cpus': count}) or None when there was nothing to do."""
    if os.environ.get("AAR_NO_AFFINITY") is not None or not hasattr(os, "sched_setaffinity"):
        return None
    addr = pci_address(device_index)
    node = numa_node_of_pci(addr, sysfs) if addr else None
    if node is None:
        return None
    try:
        allowed = os.sched_getaffinity(0)
        local = cpus_of_node(node, sysfs) & allowed
        if len(local) < MIN_LOCAL_CPUS or local == allowed:
            return None  # nothing to gain, or too few local cores for the worker, IO and sampler threads
        os.sched_setaffinity(0, local)
    except OSError:
        return None
    return {"numa_node": node, "cpus": len(local)}
