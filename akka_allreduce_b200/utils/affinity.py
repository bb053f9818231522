"""CPU/NUMA placement for a rank process: run on the cores next to the GPU it drives.

The end-to-end path copies every round's input from pinned host memory to the device.  On a two-socket host
half of the GPUs hang off each socket; a rank whose pinned buffers were first touched on the far socket pays the
inter-socket link on every H2D copy (and the coordinator/worker threads pay it on every flag they poll).  Binding
the process to the GPU's NUMA node before it allocates anything keeps buffers, threads and PCIe root on one side.
The reference leaves placement to the JVM and the OS; NCCL does the same thing for its own threads.

Everything here is best effort: the GPU's NUMA node comes from sysfs, or — in a container that hides it — from the
CPU affinity NVML reports for the device; on a single-node host, or with ``AAR_NO_AFFINITY`` set, nothing is bound,
and the returned record says why.
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


def nvml_cpus_of_gpu(device_index: int) -> set[int]:
    """Cores NVML calls ideal for this GPU (what ``nvidia-smi topo -m`` prints as CPU affinity).  Works where the
    container hides the PCI device's sysfs node; empty when NVML cannot say."""
    try:
        import pynvml as nv
        import torch

        nv.nvmlInit()
        try:
            p = torch.cuda.get_device_properties(device_index)
            bus = f"{p.pci_domain_id:08x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
            handle = nv.nvmlDeviceGetHandleByPciBusId(bus.encode())
            words = (os.cpu_count() + 63) // 64
            mask = nv.nvmlDeviceGetCpuAffinity(handle, words)
            return {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        finally:
            nv.nvmlShutdown()
    except Exception:
        return set()


def bind_to_gpu(device_index: int, sysfs: str = "/sys") -> dict | None:
    """Restricts the calling thread (and every thread it starts later) to the cores next to the GPU: the cores of
    its NUMA node (sysfs), else the CPU affinity NVML reports for it.  Always says what happened:
    ``{'bound': True, 'numa_node': n | None, 'cpus': count, 'source': 'sysfs' | 'nvml'}`` or
    ``{'bound': False, 'reason': ...}``; None only when AAR_NO_AFFINITY is set or the platform has no affinity calls."""
    if os.environ.get("AAR_NO_AFFINITY") is not None or not hasattr(os, "sched_setaffinity"):
        return None
    addr = pci_address(device_index)
    node = numa_node_of_pci(addr, sysfs) if addr else None
    try:
        allowed = os.sched_getaffinity(0)
    except OSError as e:
        return {"bound": False, "reason": f"sched_getaffinity: {e}"}
    local, source = set(), None
    if node is not None:
        local, source = cpus_of_node(node, sysfs) & allowed, "sysfs"
    if not local:
        local, source = nvml_cpus_of_gpu(device_index) & allowed, "nvml"
    if not local:
        return {"bound": False, "reason": "neither sysfs nor NVML names the cores next to this GPU",
                "allowed_cpus": len(allowed)}
    if local == allowed:
        return {"bound": False, "reason": "the process is already confined to the GPU's cores", "numa_node": node,
                "cpus": len(local), "source": source}
    if len(local) < MIN_LOCAL_CPUS:  # too few local cores for the worker, IO and sampler threads
        return {"bound": False, "reason": f"only {len(local)} local cores allowed", "numa_node": node, "source": source}
    try:
        os.sched_setaffinity(0, local)
    except OSError as e:
        return {"bound": False, "reason": f"sched_setaffinity: {e}"}
    return {"bound": True, "numa_node": node, "cpus": len(local), "source": source}
