"""Host-memory placement for ranks that stage activations in pinned memory.

One process per GPU (vllm --distributed-executor-backend mp, README.md:33).  On a two-socket
B200 box four GPUs hang off each socket; a rank whose pinned buffers were first touched on the
OTHER socket pays the inter-socket link on every H2D / D2H, and all eight ranks then share
that link.  `bind_to_gpu_numa_node` pins the calling process to the CPUs of its GPU's NUMA
node BEFORE anything is allocated, so first-touch places the pinned arena (and the CUDA
context's own host allocations) on the right node.  Pure sysfs + sched_setaffinity: no
libnuma, nothing to install.  Returns a small dict for the bench log, never raises.
"""
from __future__ import annotations

import os
import subprocess
from typing import Dict, List, Optional


def _parse_cpulist(s: str) -> List[int]:
    cpus: List[int] = []
    for part in s.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.extend(range(int(a), int(b) + 1))
        else:
            cpus.append(int(part))
    return cpus


def gpu_pci_bus_id(index: int) -> Optional[str]:
    """PCI bus id of CUDA device `index` without creating a CUDA context."""
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=index,pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=20).stdout
    except (OSError, subprocess.TimeoutExpired):
        return None
    rows = {}
    for line in out.splitlines():
        parts = [p.strip() for p in line.split(",")]
        if len(parts) == 2 and parts[0].isdigit():
            rows[int(parts[0])] = parts[1]
    phys = index
    if vis:
        ids = [v.strip() for v in vis.split(",") if v.strip()]
        if index < len(ids) and ids[index].isdigit():
            phys = int(ids[index])
    return rows.get(phys)


def numa_node_of_pci(bus_id: str) -> Optional[int]:
    # nvidia-smi prints 00000000:1B:00.0, sysfs wants 0000:1b:00.0
    dom, rest = bus_id.split(":", 1)
    path = f"/sys/bus/pci/devices/{dom[-4:].lower()}:{rest.lower()}/numa_node"
    try:
        with open(path) as f:
            n = int(f.read().strip())
        return n if n >= 0 else None
    except (OSError, ValueError):
        return None


def bind_to_gpu_numa_node(device_index: int) -> Dict[str, object]:
    info: Dict[str, object] = {"bound": False}
    bus = gpu_pci_bus_id(device_index)
    if bus is None:
        info["why"] = "no PCI bus id (nvidia-smi unavailable)"
        return info
    node = numa_node_of_pci(bus)
    info["pci"] = bus
    if node is None:
        info["why"] = "sysfs reports no NUMA node for the GPU (single-node box or virtualised)"
        return info
    try:
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = _parse_cpulist(f.read())
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            info["why"] = f"no allowed CPU on NUMA node {node}"
            return info
        os.sched_setaffinity(0, cpus)
    except (OSError, ValueError) as e:
        info["why"] = f"sched_setaffinity failed: {e}"
        return info
    info.update(bound=True, node=node, cpus=len(cpus))
    return info
