"""CPU / NUMA placement of a rank next to its GPU.

The reference leaves process placement to the OS (reference resnet/main.py:74-79: no ``set_device``, no affinity).
With one process per GPU on a two-socket host that puts half of the ranks' pinned staging buffers on the wrong
socket: every host->device batch copy (154 MB per step and GPU for ImageNet-shaped fp32 input) then crosses the
socket interconnect, and at 8 GPUs the copies, not the GPUs, set the end-to-end step time.  ``pin_to_gpu_numa_node``
restricts the calling process to the CPUs of the NUMA node its GPU hangs off (from sysfs; no libnuma needed), so
first-touch allocation - including ``cudaHostAlloc`` pinned memory and DataLoader workers forked later - lands on
that node.
"""
import os


def _read(path):
    try:
        with open(path) as fh:
            return fh.read().strip()
    except OSError:
        return None


def _parse_cpulist(text):
    cpus = set()
    for part in (text or "").split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_numa_node(index):
    """NUMA node of CUDA device ``index`` (None when the platform does not say)."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(index).pci_bus_id
        dom = getattr(torch.cuda.get_device_properties(index), "pci_domain_id", 0)
        dev = torch.cuda.get_device_properties(index).pci_device_id
        addr = "%04x:%02x:%02x.0" % (dom, bus, dev)
    except Exception:  # noqa: BLE001
        return None
    node = _read("/sys/bus/pci/devices/%s/numa_node" % addr)
    if node is None:
        return None
    try:
        node = int(node)
    except ValueError:
        return None
    return node if node >= 0 else None


LAST = {"node": None, "cpus": None}   # what the most recent pin_to_gpu_numa_node() call found / did (diagnostics)


def pin_to_gpu_numa_node(index, verbose=False):
    """Restrict this process to the CPUs of the GPU's NUMA node. Returns the CPU set used (None = left alone:
    single-node host, sysfs not readable, or the node's CPUs are outside the current cgroup's allowance)."""
    if not hasattr(os, "sched_setaffinity"):
        return None
    node = gpu_numa_node(index)
    LAST["node"] = node
    if node is None:
        return None
    cpus = _parse_cpulist(_read("/sys/devices/system/node/node%d/cpulist" % node))
    try:
        allowed = os.sched_getaffinity(0)
    except OSError:
        return None
    cpus &= allowed
    if not cpus or cpus == allowed:
        return None
    try:
        os.sched_setaffinity(0, cpus)
    except OSError:
        return None
    LAST["cpus"] = len(cpus)
    if verbose:
        print("[b200ddp] cuda:%d -> NUMA node %d, %d CPUs" % (index, node, len(cpus)), flush=True)
    return cpus
