"""Binds the calling process to the CPU cores (and, through first-touch, the memory) of the NUMA node its GPU hangs
off, read from sysfs. One process per GPU launched by torchrun inherits no affinity: on a two-socket host half of
the ranks would otherwise stage their pinned host buffers on the far socket and every host<->device copy would cross
the inter-socket link. Used by bench.py's end-to-end leg and tools/h2d_bandwidth.py."""
import os


def gpu_numa_node(device_index):
    import torch

    p = torch.cuda.get_device_properties(device_index)
    bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    try:
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        cpus = open(f"/sys/bus/pci/devices/{bdf}/local_cpulist").read().strip()
    except OSError:
        return None, None, bdf
    return node, cpus, bdf


def _parse_cpulist(text):
    out = set()
    for part in text.split(","):
        if "-" in part:
            a, b = part.split("-")
            out.update(range(int(a), int(b) + 1))
        elif part:
            out.add(int(part))
    return out


def bind_to_gpu(device_index):
    """-> dict describing what was done (for the bench line)."""
    node, cpus, bdf = gpu_numa_node(device_index)
    info = {"gpu": device_index, "pci": bdf, "numa_node": node, "bound": False}
    if node is None or node < 0 or not cpus:
        return info
    try:
        allowed = os.sched_getaffinity(0)
        want = _parse_cpulist(cpus) & allowed
        if want:
            os.sched_setaffinity(0, want)
            info.update(bound=True, cpus=len(want))
    except OSError as e:
        info["error"] = str(e)
    return info
