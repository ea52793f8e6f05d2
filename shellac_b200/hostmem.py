"""Host-memory placement for the end-to-end path.

The e2e number is bound by the host->device copy of the request bytes, and on
a multi-socket box a pinned arena that lives on the OTHER socket's memory crosses the inter-socket
link before it reaches the GPU's PCIe root.  `bind_near_gpu(device)` pins the calling process to the
CPUs of the NUMA node the GPU hangs off, BEFORE the pinned arenas are allocated (pages are placed
on the node of the thread that first touches them), one process per GPU as everywhere else here.

Pure sysfs + sched_setaffinity: no libnuma, no NVML.  Every step that cannot be done (no sysfs
entry, node -1 inside a VM, empty cpulist) returns None and leaves the process alone.
"""
import ctypes as C
import os

from . import _native as N


def _read(path):
    try:
        with open(path) as f:
            return f.read().strip()
    except OSError:
        return None


def parse_cpulist(text):
    """'0-3,8,10-11' -> [0, 1, 2, 3, 8, 10, 11]"""
    cpus = []
    for part in (text or '').split(','):
        part = part.strip()
        if not part:
            continue
        if '-' in part:
            a, b = part.split('-', 1)
            cpus.extend(range(int(a), int(b) + 1))
        else:
            cpus.append(int(part))
    return cpus


def gpu_pci_address(device):
    buf = C.create_string_buffer(32)
    if N.lib().shl_device_pci_bus_id(device, buf, 32) != 0:
        return None
    return buf.value.decode().lower()


def gpu_numa_node(device, sysfs='/sys'):
    addr = gpu_pci_address(device)
    if not addr:
        return None
    text = _read(os.path.join(sysfs, 'bus/pci/devices', addr, 'numa_node'))
    try:
        node = int(text)
    except (TypeError, ValueError):
        return None
    return node if node >= 0 else None


def node_cpus(node, sysfs='/sys'):
    return parse_cpulist(_read(os.path.join(sysfs, 'devices/system/node/node%d/cpulist' % node)))


def bind_near_gpu(device, sysfs='/sys'):
    """Restrict this process to the CPUs next to `device`.  Returns (node, n_cpus) or None."""
    node = gpu_numa_node(device, sysfs)
    if node is None:
        return None
    allowed = os.sched_getaffinity(0)
    cpus = [c for c in node_cpus(node, sysfs) if c in allowed]
    if not cpus:
        return None
    os.sched_setaffinity(0, cpus)
    return node, len(cpus)
