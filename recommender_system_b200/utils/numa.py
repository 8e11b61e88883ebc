"""Host-side placement for the feeder (utils/pipeline.py): run the feeding thread, and therefore first-touch its pinned
buffers, on the NUMA node the GPU hangs off.  With one process per GPU on a two-socket host, every rank otherwise allocates
its pinned staging memory on node 0 and half of the H2D traffic crosses the socket link (round 1's 8-GPU end-to-end line
scaled at 0.52 of linear with all ranks on node 0).  No reference counterpart.  Linux only; a no-op where sysfs does not
answer (containers without the PCI tree, single-node hosts)."""
from __future__ import annotations

import os


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_numa_node(device_index):
    """-> (node or None, set of local cpus or None) for CUDA device `device_index` of this process."""
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        bus = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        base = "/sys/bus/pci/devices/" + bus
        node = int(open(base + "/numa_node").read())
        cpus = _parse_cpulist(open(base + "/local_cpulist").read())
        return (node if node >= 0 else None), (cpus or None)
    except Exception:
        return None, None


class bound_to_gpu:
    """``with bound_to_gpu(local_rank): allocate pinned buffers, feed`` -- CPU affinity of this thread narrowed to the GPU's
    local cpus for the block (restored after).  ``.info`` says what was done."""

    def __init__(self, device_index, enable=True):
        self.idx, self.enable, self.prev, self.info = device_index, enable, None, {"bound": False}

    def __enter__(self):
        if self.enable and hasattr(os, "sched_setaffinity"):
            node, cpus = gpu_numa_node(self.idx)
            allowed = os.sched_getaffinity(0)
            use = (cpus & allowed) if cpus else None
            if node is not None and use and use != allowed:
                self.prev = allowed
                os.sched_setaffinity(0, use)
                self.info = {"bound": True, "numa_node": node, "cpus": len(use)}
            else:
                self.info = {"bound": False, "numa_node": node, "cpus": len(allowed)}
        return self

    def __exit__(self, *exc):
        if self.prev is not None:
            os.sched_setaffinity(0, self.prev)
        return False
