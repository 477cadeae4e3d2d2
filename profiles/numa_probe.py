#!/usr/bin/env python3
"""Where is the GPU attached, where does this process run, and what does pinned-memory H2D bandwidth look like
when the pinned buffer is allocated from each NUMA node's CPUs?"""
import glob
import os
import time

import torch

torch.cuda.init()
prop = torch.cuda.get_device_properties(0)
bdf = "%04x:%02x:%02x.0" % (prop.pci_domain_id, prop.pci_bus_id, prop.pci_device_id)
node = open("/sys/bus/pci/devices/%s/numa_node" % bdf).read().strip() if os.path.exists(
    "/sys/bus/pci/devices/%s/numa_node" % bdf) else "?"
print("gpu0", bdf, "numa_node", node, "| allowed cpus", len(os.sched_getaffinity(0)))
nodes = sorted(glob.glob("/sys/devices/system/node/node[0-9]*"))


def cpus_of(path):
    out = set()
    for part in open(path + "/cpulist").read().strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        out.update(range(int(a), int(b or a) + 1))
    return out


allowed = os.sched_getaffinity(0)
dev = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
for n in nodes:
    cpus = cpus_of(n) & allowed
    if not cpus:
        print(os.path.basename(n), "no allowed cpus")
        continue
    os.sched_setaffinity(0, cpus)
    h = torch.empty(1 << 30, dtype=torch.uint8, pin_memory=True)
    h.fill_(1)
    for _ in range(2):
        dev.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        dev.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    print(os.path.basename(n), "cpus", len(cpus), "H2D %.1f GB/s" % (5 * (1 << 30) / (time.perf_counter() - t0) / 1e9))
    del h
    os.sched_setaffinity(0, allowed)
