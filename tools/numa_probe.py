#!/usr/bin/env python
"""Host topology next to the GPU: NUMA nodes, the GPU's node and local CPUs, this process's affinity."""
import glob, os
import torch
p = torch.cuda.get_device_properties(0)
bus = "%04x:%02x:%02x.0" % (getattr(p, "pci_domain_id", 0), getattr(p, "pci_bus_id", 0), getattr(p, "pci_device_id", 0))
print("gpu", p.name, "pci", bus)
for f in ("numa_node", "local_cpulist"):
    try:
        print(f, open("/sys/bus/pci/devices/%s/%s" % (bus, f)).read().strip())
    except Exception as e:
        print(f, "unreadable:", e)
print("nodes online", open("/sys/devices/system/node/online").read().strip())
for n in sorted(glob.glob("/sys/devices/system/node/node*/cpulist")):
    print(n, open(n).read().strip())
print("cpus", os.cpu_count(), "affinity", sorted(os.sched_getaffinity(0))[:8], "...", len(os.sched_getaffinity(0)))
