"""Diagnostic (GPU box): NUMA node and local CPUs of every visible GPU, and this process's affinity mask."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import seacurrent_b200 as scb

print("cpu_count", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)))
try:
    print("nodes", sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")))
except OSError as e:
    print("nodes: n/a", e)
for i in range(torch.cuda.device_count()):
    pr = torch.cuda.get_device_properties(i)
    bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
    try:
        node = open("/sys/bus/pci/devices/%s/numa_node" % bdf).read().strip()
    except OSError:
        node = "n/a"
    c = scb.gpu_local_cpus(i)
    print("gpu", i, bdf, "numa_node", node, "local_cpus", (min(c), max(c), len(c)) if c else None)
