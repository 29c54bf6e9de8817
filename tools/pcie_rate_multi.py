"""Concurrent pinned D2H bandwidth of all ranks (torchrun): what the host side of the box sustains when every GPU
copies at once.  usage: torchrun --nproc-per-node N tools/pcie_rate_multi.py [pin]   (pin: bind to the GPU's NUMA node first)"""
import os
import sys

import torch
import torch.distributed as dist

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if len(sys.argv) > 1 and sys.argv[1] == "pin":
    bus = torch.cuda.get_device_properties(local).pci_bus_id if hasattr(torch.cuda.get_device_properties(local), "pci_bus_id") else None
    import subprocess
    bid = subprocess.run(["nvidia-smi", f"--id={local}", "--query-gpu=pci.bus_id", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
    path = f"/sys/bus/pci/devices/{bid[4:].lower()}/local_cpulist"
    try:
        cl = open(path).read().strip()
        cpus = set()
        for part in cl.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        os.sched_setaffinity(0, cpus)
        print(f"rank {local}: {bid} -> cpus {cl}", flush=True)
    except Exception as e:
        print(f"rank {local}: no affinity ({e})", flush=True)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = (3 << 30) if "big" in sys.argv else (1 << 30)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h = torch.empty(n, dtype=torch.uint8).pin_memory()
h.copy_(d)
torch.cuda.synchronize()
dist.barrier()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(4):
    h.copy_(d, non_blocking=True)
e1.record()
torch.cuda.synchronize()
gbs = 4 * n / e0.elapsed_time(e1) / 1e6
t = torch.tensor([gbs], device="cuda")
dist.all_reduce(t)
print(f"rank {local}: D2H {gbs:.1f} GB/s" + (f"   sum over ranks {t.item():.1f} GB/s" if local == 0 else ""), flush=True)
dist.destroy_process_group()
