"""Host link bandwidth as the end-to-end leg sees it: pinned H2D, pinned D2H, and both at once — on one GPU, or
on every GPU of the box at the same time (launch with torch.distributed.run: one process per GPU, LOCAL_RANK),
optionally with the process and its pinned buffers bound to the GPU's NUMA node (PCIE_BIND_NUMA=1 in the
environment; torch.distributed.run swallows unknown flags).

    python tools/microbench/pcie.py
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/microbench/pcie.py
"""
import os
import sys
import time

import torch

local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
numa = os.environ.get("PCIE_BIND_NUMA", "") == "1"
node = None
if numa:
    # the GPU's NUMA node from sysfs; bind this process's CPUs to it BEFORE the pinned buffers are allocated
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()[-12:]
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node >= 0:
            cpus = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
            ids = set()
            for part in cpus.split(","):
                a, _, b = part.partition("-")
                ids.update(range(int(a), int(b or a) + 1))
            os.sched_setaffinity(0, ids)
    except Exception as ex:  # report and carry on unbound
        node = f"unavailable ({ex})"
torch.cuda.set_device(local)
n = 150_000_000
h_in = torch.empty(n, dtype=torch.int64).pin_memory()
h_out = torch.empty(n, dtype=torch.int64).pin_memory()
h_in.fill_(1)      # first touch on the bound node
h_out.fill_(1)
d1 = torch.empty(n, dtype=torch.int64, device="cuda")
d2 = torch.empty(n, dtype=torch.int64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("gloo")


def t(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()   # all GPUs copy at the same time
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best


gb = n * 8 / 1e9
a = t(lambda: d1.copy_(h_in, non_blocking=True))
b = t(lambda: h_out.copy_(d2, non_blocking=True))


def both():
    with torch.cuda.stream(s1):
        d1.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2):
        h_out.copy_(d2, non_blocking=True)


c = t(both)
print(f"gpu {local} of {world} numa {node}: H2D {gb / a:.1f} GB/s  D2H {gb / b:.1f} GB/s  both at once: {gb / c:.1f} GB/s each way "
      f"({c * 1e3:.1f} ms for {gb:.1f} GB each)", flush=True)
