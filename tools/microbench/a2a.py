"""What the NVSwitch fabric gives an all-to-all of peer copies, nothing else running: one process, every GPU of the
box copies `mb` MiB to each of the other GPUs at the same time (cudaMemcpyPeerAsync through torch, one stream per
(source, destination) pair, or one stream per source), plus the same with kernels doing the stores
(torch's copy kernel is not used: the copy engines are what the grouped-span exchange uses).

    python tools/microbench/a2a.py [mb]
"""
import sys
import time

import torch

mb = int(sys.argv[1]) if len(sys.argv) > 1 else 256
n = torch.cuda.device_count()
words = mb * (1 << 20) // 8
src = [[None] * n for _ in range(n)]
dst = [[None] * n for _ in range(n)]
streams = [[None] * n for _ in range(n)]
one = []
for i in range(n):
    with torch.cuda.device(i):
        one.append(torch.cuda.Stream())
        for j in range(n):
            if i != j:
                src[i][j] = torch.ones(words, dtype=torch.int64, device=f"cuda:{i}")
                streams[i][j] = torch.cuda.Stream()
for j in range(n):
    with torch.cuda.device(j):
        for i in range(n):
            if i != j:
                dst[i][j] = torch.empty(words, dtype=torch.int64, device=f"cuda:{j}")


def sync():
    for i in range(n):
        torch.cuda.synchronize(i)


def run(per_pair_streams, peers):
    """peers(i) -> destinations of source i, in issue order."""
    best = 1e9
    for _ in range(4):
        sync()
        t0 = time.perf_counter()
        for i in range(n):
            with torch.cuda.device(i):
                for j in peers(i):
                    with torch.cuda.stream(streams[i][j] if per_pair_streams else one[i]):
                        dst[i][j].copy_(src[i][j], non_blocking=True)
        sync()
        best = min(best, time.perf_counter() - t0)
    return best


gb = words * 8 / 1e9
if n < 2:
    print("one GPU: nothing to measure")
    sys.exit(0)
for name, pps in (("one stream per (source, destination)", True), ("one stream per source, owners staggered", False)):
    t = run(pps, lambda i: [(i + k) % n for k in range(1, n)])
    print(f"{n} GPUs, {mb} MiB per pair, {name}: {t * 1e3:.2f} ms -> {gb * (n - 1) / t:.0f} GB/s out of (and into) every GPU", flush=True)
# a ring only (every GPU sends to its right neighbour): the per-link figure
t = run(True, lambda i: [(i + 1) % n])
print(f"{n} GPUs, ring (one destination each): {t * 1e3:.2f} ms -> {gb / t:.0f} GB/s per GPU", flush=True)
