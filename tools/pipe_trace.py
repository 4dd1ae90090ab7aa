"""Reads a HSGPU_PIPE_TRACE file (per-item timestamps of one k_sort_pipeline launch) and says where the time went:
per kind of item the count, the mean time waiting for dependencies and the mean working time; per CTA the share
of the kernel's span spent waiting, working, and between items (fetching the next ticket)."""
import sys

import numpy as np

a = np.fromfile(sys.argv[1], dtype=np.uint64)
n = int(a[0])
r = a[4:4 + 4 * n].reshape(n, 4)
kind = (r[:, 0] & 0xFF).astype(int)
seg = ((r[:, 0] >> 8) & 0xFFFFFF).astype(int)
cta = ((r[:, 0] >> 32) & 0xFFF).astype(int)
t0, t1, t2 = r[:, 1].astype(np.int64), r[:, 2].astype(np.int64), r[:, 3].astype(np.int64)
base = t0.min()
span = t2.max() - base
print(f"items {n}, span {span / 1e3:.1f} us, CTAs {len(set(cta))}")
names = {0: "scatter", 1: "hist", 2: "leaf"}
for k in (1, 0, 2):
    m = kind == k
    if not m.any():
        continue
    w, x = (t1 - t0)[m], (t2 - t1)[m]
    print(f"  {names[k]:8s} n={m.sum():7d}  wait mean {w.mean() / 1e3:7.2f} us (p99 {np.percentile(w, 99) / 1e3:7.2f}, total {w.sum() / 1e6:8.2f} ms)"
          f"  work mean {x.mean() / 1e3:7.2f} us (total {x.sum() / 1e6:8.2f} ms)")
tot_wait = (t1 - t0).sum()
tot_work = (t2 - t1).sum()
ncta = len(set(cta))
print(f"  CTA time: waiting {tot_wait / (ncta * span):.1%}, working {tot_work / (ncta * span):.1%}, "
      f"between items {1 - (tot_wait + tot_work) / (ncta * span):.1%}")
# progress over time: which segment the leaves are at, per 10 % of the span
order = np.argsort(t2)
for q in range(1, 11):
    cut = base + span * q // 10
    m = t2 <= cut
    print(f"  by {q * 10:3d} % of the span: {m.sum():7d} items done, max leaf segment {seg[m & (kind == 2)].max() if (m & (kind == 2)).any() else -1}")
