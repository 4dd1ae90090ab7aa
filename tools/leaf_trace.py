"""Reads a HSGPU_LEAF_TRACE file (SM clock at the phase boundaries of every leaf segment of one k_leaf_sort32
launch) and prints the mean cycles per phase, overall and by segment size."""
import sys

import numpy as np

a = np.fromfile(sys.argv[1], dtype=np.int64).reshape(-1, 16)
a = a[a[:, 8] > 0]
names = ["start -> key slice landed", "keys -> packed words -> pass A (atomic) done", "pass B (stable) done",
         "tie repair / flags", "column 0 written", "column 1 written", "column 2 written", "end"]
d = np.stack([a[:, 1] - a[:, 0], a[:, 2] - a[:, 1], a[:, 3] - a[:, 2], a[:, 4] - a[:, 3], a[:, 5] - a[:, 4],
              a[:, 6] - a[:, 5], a[:, 8] - a[:, 6]], axis=1)
tot = a[:, 8] - a[:, 0]
print(f"segments {len(a)}, mean rows {a[:, 9].mean():.0f}, mean cycles per segment {tot.mean():.0f}")
lab = ["wait for the key slice", "pack + pass A (atomic rank)", "pass B (stable rank)", "flags + wait col 0 + write col 0",
       "write col 1", "write col 2", "tail (last barrier)"]
for i, l in enumerate(lab):
    print(f"  {l:36s} {d[:, i].mean():9.0f} cycles  {100 * d[:, i].mean() / tot.mean():5.1f} %")
