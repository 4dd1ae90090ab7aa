"""tools/index_check.py (the check bench.py runs on every N): accepts what the oracle builds, rejects a row in
the wrong bucket, a bucket out of order, a lost row and a torn row. CPU tensors stand in for device tensors."""
import os
import sys

import numpy as np
import pyarrow as pa

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_index_check_accepts_the_oracle_and_rejects_damage(oracle):
    import torch
    from tools import index_check as ic
    import fixtures
    cols, _ = fixtures.lineitem(5000)
    nb = 16
    perm, off = oracle.build_index(cols, 1, nb)
    ins = [torch.from_numpy(c.to_numpy().copy()) for c in cols]
    outs = [t[torch.from_numpy(perm)] for t in ins]
    r = ic.check_index(torch, oracle, outs, off, ins, 1, nb, sample=4000)
    assert r["ok"] and r["rows_all_ranks"] == len(perm) and r["sample_rows_per_rank"] == 4000, r
    # a row moved into the neighbouring bucket (offsets shifted by one)
    bad_off = off.copy()
    bad_off[3] += 1
    r = ic.check_index(torch, oracle, outs, bad_off, ins, 1, nb, sample=len(perm))
    assert not r["ok"] and not r["bucket_of_sample_by_oracle"], r
    # two rows of one bucket swapped: order broken, everything else intact
    sw = [t.clone() for t in outs]
    a = int(off[5])
    assert int(outs[0][a]) != int(outs[0][int(off[6]) - 1])
    for t in sw:
        t[[a, int(off[6]) - 1]] = t[[int(off[6]) - 1, a]]
    r = ic.check_index(torch, oracle, sw, off, ins, 1, nb, sample=len(perm))
    assert not r["ok"] and not r["sorted_in_every_bucket"] and r["rows_and_checksums_conserved"], r
    # a payload value swapped between two rows: rows torn apart
    torn = [t.clone() for t in outs]
    torn[1][[0, 1]] = torn[1][[1, 0]]
    assert int(outs[0][0]) != int(outs[0][1]) or float(outs[1][0]) != float(outs[1][1])
    r = ic.check_index(torch, oracle, torn, off, ins, 1, nb)
    assert not r["ok"] and not r["rows_and_checksums_conserved"], r
    # a rank that holds rows outside its owner range
    r = ic.check_index(torch, oracle, outs, off, ins, 1, nb, rank=0, world=2)
    assert not r["owner_ranges"], r


def test_index_check_composite_key(oracle):
    import torch
    from tools import index_check as ic
    rng = np.random.default_rng(5)
    k1 = rng.integers(0, 50, size=20000).astype(np.int32)
    k2 = rng.integers(0, 1 << 40, size=20000)
    v = rng.random(20000)
    cols = [pa.array(k1), pa.array(k2), pa.array(v)]
    perm, off = oracle.build_index(cols, 2, 64)
    ins = [torch.from_numpy(x) for x in (k1, k2, v)]
    outs = [t[torch.from_numpy(perm)] for t in ins]
    assert ic.check_index(torch, oracle, outs, off, ins, 2, 64)["ok"]
