"""Full-size checks (BASELINE.json configs[1], 100 M rows) through size-independent properties:
column checksums are preserved, every bucket is sorted, every row sits in the bucket its key hashes to,
and the lineitem x orders join returns exactly one order per lineitem row with matching keys."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
N_ROWS = int(os.environ.get("HS_FULLSIZE_ROWS", "100000000"))


def _dev_tensor(torch, col, dtype, n):
    """wrap a DEVICE hs_column buffer as a torch tensor (no copy)"""
    class _Arr:
        pass
    a = _Arr()
    itemsize = torch.empty(0, dtype=dtype).element_size()
    a.__cuda_array_interface__ = {"shape": (n,), "typestr": {torch.int64: "<i8", torch.float64: "<f8",
                                  torch.int32: "<i4"}[dtype], "data": (col.data, False), "version": 2}
    return torch.as_tensor(a, device="cuda")


def test_cfg2_build_and_cfg3_join_properties(hs, oracle):
    import torch
    import bench
    from hyperspace_1_b200 import abi
    dev = torch.device("cuda", 0)
    n = N_ROWS
    key, price, disc, okeys = bench.gen_lineitem_device(torch, n, 20240601, dev)
    torch.cuda.synchronize()
    ctx = hs.Context(0)
    ctx.set_output_location(abi.HS_MEM_DEVICE)
    cs = abi.DeviceColumnSet([{"type": abi.HS_INT64, "length": n, "data": key.data_ptr()},
                              {"type": abi.HS_FLOAT64, "length": n, "data": price.data_ptr()},
                              {"type": abi.HS_FLOAT64, "length": n, "data": disc.data_ptr()}])
    t = ctx.build_index_table(cs, 1, 200)
    assert t.nrows == n
    boff = t.bucket_offsets()
    assert boff[0] == 0 and boff[-1] == n and (np.diff(boff) >= 0).all()
    k = _dev_tensor(torch, t.column(0), torch.int64, n)
    p = _dev_tensor(torch, t.column(1), torch.float64, n)
    d = _dev_tensor(torch, t.column(2), torch.float64, n)
    # (1) checksums: nothing lost, nothing duplicated, rows stay intact
    assert int(k.sum()) == int(key.sum())
    assert int((k ^ (k >> 7)).sum()) == int((key ^ (key >> 7)).sum())
    assert int(p.view(torch.int64).sum()) == int(price.view(torch.int64).sum())
    row_in = (key * 1000003 + price.view(torch.int64) * 7 + disc.view(torch.int64)).sum()
    row_out = (k * 1000003 + p.view(torch.int64) * 7 + d.view(torch.int64)).sum()
    assert int(row_in) == int(row_out), "rows were torn apart"
    # (2) sorted inside every bucket
    desc = (k[1:] < k[:-1]).nonzero().flatten() + 1
    starts = torch.as_tensor(boff[1:-1], device=dev)
    assert bool(torch.isin(desc, starts).all()), "a bucket is not sorted"
    # (3) every row is in the bucket its key hashes to: bucket ids recomputed by the ORACLE on the host over the
    # whole key column of the result (the invariant of HT/index/DataFrameWriterExtensionsTest.scala:136-145)
    import pyarrow as pa
    bid = torch.from_numpy(oracle.bucket_ids([pa.array(k.cpu().numpy())], 200)).to(dev)
    want = torch.repeat_interleave(torch.arange(200, dtype=torch.int32, device=dev),
                                   torch.as_tensor(np.diff(boff), device=dev))
    assert bool((bid == want).all()), "a row is in the wrong bucket"
    del bid, want, desc

    # cfg3: orders index + join
    n_o = int(okeys.numel())
    g = torch.Generator(device=dev)
    g.manual_seed(20240602)
    ok = okeys[torch.randperm(n_o, generator=g, device=dev)].contiguous()
    tot = (ok % 1000).to(torch.float64) + 0.25   # o_totalprice is a function of the key: checkable
    torch.cuda.synchronize()   # the library runs on its own stream: torch's writes must have landed
    ocs = abi.DeviceColumnSet([{"type": abi.HS_INT64, "length": n_o, "data": ok.data_ptr()},
                               {"type": abi.HS_FLOAT64, "length": n_o, "data": tot.data_ptr()}])
    ot = ctx.build_index_table(ocs, 1, 200)

    def view(tbl, ncols):
        arr = (abi.hs_column * ncols)(*[tbl.column(i) for i in range(ncols)])
        dcs = abi.DeviceColumnSet([])
        dcs.c, dcs.n = arr, ncols
        return hs.IndexView(dcs, 1, tbl.bucket_offsets())
    jt = ctx.join_table(view(t, 3), view(ot, 2), [0, 1, 2], [1])
    assert jt.nrows == n, "every lineitem row matches exactly one order"
    jk = _dev_tensor(torch, jt.column(0), torch.int64, n)
    jp = _dev_tensor(torch, jt.column(1), torch.float64, n)
    jtot = _dev_tensor(torch, jt.column(3), torch.float64, n)
    assert int(jk.sum()) == int(key.sum())
    assert int(jp.view(torch.int64).sum()) == int(price.view(torch.int64).sum())
    assert bool((jtot == (jk % 1000).to(torch.float64) + 0.25).all()), "joined a row with the wrong order"
    assert np.array_equal(jt.bucket_offsets(), boff)
    jt.free()
    ot.free()
    t.free()
    ctx.close()


def test_cfg5_shape_composite_key_properties(hs, oracle):
    """cfg5 per-GPU shape at 50 M rows: (k1:int32 < 2^20, k2:int64 < 2^40, v:float64), 8192 buckets —
    lexicographic order inside every bucket, bucket membership, checksums."""
    import torch
    from hyperspace_1_b200 import abi
    dev = torch.device("cuda", 0)
    n = int(os.environ.get("HS_CFG5_ROWS", "50000000"))
    g = torch.Generator(device=dev)
    g.manual_seed(20240604)
    k1 = torch.randint(0, 1 << 20, (n,), generator=g, device=dev, dtype=torch.int32)
    k2 = torch.randint(0, 1 << 40, (n,), generator=g, device=dev, dtype=torch.int64)
    v = torch.rand(n, generator=g, device=dev, dtype=torch.float64)
    torch.cuda.synchronize()
    ctx = hs.Context(0)
    ctx.set_output_location(abi.HS_MEM_DEVICE)
    cs = abi.DeviceColumnSet([{"type": abi.HS_INT32, "length": n, "data": k1.data_ptr()},
                              {"type": abi.HS_INT64, "length": n, "data": k2.data_ptr()},
                              {"type": abi.HS_FLOAT64, "length": n, "data": v.data_ptr()}])
    nb = 8192   # cfg5's bucket count: more than 2048 buckets take the two-level bucket scatter
    t = ctx.build_index_table(cs, 2, nb)
    boff = t.bucket_offsets()
    assert t.nrows == n and boff[-1] == n
    o1 = _dev_tensor(torch, t.column(0), torch.int32, n)
    o2 = _dev_tensor(torch, t.column(1), torch.int64, n)
    ov = _dev_tensor(torch, t.column(2), torch.float64, n)
    mix_in = (k1.to(torch.int64) * 1000003 + k2 * 7 + v.view(torch.int64)).sum()
    mix_out = (o1.to(torch.int64) * 1000003 + o2 * 7 + ov.view(torch.int64)).sum()
    assert int(mix_in) == int(mix_out), "rows were torn apart or lost"
    bad = ((o1[1:] < o1[:-1]) | ((o1[1:] == o1[:-1]) & (o2[1:] < o2[:-1]))).nonzero().flatten() + 1
    starts = torch.as_tensor(boff[1:-1], device=dev)
    assert bool(torch.isin(bad, starts).all()), "a bucket is not sorted on (k1, k2)"
    import pyarrow as pa
    bid = torch.from_numpy(oracle.bucket_ids([pa.array(o1.cpu().numpy()), pa.array(o2.cpu().numpy())], nb)).to(dev)
    want = torch.repeat_interleave(torch.arange(nb, dtype=torch.int32, device=dev), torch.as_tensor(np.diff(boff), device=dev))
    assert bool((bid == want).all()), "a row is in the wrong bucket"
    t.free()
    ctx.close()


def test_cfg4_shape_string_keys_properties(hs, oracle):
    """cfg4 shape at 10 M rows (host buffers): variable-length keys from a 64-symbol alphabet, 10 % sharing one of
    1000 16-byte prefixes, 1000 buckets. Every output row is checked: bucket membership (oracle hash), order
    inside the bucket, and (key, row id) pairing against the input."""
    import pyarrow as pa
    import pyarrow.compute as pc
    m = int(os.environ.get("HS_CFG4_ROWS", "10000000"))
    rng = np.random.default_rng(20240603)
    lens = np.clip(8 + rng.geometric(1 / 17, size=m), 1, 96).astype(np.int32)
    off = np.zeros(m + 1, dtype=np.int32)
    np.cumsum(lens, out=off[1:])
    alphabet = np.frombuffer(b"abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789-_", dtype=np.uint8)
    data = alphabet[rng.integers(0, 64, size=int(off[-1]))]
    pref = alphabet[rng.integers(0, 64, size=(1000, 16))]
    sel = np.nonzero((rng.random(m) < 0.10) & (lens >= 16))[0]
    pi = rng.integers(0, 1000, size=len(sel))
    for j in range(16):
        data[off[sel] + j] = pref[pi, j]
    s = pa.Array.from_buffers(pa.string(), m, [None, pa.py_buffer(off.tobytes()), pa.py_buffer(data.tobytes())])
    rid = pa.array(np.arange(m, dtype=np.int64))
    ctx = hs.Context(0)
    nb = 1000
    got, boff = ctx.build_index([s, rid], 1, nb)
    ctx.close()
    assert len(got[0]) == m and boff[-1] == m
    # pairing: the key that travelled with row id r is the input key of row r; every row id appears once
    r = got[1].to_numpy()
    assert np.array_equal(np.sort(r), np.arange(m))
    assert got[0].equals(s.take(pa.array(r)))
    # membership: the oracle's bucket of every output key is the bucket the row sits in
    b = oracle.bucket_ids([got[0]], nb)
    assert np.array_equal(b, np.repeat(np.arange(nb, dtype=np.int32), np.diff(boff)))
    # order: inside a bucket keys never decrease (byte-wise)
    a0, a1 = got[0].slice(0, m - 1).cast(pa.binary()), got[0].slice(1).cast(pa.binary())
    dec = np.nonzero(pc.greater(a0, a1).to_numpy(zero_copy_only=False))[0] + 1
    assert np.isin(dec, boff[1:-1]).all(), "a bucket is not sorted"


def test_skewed_key_distributions(hs):
    """20 M rows per case (device-resident): constant keys, a handful of distinct keys (whole buckets of equal
    keys: the copy path of the segmented sort), power-law keys (several MSD levels on few buckets' worth of key
    range), sorted input, two far-apart clusters, the full signed range. Every case: buckets sorted, every key
    still paired with its row id, every row id present once."""
    import torch
    from hyperspace_1_b200 import abi
    dev = torch.device("cuda", 0)
    n = int(os.environ.get("HS_SKEW_ROWS", "20000000"))
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    ctx = hs.Context(0)
    ctx.set_output_location(abi.HS_MEM_DEVICE)
    u = lambda hi: torch.randint(0, hi, (n,), generator=g, device=dev)
    cases = {
        "all equal": torch.full((n,), 7, device=dev, dtype=torch.int64),
        "10 distinct": u(10),
        "1000 distinct": u(1000),
        "power law": (torch.rand(n, generator=g, device=dev).pow(8) * 1e9).to(torch.int64),
        "sorted input": torch.arange(n, device=dev, dtype=torch.int64),
        "two clusters": torch.where(torch.rand(n, generator=g, device=dev) < 0.5, u(1000000), (1 << 62) + u(1000000)),
        "full signed range": torch.randint(-(1 << 62), 1 << 62, (n,), generator=g, device=dev),
    }
    rid = torch.arange(n, device=dev, dtype=torch.int64)
    for name, key in cases.items():
        torch.cuda.synchronize()
        cs = abi.DeviceColumnSet([{"type": abi.HS_INT64, "length": n, "data": key.data_ptr()},
                                  {"type": abi.HS_INT64, "length": n, "data": rid.data_ptr()}])
        t = ctx.build_index_table(cs, 1, 200)
        boff = t.bucket_offsets()
        ok = _dev_tensor(torch, t.column(0), torch.int64, n)
        ov = _dev_tensor(torch, t.column(1), torch.int64, n)
        bad = (ok[1:] < ok[:-1]).nonzero().flatten() + 1
        assert bool(torch.isin(bad, torch.as_tensor(boff[1:-1], device=dev)).all()), f"{name}: a bucket is not sorted"
        assert bool((key[ov] == ok).all()), f"{name}: a key lost its row"
        assert int(ov.sum()) == n * (n - 1) // 2 and int(ov.max()) == n - 1, f"{name}: row ids are not a permutation"
        t.free()
    ctx.close()
