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


def test_cfg2_build_and_cfg3_join_properties(hs):
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
    # (3) every row is in the bucket its key hashes to
    bid = torch.empty(n, dtype=torch.int32, device=dev)
    kcs = abi.DeviceColumnSet([{"type": abi.HS_INT64, "length": n, "data": t.column(0).data}])
    st = ctx._lib.hs_hash_bucket(ctx._h, kcs.ptr(), C.c_int32(1), C.c_int32(200), C.c_void_p(bid.data_ptr()))
    assert st == 0
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
