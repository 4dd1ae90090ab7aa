"""Checks of a built index that hold at any size and on any number of GPUs — the invariant of the reference's
HT/index/DataFrameWriterExtensionsTest.scala:93-158 (every row in the bucket its key hashes to, rows sorted
inside a bucket, the union of the buckets is the input) restated for device-resident tables:

  * this rank holds rows only in the bucket range it owns (owner(b) = b * N / numBuckets) and its
    bucket_offsets are consistent with its row count;
  * the key tuple is non-decreasing inside every bucket (checked on the device, every row);
  * the bucket of a sample of rows (>= 1 M per rank by default), recomputed by the ORACLE on the host from
    the keys found in the result, is the bucket the row sits in;
  * the all-reduced row count and two order-independent row checksums equal the input's, so nothing was
    lost, duplicated or torn apart on the way (over all ranks).

Used by bench.py (outside every timed region) and tools/bench_configs.py; the oracle is only the checker."""
import numpy as np

_MIX = (1000003, 7, 31, 131, 8191, 524287)


def _as_i64(torch, t):
    if t.dtype == torch.float64:
        return t.view(torch.int64)
    if t.dtype == torch.float32:
        return t.view(torch.int32).to(torch.int64)
    return t.to(torch.int64)


def row_checksums(torch, cols):
    """Two sums over rows of a mix of all columns (wrapping int64): equal for any row order."""
    mix = None
    for i, c in enumerate(cols):
        v = _as_i64(torch, c) * _MIX[i % len(_MIX)]
        mix = v if mix is None else mix + v
    return int(mix.sum().item()), int((mix ^ (mix >> 11)).sum().item())


def check_index(torch, oracle, out_cols, boff, in_cols, n_keys, num_buckets, rank=0, world=1, dist=None,
                sample=1 << 20, seed=1):
    """out_cols / in_cols: lists of device tensors (this rank's result / this rank's input), key columns first.
    boff: this rank's bucket offsets (numpy int64, num_buckets + 1). Returns a dict with ok and the evidence."""
    import pyarrow as pa
    dev = out_cols[0].device
    boff = np.asarray(boff, dtype=np.int64)
    n_out = int(out_cols[0].numel())
    # contiguous owner ranges: owner(b) = b * world // num_buckets  <=>  ceil(r*nb/world) <= b < ceil((r+1)*nb/world)
    lo = -(-rank * num_buckets // world)
    hi = -(-(rank + 1) * num_buckets // world)
    cnt = np.diff(boff)
    res = {"rows_out_this_rank": n_out}
    owner_ok = bool(boff[0] == 0 and boff[-1] == n_out and (cnt >= 0).all() and
                    cnt[:lo].sum() == 0 and cnt[hi:].sum() == 0)
    # sorted inside every bucket (lexicographic over the key tuple)
    sorted_ok = True
    if n_out > 1:
        less = None   # row i+1 < row i
        eq = None
        for k in range(n_keys):
            a, b = out_cols[k][:-1], out_cols[k][1:]
            lt_k, eq_k = b < a, b == a
            less = lt_k if less is None else (less | (eq & lt_k))
            eq = eq_k if eq is None else (eq & eq_k)
        desc = less.nonzero().flatten() + 1
        starts = torch.as_tensor(boff[1:-1], device=dev)
        sorted_ok = bool(torch.isin(desc, starts).all().item())
        del less, eq, desc
    # bucket membership of a sample, recomputed by the oracle on the host
    m = min(sample, n_out)
    member_ok = True
    if m > 0:
        g = torch.Generator(device=dev)
        g.manual_seed(seed + rank)
        pos = torch.randint(0, n_out, (m,), generator=g, device=dev) if m < n_out else torch.arange(n_out, device=dev)
        keys = [pa.array(out_cols[k][pos].cpu().numpy()) for k in range(n_keys)]
        got = oracle.bucket_ids(keys, num_buckets)
        want = np.searchsorted(boff, pos.cpu().numpy(), side="right") - 1
        member_ok = bool(np.array_equal(got.astype(np.int64), want))
    # conservation over all ranks
    sums_in = (int(in_cols[0].numel()),) + row_checksums(torch, in_cols)
    sums_out = (n_out,) + row_checksums(torch, out_cols)
    if dist is not None and world > 1:
        # 64-bit wrapping sums: reduce the halves separately so that the all-reduce itself cannot overflow
        def red(vals):
            t = torch.tensor([[(v >> 32) & 0xFFFFFFFF, v & 0xFFFFFFFF] for v in vals], dtype=torch.int64, device=dev)
            dist.all_reduce(t)
            return tuple(int((((int(h) << 32) + int(l)) & 0xFFFFFFFFFFFFFFFF)) for h, l in t.tolist())
        sums_in, sums_out = red(sums_in), red(sums_out)
        flags = torch.tensor([owner_ok, sorted_ok, member_ok], dtype=torch.int64, device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        owner_ok, sorted_ok, member_ok = (bool(x) for x in flags.tolist())
    else:
        sums_in = tuple(v & 0xFFFFFFFFFFFFFFFF for v in sums_in)
        sums_out = tuple(v & 0xFFFFFFFFFFFFFFFF for v in sums_out)
    conserved = sums_in == sums_out
    res.update({"ok": bool(owner_ok and sorted_ok and member_ok and conserved), "owner_ranges": owner_ok,
                "sorted_in_every_bucket": sorted_ok, "bucket_of_sample_by_oracle": member_ok,
                "sample_rows_per_rank": m, "rows_and_checksums_conserved": conserved, "rows_all_ranks": sums_out[0]})
    return res


def table_tensors(torch, table, dtypes, device):
    """The columns of a DEVICE hs_table as torch tensors (no copy). dtypes: one torch dtype per column."""
    out = []
    n = table.nrows
    for i, dt in enumerate(dtypes):
        col = table.column(i)
        ts = {torch.int64: "<i8", torch.float64: "<f8", torch.int32: "<i4", torch.float32: "<f4", torch.int16: "<i2"}[dt]

        class _Arr:
            pass
        a = _Arr()
        a.__cuda_array_interface__ = {"shape": (n,), "typestr": ts, "data": (col.data, False), "version": 2}
        out.append(torch.as_tensor(a, device=device))
    return out
