#!/usr/bin/env python
"""Side measurements for the BASELINE.json configs that are not the bench.py headline line:
cfg4 (utf8 keys, 1000 buckets) and cfg5 (composite int32+int64 key, 8192 buckets, then a 10 % append =
refresh-incremental and a join over two runs per bucket). Device-resident in -> device-resident out, CUDA events,
per-stage times from the library's own stage timers. One JSON line per config on stdout.

    python tools/bench_configs.py --config cfg4 --rows 80000000
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_configs.py --config cfg5
"""
import argparse
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))


def gen_cfg4_device(torch, m, seed, dev):
    """cfg4 rows on the device: length ~ 8 + Geometric(1/17) clipped to [1, 96], bytes from a 64-symbol alphabet,
    10 % of the keys (those of >= 16 bytes) start with one of 1000 fixed 16-byte prefixes; v = row id."""
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    lens = torch.empty(m, device=dev, dtype=torch.float32).geometric_(1.0 / 17.0, generator=g).to(torch.int64) + 8
    lens.clamp_(1, 96)
    off = torch.zeros(m + 1, dtype=torch.int64, device=dev)
    torch.cumsum(lens, 0, out=off[1:])
    total = int(off[-1])
    alphabet = torch.frombuffer(bytearray(b"abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789-_"),
                                dtype=torch.uint8).to(dev)
    data = torch.empty(total, dtype=torch.uint8, device=dev)
    step = 1 << 28
    for s in range(0, total, step):
        e = min(total, s + step)
        data[s:e] = alphabet[torch.randint(0, 64, (e - s,), generator=g, device=dev)]
    pref = alphabet[torch.randint(0, 64, (1000, 16), generator=g, device=dev)]
    sel = ((torch.rand(m, generator=g, device=dev) < 0.10) & (lens >= 16)).nonzero().flatten()
    pi = torch.randint(0, 1000, (sel.numel(),), generator=g, device=dev)
    pos = off[sel]
    for j in range(16):
        data[pos + j] = pref[pi, j]
    rid = torch.arange(m, dtype=torch.int64, device=dev)
    return off, data, rid, total


def cfg4_measure(torch, native, abi, dev, m, steps, warmup, large_offsets=False, check=True, ctx=None):
    """cfg4 on one GPU: build of m string-keyed rows (device-resident in and out). Returns the result line.
    check: bucket membership of a sample of output rows recomputed by the oracle, and byte-wise order of
    sampled neighbours inside buckets, both on the host."""
    nb = 1000
    off64, data, rid, total = gen_cfg4_device(torch, m, 20240603, dev)
    wide = total >= (1 << 31) or large_offsets
    off = off64 if wide else off64.to(torch.int32)
    torch.cuda.synchronize()
    own_ctx = ctx is None
    if own_ctx:
        ctx = native.Context(dev.index or 0)
    ctx.set_output_location(abi.HS_MEM_DEVICE)
    cs = abi.DeviceColumnSet([
        {"type": getattr(abi, "HS_LARGE_UTF8", -1) if wide else abi.HS_UTF8, "length": m, "offsets": off.data_ptr(),
         "data": data.data_ptr()},
        {"type": abi.HS_INT64, "length": m, "data": rid.data_ptr()}])
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    for _ in range(warmup):
        ctx.build_index_table(cs, 1, nb).free()
    ctx.enable_timing(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage_ms = {}
    e0.record(stream)
    for _ in range(steps):
        t = ctx.build_index_table(cs, 1, nb)
        for name, ms, launches in ctx.last_timings():
            a = stage_ms.setdefault(name, [0.0, 0])
            a[0] += ms
            a[1] += launches
        t.free()
    e1.record(stream)
    torch.cuda.synchronize()
    ctx.enable_timing(False)
    ms_step = e0.elapsed_time(e1) / steps
    row_bytes = total / m + (8 if wide else 4) + 8
    key_bytes = total / m + (8 if wide else 4)
    algo = key_bytes + 4 * row_bytes
    line = {"config": "cfg4", "rows": m, "key_bytes_total": total, "num_buckets": nb, "offsets": "int64" if wide else "int32",
            "ms_per_step": ms_step, "rows_per_s": m / (ms_step * 1e-3), "algorithmic_bytes_per_row": algo,
            "achieved_gbs": algo * m / (ms_step * 1e-3) / 1e9,
            "stages_ms": {k: round(v[0] / steps, 3) for k, v in stage_ms.items()},
            "launches_per_step": sum(v[1] for v in stage_ms.values()) / steps}
    if check:
        line["check"] = cfg4_check(torch, abi, ctx, cs, nb, m, dev, off64, data)
    if own_ctx:
        ctx.close()
    return line


def cfg4_check(torch, abi, ctx, cs, nb, m, dev, in_off, in_data, sample=200_000):
    """One more build; on a sample of output positions: the row's bucket recomputed by the oracle from the key
    bytes found in the result, the order against the next row of the same bucket (unsigned byte-wise, shorter
    first), and the key bytes against the input row the v column (= source row id) names."""
    import numpy as np
    import pyarrow as pa
    from oracle import oracle   # the checker only
    t = ctx.build_index_table(cs, 1, nb)
    torch.cuda.synchronize()
    boff = t.bucket_offsets()
    kc, vc = t.column(0), t.column(1)
    large = kc.type in (abi.HS_LARGE_UTF8, abi.HS_LARGE_BINARY)

    def dev_view(addr, n, dt, ts):
        class _A:
            pass
        a = _A()
        a.__cuda_array_interface__ = {"shape": (n,), "typestr": ts, "data": (addr, False), "version": 2}
        return torch.as_tensor(a, device=dev)
    ooff = dev_view(kc.offsets, m + 1, None, "<i8" if large else "<i4").to(torch.int64)
    nbytes = int(ooff[-1])
    odata = dev_view(kc.data, nbytes, None, "|u1")
    orid = dev_view(vc.data, m, None, "<i8")
    g = torch.Generator(device=dev)
    g.manual_seed(7)
    pos = torch.randint(0, m - 1, (min(sample, m - 1),), generator=g, device=dev)

    def strings(offs, dat, idx):
        s0, s1 = offs[idx].cpu().numpy(), offs[idx + 1].cpu().numpy()
        # gather the bytes of the sampled rows with one device gather
        lens = s1 - s0
        tot = int(lens.sum())
        starts = torch.as_tensor(np.repeat(s0 - np.concatenate([[0], np.cumsum(lens)[:-1]]), lens), device=dev)
        flat = dat[starts + torch.arange(tot, device=dev)].cpu().numpy().tobytes()
        cut = np.concatenate([[0], np.cumsum(lens)])
        return [flat[cut[i]:cut[i + 1]] for i in range(len(lens))]
    here, nxt = strings(ooff, odata, pos), strings(ooff, odata, pos + 1)
    bucket = np.searchsorted(boff, pos.cpu().numpy(), side="right") - 1
    bucket_next = np.searchsorted(boff, (pos + 1).cpu().numpy(), side="right") - 1
    got = oracle.bucket_ids([pa.array(here, pa.binary())], nb)
    member_ok = bool(np.array_equal(got.astype(np.int64), bucket))
    order_ok = all(a <= b for a, b, x, y in zip(here, nxt, bucket, bucket_next) if x == y)   # bytes compare unsigned
    src = strings(in_off, in_data, orid[pos])
    pairing_ok = src == here
    rows_ok = bool(t.nrows == m and boff[-1] == m and int(orid.sum()) == m * (m - 1) // 2)
    t.free()
    return {"ok": bool(member_ok and order_ok and pairing_ok and rows_ok), "bucket_of_sample_by_oracle": member_ok,
            "sampled_neighbours_in_order": bool(order_ok), "key_matches_source_row": bool(pairing_ok),
            "row_ids_conserved": rows_ok, "sample_rows": int(len(here))}


def run_cfg4(args):
    import torch
    from hyperspace_1_b200 import abi, native
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    line = cfg4_measure(torch, native, abi, dev, args.rows or 80_000_000, args.steps, args.warmup, args.large_offsets,
                        check=not args.no_check)
    print(json.dumps(line), flush=True)


def run_skew(args):
    """Skewed keys — the case the reference's FAQ warns about (docs/_docs/04-ug-faqs.md:111-125: low-cardinality
    indexed columns skew the buckets): cfg2-shaped rows whose key column is (a) uniform, (b) half the rows one hot
    key, (c) ten distinct keys, (d) a heavy-tailed distribution. Build time and the usual index check per shape."""
    import torch
    from hyperspace_1_b200 import abi, native
    from oracle import oracle as _oracle   # the checker only
    from tools import index_check as ic
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    n, nb = args.rows or 100_000_000, 200
    g = torch.Generator(device=dev)
    g.manual_seed(20240606)
    price = torch.rand(n, generator=g, device=dev, dtype=torch.float64)
    disc = torch.rand(n, generator=g, device=dev, dtype=torch.float64)
    uni = torch.randint(0, 1 << 40, (n,), generator=g, device=dev, dtype=torch.int64)
    shapes = {
        "uniform": lambda: uni,
        "half_the_rows_one_key": lambda: torch.where(torch.rand(n, generator=g, device=dev) < 0.5, torch.full_like(uni, 77), uni),
        "ten_distinct_keys": lambda: torch.randint(0, 10, (n,), generator=g, device=dev, dtype=torch.int64) * 1_000_003,
        "heavy_tail": lambda: (1.0 / (torch.rand(n, generator=g, device=dev, dtype=torch.float64) + 1e-9)).to(torch.int64),
    }
    ctx = native.Context(0)
    ctx.set_output_location(abi.HS_MEM_DEVICE)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    out = {"config": "skew", "rows": n, "num_buckets": nb}
    for name, make in shapes.items():
        key = make().contiguous()
        torch.cuda.synchronize()
        cs = abi.DeviceColumnSet([{"type": abi.HS_INT64, "length": n, "data": key.data_ptr()},
                                  {"type": abi.HS_FLOAT64, "length": n, "data": price.data_ptr()},
                                  {"type": abi.HS_FLOAT64, "length": n, "data": disc.data_ptr()}])
        for _ in range(args.warmup):
            ctx.build_index_table(cs, 1, nb).free()
        ctx.enable_timing(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            t = ctx.build_index_table(cs, 1, nb)
            stages = {nm: round(ms, 3) for nm, ms, _ in ctx.last_timings() if ms > 0.01}
            t.free()
        e1.record(stream)
        torch.cuda.synchronize()
        ctx.enable_timing(False)
        ms = e0.elapsed_time(e1) / args.steps
        t = ctx.build_index_table(cs, 1, nb)
        torch.cuda.synchronize()
        boff = t.bucket_offsets()
        chk = ic.check_index(torch, _oracle, ic.table_tensors(torch, t, [torch.int64, torch.float64, torch.float64], dev),
                             boff, [key, price, disc], 1, nb)
        import numpy as np
        out[name] = {"ms_per_build": ms, "rows_per_s": n / (ms * 1e-3), "largest_bucket_rows": int(np.diff(boff).max()),
                     "stages_ms": stages, "check_ok": chk["ok"]}
        t.free()
        del key
    print(json.dumps(out), flush=True)
    ctx.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", choices=["cfg4", "cfg5", "parquet", "skew"], required=True)
    ap.add_argument("--rows", type=int, default=0, help="rows (per GPU for cfg5); default = the config's size that fits")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--large-offsets", action="store_true", help="cfg4: int64 offsets even below 2 GiB of key bytes")
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()
    if args.config == "cfg4":
        run_cfg4(args)
    elif args.config == "parquet":
        run_parquet(args)
    elif args.config == "skew":
        run_skew(args)
    else:
        run_cfg5(args)


def run_parquet(args):
    """K4 beside the kernels (timed separately): the files of a cfg2-shaped index written by the
    host library (pyarrow, snappy, what index_writer does by default) and by the device encoder
    (hs_parquet_plain_pages / hs_parquet_encode_pages + parquet_plain.py framing: PLAIN uncompressed, with
    dictionaries, with dictionaries and snappy), same rows, to a tmpfs directory."""
    import shutil
    import tempfile
    import numpy as np
    import pyarrow as pa
    import pyarrow.parquet as pq
    from hyperspace_1_b200 import index_writer, native, parquet_plain
    n, nb = args.rows or 20_000_000, 200
    rng = np.random.default_rng(20240601)
    cols = [pa.array(rng.integers(1, n, size=n)), pa.array(np.round(rng.random(n) * 104100 + 900, 2)),
            pa.array(rng.integers(0, 11, size=n) / 100.0)]
    names = ["l_orderkey", "l_extendedprice", "l_discount"]
    ctx = native.Context(0)
    arrays, boff = ctx.build_index(cols, 1, nb)
    root = tempfile.mkdtemp(dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    out = {"config": "parquet", "rows": n, "num_buckets": nb, "raw_bytes": n * 24}
    try:
        for name, fn in (("pyarrow_snappy", lambda d: index_writer.write_buckets(arrays, names, boff, d)[0]),
                         ("gpu_plain", lambda d: parquet_plain.write_buckets_plain(ctx, arrays, names, boff, d)),
                         ("gpu_dictionary", lambda d: parquet_plain.write_buckets_plain(ctx, arrays, names, boff, d, max_dict=4096)),
                         ("gpu_snappy", lambda d: parquet_plain.write_buckets_plain(ctx, arrays, names, boff, d, max_dict=4096, codec=1))):
            best, files = 1e9, []
            for rep in range(max(1, args.steps)):
                d = os.path.join(root, f"{name}_{rep}")
                t0 = time.perf_counter()
                files = fn(d)
                best = min(best, time.perf_counter() - t0)
                if rep + 1 < max(1, args.steps):
                    shutil.rmtree(d)
            size = sum(os.path.getsize(f) for f in files)
            t0 = time.perf_counter()
            rows = sum(pq.ParquetFile(f).read().num_rows for f in files)
            rd = time.perf_counter() - t0
            assert rows == n
            # the same files decoded on the device (hs_parquet_decode_pages; host parses the framing only)
            from hyperspace_1_b200 import parquet_read
            parquet_read.decode_files(ctx, sorted(files))   # warm
            t0 = time.perf_counter()
            got, _, rws = parquet_read.decode_files(ctx, sorted(files))
            gd = time.perf_counter() - t0
            assert sum(rws) == n and len(got[0]) == n
            out[name] = {"write_s": best, "rows_per_s": n / best, "file_bytes": size, "files": len(files),
                         "pyarrow_read_back_s": rd, "device_decode_read_back_s": gd}
    finally:
        shutil.rmtree(root, ignore_errors=True)
    print(json.dumps(out), flush=True)
    ctx.close()


def cfg5_measure(torch, dist, native, abi, ctx, rank, world, dev, n, steps, warmup):
    """cfg5 on `world` GPUs with an initialised context (comm set up when world > 1): n rows per GPU generated on
    the device, distributed build, build of a 10 % append, merge of the two runs, join; every index checked with
    tools/index_check.py. Rank 0 gets the result line, the others None."""
    ROOT = os.path.dirname(HERE)
    nb = 8192
    nd = n // 10

    def gen(m, seed):
        g = torch.Generator(device=dev)
        g.manual_seed(seed)
        k1 = torch.randint(0, 1 << 20, (m,), generator=g, device=dev, dtype=torch.int32)
        k2 = torch.randint(0, 1 << 40, (m,), generator=g, device=dev, dtype=torch.int64)
        v = torch.rand(m, generator=g, device=dev, dtype=torch.float64)
        return k1, k2, v

    def cols(t):
        return abi.DeviceColumnSet([{"type": abi.HS_INT32, "length": t[0].numel(), "data": t[0].data_ptr()},
                                    {"type": abi.HS_INT64, "length": t[1].numel(), "data": t[1].data_ptr()},
                                    {"type": abi.HS_FLOAT64, "length": t[2].numel(), "data": t[2].data_ptr()}])

    base, delta = gen(n, 20240604 + rank), gen(nd, 20240605 + rank)
    torch.cuda.synchronize()
    ctx.set_output_location(abi.HS_MEM_DEVICE)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    use_dist = world > 1

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(fn, steps, warmup, keep_last=False):
        out = None
        for _ in range(warmup):
            fn().free()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(steps):
            t = fn()
            if keep_last and i == steps - 1:
                out = t
            else:
                t.free()
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1) / steps
        if world > 1:
            tt = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        return ms, out

    def total(x):
        if world == 1:
            return int(x)
        tt = torch.tensor([x], dtype=torch.int64, device=dev)
        dist.all_reduce(tt)
        return int(tt.item())

    cb, cd = cols(base), cols(delta)
    ms_base, tb = timed(lambda: ctx.build_index_table(cb, 2, nb, dist=use_dist), steps, warmup, True)
    xb = ctx.comm_last_exchange() if use_dist else None
    ctx.enable_timing(True)
    for _ in range(2):   # the second call's numbers are reported (the first may still grow the block cache)
        ctx.build_index_table(cb, 2, nb, dist=use_dist).free()
        stages = {name: round(ms, 3) for name, ms, _ in ctx.last_timings()}
    ctx.enable_timing(False)
    ms_delta, td = timed(lambda: ctx.build_index_table(cd, 2, nb, dist=use_dist), steps, warmup, True)
    rows_owned = total(tb.nrows)
    vb, vd = tb.view(2), td.view(2)
    ms_merge, tm = timed(lambda: ctx.merge_tables([vb, vd]), steps, 1, True)
    vm = tm.view(2)
    ms_join, tj = timed(lambda: ctx.join_table(vm, vd, [0, 1, 2], [2]), steps, 1, True)
    out_rows = total(tj.nrows)
    merged_rows = total(tm.nrows)
    # every delta row finds itself in the merged index ...
    ok = out_rows >= nd * world and merged_rows == (n + nd) * world and rows_owned == n * world
    # ... and the three indexes hold what they should (tools/index_check.py: owner ranges, order inside every
    # bucket, bucket of a 1 M-row sample per rank recomputed by the oracle, row count and checksums over all ranks)
    from oracle import oracle as _oracle   # the checker only
    from tools import index_check as ic
    dts = [torch.int32, torch.int64, torch.float64]
    dd = dist if world > 1 else None
    torch.cuda.synchronize()
    checks = {
        "base": ic.check_index(torch, _oracle, ic.table_tensors(torch, tb, dts, dev), tb.bucket_offsets(), list(base),
                               2, nb, rank, world, dd),
        "refresh_delta": ic.check_index(torch, _oracle, ic.table_tensors(torch, td, dts, dev), td.bucket_offsets(),
                                        list(delta), 2, nb, rank, world, dd),
        "merged": ic.check_index(torch, _oracle, ic.table_tensors(torch, tm, dts, dev), tm.bucket_offsets(),
                                 [torch.cat([a, b]) for a, b in zip(base, delta)], 2, nb, rank, world, dd)}
    ok = ok and all(c["ok"] for c in checks.values())
    for t in (tj, tm, td, tb):
        t.free()
    line = None
    if rank == 0:
        row_bytes, key_bytes = 20, 12
        line = {"config": "cfg5", "n_gpus": world, "rows_per_gpu": n, "delta_rows_per_gpu": nd, "num_buckets": nb,
                "build_ms": ms_base, "build_rows_per_s": n * world / (ms_base * 1e-3),
                "build_algorithmic_bytes_per_row": key_bytes + 4 * row_bytes,
                "build_achieved_gbs_per_gpu": (key_bytes + 4 * row_bytes) * n / (ms_base * 1e-3) / 1e9,
                "refresh_incremental_ms": ms_delta, "refresh_rows_per_s": nd * world / (ms_delta * 1e-3),
                "merge_two_runs_ms": ms_merge, "merge_rows_per_s": (n + nd) * world / (ms_merge * 1e-3),
                "join_merged_x_delta_ms": ms_join, "join_output_rows": out_rows,
                "join_output_rows_per_s": out_rows / (ms_join * 1e-3), "checks_ok": bool(ok),
                "index_checks": checks}
        line["build_stages_ms_rank0"] = stages
        if xb:
            line["exchange_bytes_ms_rank0"] = xb
    return line


def run_cfg5(args):
    """cfg5: (k1:int32 < 2^20, k2:int64 < 2^40, v:float64), 8192 buckets, rows generated on the device per rank;
    distributed build of the base, distributed build of a 10 % append (refresh-incremental builds the delta only),
    then on every rank, over its own buckets and without communication: merge of the two runs per bucket and a
    join of the merged index with the delta index on (k1, k2)."""
    import torch
    import torch.distributed as dist
    from hyperspace_1_b200 import abi, native
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29517")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    ctx = native.Context(local)
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            uid = torch.frombuffer(bytearray(ctx.comm_unique_id()), dtype=torch.uint8).to(dev)
        dist.broadcast(uid, 0)
        ctx.comm_init(rank, world, bytes(uid.cpu().numpy().tobytes()))
    line = cfg5_measure(torch, dist, native, abi, ctx, rank, world, dev, args.rows or 500_000_000, args.steps, args.warmup)
    if rank == 0:
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    t0 = time.time()
    main()
