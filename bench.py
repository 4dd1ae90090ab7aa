#!/usr/bin/env python
"""bench.py — index-build throughput (rows/s) of the covering-index hot path on B200.

One step = one pass of the build (K1 hash+histogram, K2 bucket scatter, K3 per-bucket sort) over one
batch of synthetic rows. Workload at N=1: BASELINE.json configs[1]: 100 M rows of
TPC-H-shaped lineitem {l_orderkey:int64, l_extendedprice:float64, l_discount:float64}, CoveringIndex on
l_orderkey, 200 buckets. N>1: every rank builds its own 100 M rows (weak scaling) and rows reach the GPU
that owns their bucket through the NCCL all-to-all inside hs_build_index_dist.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--rows R]

Prints ONE JSON line (see README / DESIGN.md for the keys).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BUCKETS_PER_GPU = 200
NUM_BUCKETS = BUCKETS_PER_GPU
ROW_BYTES = 24          # R: int64 key + two float64
KEY_BYTES = 8           # Kb
ALGO_BYTES_PER_ROW = KEY_BYTES + 4 * ROW_BYTES   # hash Kb + scatter 2R + sort 2R = 104
METRIC = "index_build_rows_per_s"
UNIT = "rows/s"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def wait_first_sample(self, timeout=5.0):
        # nvidia-smi takes the driver lock while it starts: let it settle before anything is timed
        t0 = time.perf_counter()
        while self.proc and not self.lines and time.perf_counter() - t0 < timeout:
            time.sleep(0.02)

    def stop(self, t_begin=None, t_end=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        lines = self.lines
        window = "timed region"
        if t_begin is not None:
            inside = [x for x in lines if t_begin <= x[0] <= t_end + 0.05]
            if inside:
                lines = inside
            else:
                window = "warm-up + timed region (timed region shorter than the sampling period)"
        for _, ln in lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "window": window, "reasons": sorted(reasons)}


def gen_lineitem_device(torch, n_rows, seed, device):
    """cfg2-shaped rows generated on the device: sparse TPC-H order keys repeated 1..7 times (mean 4),
    shuffled; prices U[900,105000) to 2 dp; discounts in {0.00..0.10}."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    n_orders = n_rows // 4
    o = torch.arange(n_orders, dtype=torch.int64, device=device)
    keys = (o // 8) * 32 + o % 8 + 1
    reps = torch.randint(1, 8, (n_orders,), generator=g, device=device)
    # adjust the counts so that the total is exactly n_rows
    diff = int(n_rows - int(reps.sum().item()))
    step = 1 if diff > 0 else -1
    i = 0
    while diff != 0:
        lo, hi = (1, 6) if step > 0 else (2, 7)
        cand = ((reps >= lo) & (reps <= hi)).nonzero().flatten()
        take = min(abs(diff), int(cand.numel()))
        reps[cand[:take]] += step
        diff -= step * take
        i += 1
        assert i < 16
    l_orderkey = torch.repeat_interleave(keys, reps)
    del reps, o
    perm = torch.randperm(n_rows, generator=g, device=device)
    l_orderkey = l_orderkey[perm].contiguous()
    del perm
    price = torch.round((900.0 + torch.rand(n_rows, generator=g, device=device, dtype=torch.float64) * 104100.0) * 100) / 100
    disc = torch.randint(0, 11, (n_rows,), generator=g, device=device).to(torch.float64) / 100.0
    return l_orderkey, price, disc, keys


def workload_config(n, world, num_buckets):
    """The `config` object of the line: the same for the native and the reference arm."""
    return {"workload": "cfg2: TPC-H-shaped lineitem, CoveringIndex on int64 l_orderkey including "
                        "l_extendedprice,l_discount, 200 buckets per GPU "
                        "(100 M rows and 200 buckets are added with every GPU: constant bucket size)",
            "rows_per_gpu": n, "num_buckets": num_buckets, "row_bytes": ROW_BYTES,
            "l2_policy": "inputs (2.4 GB per step) are larger than L2; no flush needed",
            "parallelism": f"buckets sharded over {world} GPU(s); rows reach their owners over NVLink peer memory"
            if world > 1 else "single GPU"}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def reference_arm(args, out):
    """--impl reference: the reference's CPU path for this metric on the host cores. The reference is
    Scala on Spark 2.4.2 and cannot run here (no JVM/Spark in the image), so the timed code is the
    oracle port of that path (oracle/hs_oracle.c) with every host thread. One step = one build of rank 0's
    rows of the native arm's workload (rows_per_gpu rows; --ref-rows bounds the sample if a host is too slow)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    from oracle import oracle
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import fixtures
    cores = oracle.set_threads(host_threads())   # torch.distributed.run exports OMP_NUM_THREADS=1
    n = args.ref_rows if args.ref_rows > 0 else args.rows
    cols, _ = fixtures.lineitem(max(1, n // 4), rows=n)
    assert len(cols[0]) == n
    arrays = [c.to_numpy(zero_copy_only=False) for c in cols]
    outs = [np.empty_like(a) for a in arrays]

    def one():
        perm, off = oracle.build_index(cols, 1, NUM_BUCKETS)
        for a, o in zip(arrays, outs):   # the build also materialises the index columns in their new order
            oracle.take_fixed(a, perm, o)
    for _ in range(max(0, min(args.warmup, 1))):
        one()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one()
    dt = time.perf_counter() - t0
    v = n * args.steps / dt
    sample = (f"{n} rows per step = " + ("one GPU's share of the workload (rows_per_gpu)" if n == args.rows else
              f"a bounded sample of the {args.rows} rows_per_gpu") + "; hash, bucket, per-bucket sort, take; "
              f"{cores} OpenMP threads")
    line = {"metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic", "impl": "reference",
            "config": workload_config(args.rows, max(1, args.gpus), NUM_BUCKETS),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference Spark local[*]: unavailable on this box (no JVM/Spark); timed = oracle port"}
    print(json.dumps(line), file=out, flush=True)


def _claim_stdout():
    """Libraries (NCCL prints its version banner) write to fd 1; the contract is ONE JSON line there. Send fd 1
    to stderr for the run and keep the real stdout for the result line."""
    real = os.fdopen(os.dup(1), "w")
    sys.stdout.flush()
    os.dup2(2, 1)
    return real


def main():
    real_stdout = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native")
    ap.add_argument("--rows", type=int, default=100_000_000, help="rows per GPU per step")
    ap.add_argument("--ref-rows", type=int, default=0,
                    help="rows per step of the CPU arms (0 = the same rows per step as one GPU of the native arm)")
    ap.add_argument("--e2e-steps", type=int, default=8, help="end-to-end steps per host thread")
    ap.add_argument("--e2e-threads", type=int, default=1, help="host threads (contexts) of the end-to-end leg at N=1")
    ap.add_argument("--e2e-sync", action="store_true", help="end-to-end leg without deferred results")
    ap.add_argument("--buckets", type=int, default=0,
                    help="bucket count of the index; default 200 per GPU (the bucket size, i.e. the index file size, "
                         "stays that of the 1-GPU configuration as rows are added with GPUs)")
    ap.add_argument("--no-join", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip the untimed check of one more build")
    ap.add_argument("--no-extras", action="store_true", help="skip the cfg4 (N=1) / cfg5 (N=8) side measurements")
    ap.add_argument("--cfg4-rows", type=int, default=80_000_000)
    ap.add_argument("--cfg5-rows", type=int, default=500_000_000, help="rows per GPU of the cfg5 side measurement")
    args = ap.parse_args()
    global NUM_BUCKETS
    NUM_BUCKETS = args.buckets if args.buckets > 0 else BUCKETS_PER_GPU * max(1, args.gpus)
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        reference_arm(args, real_stdout)
        return

    import numpy as np
    import torch
    from hyperspace_1_b200 import abi, native

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    ctx = native.Context(local)
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            uid = torch.frombuffer(bytearray(ctx.comm_unique_id()), dtype=torch.uint8).to(dev)
        dist.broadcast(uid, 0)
        ctx.comm_init(rank, world, bytes(uid.cpu().numpy().tobytes()))

    n = args.rows
    key, price, disc, okeys = gen_lineitem_device(torch, n, 20240601 + rank, dev)
    torch.cuda.synchronize()

    def dev_cols():
        return abi.DeviceColumnSet([
            {"type": abi.HS_INT64, "length": n, "data": key.data_ptr()},
            {"type": abi.HS_FLOAT64, "length": n, "data": price.data_ptr()},
            {"type": abi.HS_FLOAT64, "length": n, "data": disc.data_ptr()}])

    cs = dev_cols()
    use_dist = world > 1
    ctx.set_output_location(abi.HS_MEM_DEVICE)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)

    def step():
        t = ctx.build_index_table(cs, 1, NUM_BUCKETS, dist=use_dist)
        return t

    sampler = ClockSampler(local)
    sampler.start()
    sampler.wait_first_sample()
    for _ in range(max(args.warmup, 3) if args.warmup > 0 else 0):
        step().free()
    # ---- timed region: device-resident in -> device-resident out
    ctx.enable_timing(True)
    barrier()
    t_begin = time.perf_counter()
    l0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage_ms = {}
    xchg = []
    e0.record(stream)
    for _ in range(args.steps):
        t = step()
        for name, ms, launches in ctx.last_timings():
            a = stage_ms.setdefault(name, [0.0, 0])
            a[0] += ms
            a[1] += launches
        if use_dist:
            xchg.append(ctx.comm_last_exchange())
        t.free()
    e1.record(stream)
    barrier()
    launches = ctx.launch_count() - l0
    clocks = sampler.stop(t_begin, time.perf_counter())
    ms_total = e0.elapsed_time(e1)
    ctx.enable_timing(False)
    if dist is not None:
        tt = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total = float(tt.item())
    ms_step = ms_total / args.steps
    value = n * world / (ms_step * 1e-3)

    # ---- dominant kernel and its roofline (device time of its launches inside the timed region)
    peak, peak_kind = peaks()
    algo = {"hash_hist": KEY_BYTES, "bucket_scatter": 2 * ROW_BYTES, "msd_scatter": 2 * ROW_BYTES,
            "leaf_sort": 2 * ROW_BYTES, "msd_hist": KEY_BYTES}
    kernel_of = {"hash_hist": "k_hash_i64x2", "bucket_scatter": "k_partition<0>", "msd_scatter": "k_partition<1>",
                 "leaf_sort": "k_leaf_sort32<KS_I64>", "msd_hist": "k_msd_hist"}
    per_stage = {k: {"ms_per_step": v[0] / args.steps, "launches_per_step": v[1] / args.steps}
                 for k, v in stage_ms.items()}
    dom = max((k for k in per_stage if k in algo), key=lambda k: per_stage[k]["ms_per_step"])
    dom_ms = per_stage[dom]["ms_per_step"]
    achieved = algo[dom] * n / (dom_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": kernel_of[dom], "stage": dom, "achieved": achieved, "peak": peak,
                "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peak_kind})", "unit": "GB/s", "frac": achieved / peak,
                "algorithmic_bytes_per_row": algo[dom], "rows_per_launch": n, "ms_per_launch": dom_ms,
                "traffic": None}
    # DRAM bytes per launch of that kernel from an ncu capture (profiles/dominant_kernel_traffic.json). The file
    # names the sha256 of the kernel's source file it was taken with: when the source has changed since, the
    # figure is stale and is left out rather than repeated.
    tr_path = os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")
    if os.path.exists(tr_path):
        try:
            import hashlib
            tr = json.load(open(tr_path))
            src = os.path.join(ROOT, tr.get("source_file", ""))
            sha = hashlib.sha256(open(src, "rb").read()).hexdigest() if os.path.isfile(src) else None
            if tr.get("kernel") == kernel_of[dom] and tr.get("rows") == n and sha and sha == tr.get("source_sha256"):
                roofline["traffic"] = tr["dram_bytes_per_launch"]
            else:
                roofline["traffic_note"] = "no ncu capture for this kernel at this source revision"
        except Exception:
            pass
    build_gbs = ALGO_BYTES_PER_ROW * n / (ms_step * 1e-3) / 1e9
    build_roofline = {"algorithmic_bytes_per_row": ALGO_BYTES_PER_ROW, "achieved": build_gbs, "unit": "GB/s",
                      "frac_of_measured_peak": build_gbs / peak, "frac_of_8TBs": build_gbs / 8000.0}

    # ---- outside the timing: one more build, checked (every rank; results all-reduced)
    index_check = None
    if not args.no_check:
        from oracle import oracle as _oracle   # the checker only
        from tools import index_check as ic
        tchk = step()
        torch.cuda.synchronize()
        outs = ic.table_tensors(torch, tchk, [torch.int64, torch.float64, torch.float64], dev)
        index_check = ic.check_index(torch, _oracle, outs, tchk.bucket_offsets(), [key, price, disc], 1, NUM_BUCKETS,
                                     rank, world, dist)
        del outs
        tchk.free()

    # ---- e2e: the same build through the C ABI with HOST buffers (pinned in, pinned out)
    e2e = None
    if args.e2e_steps > 0:
        hk, hp, hd = key.cpu().pin_memory(), price.cpu().pin_memory(), disc.cpu().pin_memory()
        hcs = abi.DeviceColumnSet([
            {"type": abi.HS_INT64, "length": n, "data": hk.data_ptr()},
            {"type": abi.HS_FLOAT64, "length": n, "data": hp.data_ptr()},
            {"type": abi.HS_FLOAT64, "length": n, "data": hd.data_ptr()}])
        ctx.set_output_location(abi.HS_MEM_HOST)
        ctx.build_index_table(hcs, 1, NUM_BUCKETS, dist=use_dist).free()   # warm the pinned pool
        # One caller thread. With deferred results (hs_ctx_set_deferred_results) hs_build_index returns once the
        # result's download is queued; the thread starts the next build — its upload runs beside that download,
        # PCIe carries both directions at once — and then collects the previous table (hs_table_wait) and reads
        # it. Every step still copies its own inputs in and its own result out inside the timed region.
        n_threads = max(1, args.e2e_threads) if world == 1 else 1
        deferred = not args.e2e_sync
        extra = [native.Context(local) for _ in range(n_threads - 1)]
        for cx in [ctx] + extra:
            cx.set_output_location(abi.HS_MEM_HOST)
            cx.set_deferred_results(deferred)
            for _ in range(2):      # two results are alive at a time: warm two sets of pinned buffers
                a = cx.build_index_table(hcs, 1, NUM_BUCKETS, dist=use_dist if cx is ctx else False)
                b = cx.build_index_table(hcs, 1, NUM_BUCKETS, dist=use_dist if cx is ctx else False)
                a.free()
                b.free()
        results = {}

        def worker(cx, k):
            fk, prev = 0, None
            for _ in range(args.e2e_steps):
                t = cx.build_index_table(hcs, 1, NUM_BUCKETS, dist=use_dist)
                if prev is not None:
                    prev.wait()
                    fk = C.c_int64.from_address(prev.column(0).data).value if prev.nrows else 0  # read the result
                    prev.free()
                prev = t
            prev.wait()
            fk = C.c_int64.from_address(prev.column(0).data).value if prev.nrows else 0
            prev.free()
            results[k] = fk
        barrier()
        t0 = time.perf_counter()
        threads = [threading.Thread(target=worker, args=(cx, i + 1)) for i, cx in enumerate(extra)]
        for th in threads:
            th.start()
        worker(ctx, 0)
        for th in threads:
            th.join()
        barrier()
        total_steps = args.e2e_steps * n_threads
        dt = (time.perf_counter() - t0) / total_steps
        first_key = results[0]
        for cx in extra:
            cx.close()
        ctx.set_deferred_results(False)
        if dist is not None:
            tt = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
        e2e = {"value": n * world / dt, "unit": UNIT, "h2d_bytes_per_step": n * ROW_BYTES,
               "d2h_bytes_per_step": n * ROW_BYTES, "ms_per_step": dt * 1e3, "steps": total_steps,
               "host_threads": n_threads, "deferred_results": deferred,
               "note": "pinned host buffers in and out through hs_build_index on one caller thread; the download of "
                       "build k (hs_ctx_set_deferred_results / hs_table_wait) runs beside the upload of build k+1",
               "first_key": int(first_key)}
        del hk, hp, hd
        ctx.set_output_location(abi.HS_MEM_DEVICE)

    # ---- indexed join (cfg3): lineitem x orders over two indexes, bucket-local, no communication
    join = None
    filt = None
    if not args.no_join:
        jerr, jlaunches = "", 0
        flocal = [0.0, 0.0, 0.0, 0.0]
        try:
            g = torch.Generator(device=dev)
            g.manual_seed(20240602 + rank)
            if world > 1:
                # weak scaling: this rank contributes the orders of its own lineitem keys; keys of different
                # ranks are made disjoint so that every lineitem row still has exactly one order
                okeys = okeys * world + rank
                key = key * world + rank
                torch.cuda.synchronize()
                cs = dev_cols()
            n_o = int(okeys.numel())
            ok = okeys[torch.randperm(n_o, generator=g, device=dev)].contiguous()
            ototal = 850.0 + torch.rand(n_o, generator=g, device=dev, dtype=torch.float64) * 559150.0
            torch.cuda.synchronize()   # the library has its own stream: torch's writes must have landed
            ocs = abi.DeviceColumnSet([{"type": abi.HS_INT64, "length": n_o, "data": ok.data_ptr()},
                                       {"type": abi.HS_FLOAT64, "length": n_o, "data": ototal.data_ptr()}])
            lt = ctx.build_index_table(cs, 1, NUM_BUCKETS, dist=use_dist)
            ot = ctx.build_index_table(ocs, 1, NUM_BUCKETS, dist=use_dist)

            def view(tbl, ncols):
                arr = (abi.hs_column * ncols)(*[tbl.column(i) for i in range(ncols)])
                dcs = abi.DeviceColumnSet([])
                dcs.c, dcs.n = arr, ncols
                return native.IndexView(dcs, 1, tbl.bucket_offsets())
            lv, ov = view(lt, 3), view(ot, 2)
            for _ in range(2):
                ctx.join_table(lv, ov, [0, 1, 2], [1]).free()
            torch.cuda.synchronize()   # the join needs no communication: ranks are timed independently, max taken
            j0, j1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            jl0 = ctx.launch_count()
            j0.record(stream)
            jsteps = max(1, min(args.steps, 5))
            for _ in range(jsteps):
                jt = ctx.join_table(lv, ov, [0, 1, 2], [1])
                n_out = jt.nrows
                jt.free()
            j1.record(stream)
            torch.cuda.synchronize()
            jlocal = [j0.elapsed_time(j1) / jsteps, float(n_out), float(lt.nrows), float(ot.nrows), 1.0]
            jlaunches = ctx.launch_count() - jl0
            # ---- filtered index scan over the same lineitem index (FilterIndexRule's plan, K7)
            thr = int(key[: min(n, 1_000_000)].median().item())
            probe = int(key[0].item())
            for terms, slot in (([("ge", 0, thr)], 0), ([("eq", 0, probe)], 1)):
                ctx.filter_table(lv, terms, [0, 1, 2]).free()
                torch.cuda.synchronize()
                f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                f0.record(stream)
                for _ in range(jsteps):
                    ft = ctx.filter_table(lv, terms, [0, 1, 2])
                    flocal[2 * slot + 1] = float(ft.nrows)
                    ft.free()
                f1.record(stream)
                torch.cuda.synchronize()
                flocal[2 * slot] = f0.elapsed_time(f1) / jsteps
            lt.free()
            ot.free()
            del ok, ototal
        except Exception as ex:  # the join line is extra; the build line must still print
            jlocal = [0.0, 0.0, 0.0, 0.0, 0.0]
            jerr = str(ex)[:200]
        # every rank reaches the reductions, whether its join worked or not
        fmx = torch.tensor(flocal, dtype=torch.float64, device=dev)
        fsum = fmx.clone()
        if dist is not None:
            dist.all_reduce(fmx, op=dist.ReduceOp.MAX)
            dist.all_reduce(fsum, op=dist.ReduceOp.SUM)
        if float(fmx[0].item()) > 0:
            rows_all = n * world
            m_range, m_point = int(fsum[1].item()), int(fsum[3].item())
            ms_range, ms_point = float(fmx[0].item()), float(fmx[2].item())
            fbytes = rows_all * KEY_BYTES + 2 * m_range * ROW_BYTES
            filt = {"range_scan": {"predicate": "l_orderkey >= median", "ms_per_step": ms_range, "rows_scanned": rows_all,
                                   "rows_out": m_range, "scanned_rows_per_s": rows_all / (ms_range * 1e-3),
                                   "algorithmic_bytes": fbytes, "achieved_GBps": fbytes / (ms_range * 1e-3) / 1e9,
                                   "frac_of_measured_peak_per_gpu": fbytes / (ms_range * 1e-3) / 1e9 / peak / world},
                    "point_lookup": {"predicate": "l_orderkey = literal (one bucket, binary search of the sorted run)",
                                     "ms_per_step": ms_point, "rows_out": m_point}}
        tot = torch.tensor(jlocal, dtype=torch.float64, device=dev)
        mx = tot.clone()
        if dist is not None:
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        if int(tot[4].item()) != world:
            join = {"error": jerr if jlocal[4] == 0.0 else "the join failed on another rank"}
        else:
            jms = float(mx[0].item())
            n_out_all, n_left_all, n_right_all = int(tot[1].item()), int(tot[2].item()), int(tot[3].item())
            jbytes = n_left_all * ROW_BYTES + n_right_all * 16 + n_out_all * 32
            join = {"metric": "indexed_join_output_rows_per_s", "value": n_out_all / (jms * 1e-3), "unit": UNIT,
                    "ms_per_step": jms, "output_rows": n_out_all, "left_rows": n_left_all, "right_rows": n_right_all,
                    "algorithmic_bytes": jbytes, "achieved_GBps": jbytes / (jms * 1e-3) / 1e9,
                    "frac_of_measured_peak_per_gpu": jbytes / (jms * 1e-3) / 1e9 / peak / world,
                    "communication": "none (bucket-local; both indexes share the owner map)",
                    "gpu_launches": int(jlaunches)}

    # ---- CPU baseline beside it: the oracle port on a bounded sample (rank 0, N=1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        import pyarrow as pa
        from oracle import oracle
        m = min(args.ref_rows, n) if args.ref_rows > 0 else n
        cores = oracle.set_threads(host_threads())
        sample = [pa.array(key[:m].cpu().numpy()), pa.array(price[:m].cpu().numpy()), pa.array(disc[:m].cpu().numpy())]
        arrays = [c.to_numpy(zero_copy_only=False) for c in sample]
        t0 = time.perf_counter()
        perm, _ = oracle.build_index(sample, 1, NUM_BUCKETS)
        for a in arrays:
            oracle.take_fixed(a, perm)
        dt = time.perf_counter() - t0
        cpu = {"value": m / dt, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"first {m} rows of the same workload, one pass (hash, bucket, per-bucket sort, take)",
               "seconds": dt}

    # ---- the other BASELINE.json configs, measured beside the headline (tools/bench_configs.py): cfg4 (500 M-row
    # shape, string keys, 1000 buckets) at 80 M rows on one GPU; cfg5 (composite key, 8192 buckets, 10 % refresh,
    # merge and join over two runs per bucket) at 8 GPUs x 500 M rows. Each carries its own checks.
    extra = {}
    if not args.no_extras:
        del key, price, disc
        torch.cuda.empty_cache()
        from tools import bench_configs as bc
        try:
            if world == 1:
                x = bc.cfg4_measure(torch, native, abi, dev, args.cfg4_rows, 3, 2, ctx=ctx)
                x["frac_of_measured_peak"] = x["achieved_gbs"] / peak
                extra["cfg4"] = x
            elif world == 8:
                x = bc.cfg5_measure(torch, dist, native, abi, ctx, rank, world, dev, args.cfg5_rows, 2, 1)
                if x is not None:
                    x["build_frac_of_measured_peak_per_gpu"] = x["build_achieved_gbs_per_gpu"] / peak
                    extra["cfg5"] = x
        except Exception as ex:   # an extra must not cost the headline line
            extra["error"] = str(ex)[:300]

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "int64", "data": "synthetic",
                "config": workload_config(n, world, NUM_BUCKETS),
                "roofline": roofline, "build_roofline": build_roofline, "stages": per_stage,
                "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "join": join,
                "filter": filt, "dist_check" if world > 1 else "index_check": index_check, "extra": extra}
        if use_dist and xchg:
            b = sum(x[0] for x in xchg) / len(xchg)
            ms = sum(x[1] for x in xchg) / len(xchg)
            line["all_to_all"] = {"bytes_sent_per_gpu": b, "ms": ms, "bus_GBps": b / (ms * 1e-3) / 1e9 if ms else None,
                                  "frac_of_900": (b / (ms * 1e-3) / 1e9 / 900.0) if ms else None}
        print(json.dumps(line), file=real_stdout, flush=True)
    ctx.comm_destroy()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
