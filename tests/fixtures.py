"""Shared fixtures: the reference's sample tables and seeded synthetic generators."""
import numpy as np
import pyarrow as pa

# examples/scala/src/main/scala/App.scala:33-54
DEPARTMENTS = [(10, "Accounting", "New York"), (20, "Research", "Dallas"),
               (30, "Sales", "Chicago"), (40, "Operations", "Boston")]
EMPLOYEES = [(7369, "SMITH", 20), (7499, "ALLEN", 30), (7521, "WARD", 30), (7566, "JONES", 20),
             (7698, "BLAKE", 30), (7782, "CLARK", 10), (7788, "SCOTT", 20), (7839, "KING", 10),
             (7844, "TURNER", 30), (7876, "ADAMS", 20), (7900, "JAMES", 30), (7934, "MILLER", 10),
             (7902, "FORD", 20), (7654, "MARTIN", 30)]

# src/test/scala/com/microsoft/hyperspace/SampleData.scala:25-35
SAMPLE_DATA = [
    ("2017-09-03", "810a20a2baa24ff3ad493bfbf064569a", "donde", 2, 1000),
    ("2017-09-03", "fd093f8a05604515957083e70cb3dceb", "facebook", 1, 3000),
    ("2017-09-03", "af3ed6a197a8447cba8bc8ea21fad208", "facebook", 1, 3000),
    ("2017-09-03", "975134eca06c4711a0406d0464cbe7d6", "facebook", 1, 4000),
    ("2018-09-03", "e90a6028e15b4f4593eef557daf5166d", "ibraco", 2, 3000),
    ("2018-09-03", "576ed96b0d5340aa98a47de15c9f87ce", "facebook", 2, 3000),
    ("2018-09-03", "50d690516ca641438166049a6303650c", "ibraco", 2, 1000),
    ("2019-10-03", "380786e6495d4cd8a5dd4cc8d3d12917", "facebook", 2, 3000),
    ("2019-10-03", "ff60e4838b92421eafc3e6ee59a9e9f1", "miperro", 2, 2000),
    ("2019-10-03", "187696fe0a6a40cc9516bc6e47c70bc1", "facebook", 4, 3000)]


def departments_table():
    return pa.table({"deptId": pa.array([d[0] for d in DEPARTMENTS], pa.int32()),
                     "deptName": pa.array([d[1] for d in DEPARTMENTS]),
                     "location": pa.array([d[2] for d in DEPARTMENTS])})


def employees_table():
    return pa.table({"empId": pa.array([e[0] for e in EMPLOYEES], pa.int32()),
                     "empName": pa.array([e[1] for e in EMPLOYEES]),
                     "deptId": pa.array([e[2] for e in EMPLOYEES], pa.int32())})


def sample_table():
    return pa.table({"Date": pa.array([r[0] for r in SAMPLE_DATA]),
                     "RGUID": pa.array([r[1] for r in SAMPLE_DATA]),
                     "Query": pa.array([r[2] for r in SAMPLE_DATA]),
                     "imprs": pa.array([r[3] for r in SAMPLE_DATA], pa.int32()),
                     "clicks": pa.array([r[4] for r in SAMPLE_DATA], pa.int32())})


def lineitem(n_orders, seed=20240601, null_frac=0.0, rows=None):
    """cfg2-shaped rows: sparse TPC-H order keys, each repeated 1..7 times, shuffled.
    rows: adjust the repeat counts so that exactly that many rows come out (as the survey's 100 M)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    o = np.arange(n_orders, dtype=np.int64)
    keys = (o // 8) * 32 + o % 8 + 1
    reps = rng.integers(1, 8, size=n_orders)
    if rows is not None:
        diff = int(rows) - int(reps.sum())
        step = 1 if diff > 0 else -1
        while diff != 0:
            cand = np.nonzero((reps >= (1 if step > 0 else 2)) & (reps <= (6 if step > 0 else 7)))[0][: abs(diff)]
            assert len(cand), "cannot reach the requested row count with 1..7 repeats per order"
            reps[cand] += step
            diff -= step * len(cand)
    l_orderkey = np.repeat(keys, reps)
    rng.shuffle(l_orderkey)
    n = len(l_orderkey)
    price = np.round(rng.uniform(900.0, 105000.0, size=n), 2)
    disc = rng.integers(0, 11, size=n) / 100.0
    k = pa.array(l_orderkey)
    if null_frac > 0:
        mask = rng.random(n) < null_frac
        k = pa.array(l_orderkey, mask=mask)
    return [k, pa.array(price), pa.array(disc)], keys


def orders(keys, seed=20240602):
    rng = np.random.Generator(np.random.PCG64(seed))
    k = keys.copy()
    rng.shuffle(k)
    total = rng.uniform(850.0, 560000.0, size=len(k))
    return [pa.array(k), pa.array(total)]


def string_keys(n, seed=20240603):
    """cfg4-shaped keys: variable length, 64-symbol alphabet, shared 16-byte prefixes, some multi-byte."""
    rng = np.random.Generator(np.random.PCG64(seed))
    alphabet = np.frombuffer(b"abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789-_", dtype=np.uint8)
    lens = np.clip(8 + rng.geometric(1.0 / 17.0, size=n), 1, 96)
    prefixes = ["".join(chr(c) for c in alphabet[rng.integers(0, 64, size=16)]) for _ in range(50)]
    out = []
    for i in range(n):
        s = "".join(chr(c) for c in alphabet[rng.integers(0, 64, size=lens[i])])
        r = rng.random()
        if r < 0.10:
            s = prefixes[int(rng.integers(0, len(prefixes)))] + s[16:]
        elif r < 0.12:
            s = s[: len(s) // 2] + "é日" + s[len(s) // 2:]
        out.append(s)
    return pa.array(out), pa.array(np.arange(n, dtype=np.int64))
