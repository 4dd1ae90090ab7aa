"""End to end through the reference-shaped API on the GPU: quick start (examples/scala/src/main/scala/App.scala),
the verifyIndexUsage invariant (HT/index/E2EHyperspaceRulesTest.scala:1004-1019: result with indexes == result
without), incremental refresh leaving several files per bucket (:411-480), optimize, and the lifecycle states."""
import os

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

from fixtures import EMPLOYEES, departments_table, employees_table, lineitem, orders, sample_table

pytestmark = pytest.mark.gpu


def _rows(t):
    return sorted(zip(*[t[c].to_pylist() for c in t.column_names]), key=lambda r: tuple((x is None, x) for x in r))


def _verify_index_usage(session, df, expect_indexes):
    session.disable_hyperspace()
    base = df.collect()
    assert "IndexScan" not in df.explain_string()
    session.enable_hyperspace()
    plan = df.explain_string()
    for name in expect_indexes:
        assert f"Name: {name}," in plan, plan
    got = df.collect()
    assert got.column_names == base.column_names
    assert _rows(got) == _rows(base)
    return got


@pytest.fixture()
def env(tmp_path, hs):
    from hyperspace_1_b200 import Session
    s = Session(system_path=str(tmp_path / "indexes"))
    yield s, tmp_path
    if s._ctx is not None:
        s._ctx.close()


def test_quick_start(env):
    from hyperspace_1_b200 import Hyperspace, IndexConfig, log_entry as le
    s, tmp = env
    os.makedirs(tmp / "departments")
    os.makedirs(tmp / "employees")
    pq.write_table(departments_table(), tmp / "departments" / "part-0.parquet")
    pq.write_table(employees_table(), tmp / "employees" / "part-0.parquet")
    dept, emp = s.read_parquet(str(tmp / "departments")), s.read_parquet(str(tmp / "employees"))
    hs = Hyperspace(s)
    hs.createIndex(dept, IndexConfig("deptIndex", ["deptId"], ["deptName"]))
    hs.createIndex(emp, IndexConfig("empIndex", ["deptId"], ["empName"]))
    # on-disk layout: log 0 = CREATING, 1 = ACTIVE, latestStable, v__=0 with 3 bucket files
    d = tmp / "indexes" / "deptIndex"
    lm = le.IndexLogManager(str(d))
    assert lm.get_log(0)["state"] == "CREATING" and lm.get_log(1)["state"] == "ACTIVE"
    assert os.path.exists(d / "_hyperspace_log" / "latestStable")
    names = sorted(os.listdir(d / "v__=0"))
    assert [n[:10] for n in names] == ["part-00053", "part-00122", "part-00181"]
    assert all(n.endswith(".c000.snappy.parquet") for n in names)
    b53 = pq.ParquetFile(d / "v__=0" / names[0]).read()
    assert b53.column_names == ["deptId", "deptName"] and b53["deptId"].to_pylist() == [20, 40]
    listed = [os.path.basename(f["name"]) for f in le.content_files(lm.get_log(1)["content"])]
    assert sorted(listed) == names
    assert sorted(hs.indexes()["name"].to_pylist()) == ["deptIndex", "empIndex"]
    with pytest.raises(Exception):
        hs.createIndex(dept, IndexConfig("DEPTINDEX", ["deptId"], []))   # names are case-insensitive
    # filter and join of the quick start
    f = _verify_index_usage(s, dept.filter("deptId = 20").select("deptName"), ["deptIndex"])
    assert f["deptName"].to_pylist() == ["Research"]
    j = emp.join(dept, emp["deptId"] == dept["deptId"]).select(emp["empName"], dept["deptName"])
    got = _verify_index_usage(s, j, ["empIndex", "deptIndex"])
    deptname = {10: "Accounting", 20: "Research", 30: "Sales"}
    assert _rows(got) == sorted((n, deptname[k]) for _, n, k in EMPLOYEES)
    out = []
    hs.explain(j, redirectFunc=out.append)
    assert "GpuBucketedSortMergeJoin" in out[0] and "Plan without indexes" in out[0]


@pytest.mark.parametrize("how", ["left_outer", "right_outer", "full_outer", "left_semi", "left_anti"])
def test_join_types_through_the_api(env, how):
    """JoinIndexRule rewrites joins of every type (JoinIndexRule.scala:57-60): indexed result == plain result,
    with an employee in a department that does not exist, a department without employees and a NULL key."""
    from hyperspace_1_b200 import Hyperspace, IndexConfig
    s, tmp = env
    os.makedirs(tmp / "departments")
    os.makedirs(tmp / "employees")
    dept = departments_table()
    emp = employees_table()
    emp = pa.concat_tables([emp, pa.table({"empId": pa.array([9001, 9002], type=emp.schema.field("empId").type),
                                           "empName": ["NOBODY", "NULLKEY"],
                                           "deptId": pa.array([77, None], type=emp.schema.field("deptId").type)})])
    pq.write_table(dept, tmp / "departments" / "part-0.parquet")
    pq.write_table(emp, tmp / "employees" / "part-0.parquet")
    d, e = s.read_parquet(str(tmp / "departments")), s.read_parquet(str(tmp / "employees"))
    hs = Hyperspace(s)
    hs.createIndex(d, IndexConfig("deptIndex", ["deptId"], ["deptName"]))
    hs.createIndex(e, IndexConfig("empIndex", ["deptId"], ["empName"]))
    j = e.join(d, e["deptId"] == d["deptId"], how)
    j = j.select(e["empName"]) if how in ("left_semi", "left_anti") else j.select(e["empName"], d["deptName"])
    got = _verify_index_usage(s, j, ["empIndex", "deptIndex"])
    names = got["empName"].to_pylist()
    if how == "left_anti":
        assert sorted(names) == ["NOBODY", "NULLKEY"]
    if how == "left_semi":
        assert "NOBODY" not in names and len(names) == len(EMPLOYEES)
    if how in ("right_outer", "full_outer"):
        assert "Operations" in got["deptName"].to_pylist()      # department 40 has no employees
    if how in ("left_outer", "full_outer"):
        assert "NOBODY" in names and "NULLKEY" in names


def test_string_and_composite_indexes_through_the_api(env):
    from hyperspace_1_b200 import Hyperspace, IndexConfig
    s, tmp = env
    os.makedirs(tmp / "sample")
    pq.write_table(sample_table(), tmp / "sample" / "part-0.parquet")
    df = s.read_parquet(str(tmp / "sample"))
    s.conf["spark.hyperspace.index.numBuckets"] = "5"       # python/hyperspace/testing/utils.py:15
    hs = Hyperspace(s)
    hs.createIndex(df, IndexConfig("qIndex", ["Query"], ["clicks", "imprs"]))
    hs.createIndex(df, IndexConfig("ciIndex", ["clicks", "Query"], ["Date"]))
    _verify_index_usage(s, df.filter("Query = 'facebook'").select("clicks", "Query"), ["qIndex"])
    _verify_index_usage(s, df.filter("clicks >= 3000 AND Query = 'ibraco'").select("Date", "Query"), ["ciIndex"])
    # column names resolve case-insensitively to the source spelling (CreateIndexTest.scala:99-111)
    hs.createIndex(df, IndexConfig("caseIndex", ["rguid"], ["DATE"]))
    e = hs.index("caseIndex")
    assert e["derivedDataset"]["properties"]["columns"] == {"indexed": ["RGUID"], "included": ["Date"]}


def test_incremental_refresh_then_join_then_optimize(env):
    from hyperspace_1_b200 import Hyperspace, IndexConfig, log_entry as le
    s, tmp = env
    s.conf["spark.hyperspace.index.numBuckets"] = "16"
    lcols, keys = lineitem(6000)
    lt = pa.table(lcols, names=["l_orderkey", "l_extendedprice", "l_discount"])
    n = lt.num_rows
    os.makedirs(tmp / "lineitem")
    os.makedirs(tmp / "orders")
    pq.write_table(lt.slice(0, n * 9 // 10), tmp / "lineitem" / "part-0.parquet")
    pq.write_table(pa.table(orders(keys), names=["o_orderkey", "o_totalprice"]), tmp / "orders" / "part-0.parquet")
    hs = Hyperspace(s)
    li = s.read_parquet(str(tmp / "lineitem"))
    od = s.read_parquet(str(tmp / "orders"))
    hs.createIndex(li, IndexConfig("liIndex", ["l_orderkey"], ["l_extendedprice", "l_discount"]))
    hs.createIndex(od, IndexConfig("oIndex", ["o_orderkey"], ["o_totalprice"]))
    q = lambda a, b: a.join(b, a["l_orderkey"] == b["o_orderkey"]).select("l_orderkey", "l_extendedprice", "o_totalprice")
    _verify_index_usage(s, q(li, od), ["liIndex", "oIndex"])
    # append 10 % (BASELINE config 5 shape): the old index no longer matches, incremental refresh indexes the delta
    pq.write_table(lt.slice(n * 9 // 10), tmp / "lineitem" / "part-1.parquet")
    li2 = s.read_parquet(str(tmp / "lineitem"))
    s.enable_hyperspace()
    assert "liIndex" not in q(li2, od).explain_string()
    hs.refreshIndex("liIndex", "incremental")
    d = tmp / "indexes" / "liIndex"
    assert sorted(os.listdir(d)) == ["_hyperspace_log", "v__=0", "v__=1"]
    e = hs.index("liIndex")
    files = le.content_files(e["content"])
    assert {os.path.basename(os.path.dirname(f["name"])) for f in files} == {"v__=0", "v__=1"}
    assert e["state"] == "ACTIVE" and e["id"] == 3
    assert sum(pq.read_metadata(le.from_uri(f["name"])).num_rows for f in files) == n
    got = _verify_index_usage(s, q(li2, od), ["liIndex", "oIndex"])     # two runs per bucket merged on the GPU
    assert got.num_rows == n
    assert hs.refreshIndex("liIndex", "incremental") is not None       # nothing changed: logged no-op
    assert hs.index("liIndex")["id"] == 3
    # optimize: one file per bucket again, same answers
    hs.optimizeIndex("liIndex", "full")
    e = hs.index("liIndex")
    files = le.content_files(e["content"])
    per_bucket = {}
    for f in files:
        per_bucket.setdefault(le.bucket_id_of(f["name"]), []).append(f)
    assert all(len(v) == 1 for v in per_bucket.values())
    assert {os.path.basename(os.path.dirname(f["name"])) for f in files} == {"v__=2"}
    _verify_index_usage(s, q(li2, od), ["liIndex", "oIndex"])
    for f in files:   # every file sorted on the key, every row in the bucket its name says
        k = pq.ParquetFile(le.from_uri(f["name"])).read()["l_orderkey"].to_numpy()
        assert (np.diff(k) >= 0).all()
    # full refresh after a delete in the source
    os.remove(tmp / "lineitem" / "part-1.parquet")
    with pytest.raises(Exception):
        hs.refreshIndex("liIndex", "incremental")     # deletes need lineage (this index was built without)
    hs.cancel("liIndex") if hs.index("liIndex")["state"] != "ACTIVE" else None
    hs.refreshIndex("liIndex", "full")
    li3 = s.read_parquet(str(tmp / "lineitem"))
    assert _verify_index_usage(s, q(li3, od), ["liIndex", "oIndex"]).num_rows == n * 9 // 10


def _lineitem_files(tmp, n_orders=4000, parts=4):
    lcols, keys = lineitem(n_orders)
    lt = pa.table(lcols, names=["l_orderkey", "l_extendedprice", "l_discount"])
    os.makedirs(tmp / "lineitem")
    os.makedirs(tmp / "orders")
    step = lt.num_rows // parts
    for i in range(parts):
        pq.write_table(lt.slice(i * step, step if i < parts - 1 else lt.num_rows - i * step),
                       tmp / "lineitem" / f"part-{i}.parquet")
    pq.write_table(pa.table(orders(keys), names=["o_orderkey", "o_totalprice"]), tmp / "orders" / "part-0.parquet")
    return lt, step


def test_lineage_column_and_refresh_after_deletes(env):
    """spark.hyperspace.index.lineage.enabled: every index row carries the id of its source file
    (CreateIndexTest.scala:159-243: schema = indexed ++ included ++ _data_file_id, ids from 0); an incremental
    refresh after a source file was deleted rewrites the index without that file's rows
    (RefreshIndexTest.scala:58-110) — bucket files stay sorted and in their buckets."""
    from hyperspace_1_b200 import Hyperspace, IndexConfig, log_entry as le
    s, tmp = env
    s.conf["spark.hyperspace.index.numBuckets"] = "8"
    s.conf["spark.hyperspace.index.lineage.enabled"] = "true"
    lt, step = _lineitem_files(tmp)
    hs = Hyperspace(s)
    li = s.read_parquet(str(tmp / "lineitem"))
    hs.createIndex(li, IndexConfig("liIndex", ["l_orderkey"], ["l_extendedprice"]))
    e = hs.index("liIndex")
    assert e["derivedDataset"]["properties"]["properties"]["lineage"] == "true"
    assert e["derivedDataset"]["properties"]["columns"] == {"indexed": ["l_orderkey"], "included": ["l_extendedprice"]}
    src = {os.path.basename(f["name"]): f["id"] for f in
           le.content_files(e["source"]["plan"]["properties"]["relations"][0]["data"]["properties"]["content"])}
    assert sorted(src.values()) == [0, 1, 2, 3]
    files = [le.from_uri(f["name"]) for f in le.content_files(e["content"])]
    t = pa.concat_tables([pq.ParquetFile(f).read() for f in files])
    assert t.column_names == ["l_orderkey", "l_extendedprice", "_data_file_id"]
    for name, fid in src.items():      # the rows tagged with a file id are exactly that file's rows
        want = pq.ParquetFile(tmp / "lineitem" / name).read()
        got = t.filter(pa.compute.equal(t["_data_file_id"], fid))
        assert sorted(got["l_orderkey"].to_pylist()) == sorted(want["l_orderkey"].to_pylist())
    q = lambda df: df.filter("l_orderkey >= 0").select("l_orderkey", "l_extendedprice")
    _verify_index_usage(s, q(li), ["liIndex"])
    # delete one source file, append another: incremental refresh = old rows minus the deleted file + the new rows
    os.remove(tmp / "lineitem" / "part-1.parquet")
    extra = lt.slice(0, 500)
    pq.write_table(extra, tmp / "lineitem" / "part-9.parquet")
    hs.refreshIndex("liIndex", "incremental")
    e2 = hs.index("liIndex")
    files = [le.from_uri(f["name"]) for f in le.content_files(e2["content"])]
    assert {os.path.basename(os.path.dirname(f)) for f in files} == {"v__=1"}     # old files were rewritten
    t2 = pa.concat_tables([pq.ParquetFile(f).read() for f in files])
    assert src["part-1.parquet"] not in set(t2["_data_file_id"].to_pylist())
    assert t2.num_rows == lt.num_rows - step + 500
    for f in files:
        k = pq.ParquetFile(f).read()["l_orderkey"].to_numpy()
        assert (np.diff(k) >= 0).all()
    src2 = {os.path.basename(f["name"]): f["id"] for f in
            le.content_files(e2["source"]["plan"]["properties"]["relations"][0]["data"]["properties"]["content"])}
    assert all(src2[k] == v for k, v in src.items() if k in src2) and src2["part-9.parquet"] not in src.values()
    li2 = s.read_parquet(str(tmp / "lineitem"))
    got = _verify_index_usage(s, q(li2), ["liIndex"])
    assert got.num_rows == t2.num_rows


@pytest.mark.parametrize("mode", ["hybrid_conf", "quick_refresh"])
def test_hybrid_scan_appended_and_deleted_files(env, mode):
    """HybridScanSuite.scala:186-307,371-663: with source files appended and deleted after the build, queries
    still use the index — appended files are bucketed on the fly and merged bucket by bucket, rows of deleted
    files are filtered through the lineage column — and return what the plain query returns. Reached either
    through spark.hyperspace.index.hybridscan.enabled or after refreshIndex(name, "quick")."""
    from hyperspace_1_b200 import Hyperspace, IndexConfig
    s, tmp = env
    s.conf["spark.hyperspace.index.numBuckets"] = "8"
    s.conf["spark.hyperspace.index.lineage.enabled"] = "true"
    lt, step = _lineitem_files(tmp, parts=8)
    hs = Hyperspace(s)
    li, od = s.read_parquet(str(tmp / "lineitem")), s.read_parquet(str(tmp / "orders"))
    hs.createIndex(li, IndexConfig("liIndex", ["l_orderkey"], ["l_extendedprice", "l_discount"]))
    hs.createIndex(od, IndexConfig("oIndex", ["o_orderkey"], ["o_totalprice"]))
    os.remove(tmp / "lineitem" / "part-3.parquet")
    pq.write_table(lt.slice(10, 700), tmp / "lineitem" / "part-new.parquet")
    li2 = s.read_parquet(str(tmp / "lineitem"))
    fq = li2.filter("l_orderkey >= 1000").select("l_orderkey", "l_discount")
    jq = li2.join(od, li2["l_orderkey"] == od["o_orderkey"]).select("l_orderkey", "l_extendedprice", "o_totalprice")
    s.enable_hyperspace()
    assert "liIndex" not in fq.explain_string()          # signature mismatch and no hybrid scan: index unused
    if mode == "hybrid_conf":
        s.conf["spark.hyperspace.index.hybridscan.enabled"] = "true"
    else:
        hs.refreshIndex("liIndex", "quick")
        e = hs.index("liIndex")
        upd = e["source"]["plan"]["properties"]["relations"][0]["data"]["properties"]["update"]
        assert upd["appendedFiles"] is not None and upd["deletedFiles"] is not None
        assert sorted(os.listdir(tmp / "indexes" / "liIndex")) == ["_hyperspace_log", "v__=0"]   # no data written
    assert "HybridScan: +1 files -1 files" in fq.explain_string()
    got = _verify_index_usage(s, fq, ["liIndex"])
    assert got.num_rows > 0
    got = _verify_index_usage(s, jq, ["liIndex", "oIndex"])
    assert got.num_rows == lt.num_rows - step + 700
    if mode == "hybrid_conf":        # too much appended data: the index is not a candidate any more
        s.conf["spark.hyperspace.index.hybridscan.maxAppendedRatio"] = "0.01"
        assert "liIndex" not in fq.explain_string()


def test_lifecycle_states(env):
    from hyperspace_1_b200 import Hyperspace, HyperspaceException, IndexConfig
    s, tmp = env
    os.makedirs(tmp / "departments")
    pq.write_table(departments_table(), tmp / "departments" / "part-0.parquet")
    dept = s.read_parquet(str(tmp / "departments"))
    hs = Hyperspace(s)
    hs.createIndex(dept, IndexConfig("deptIndex", ["deptId"], ["deptName"]))
    q = dept.filter("deptId = 30").select("deptName")
    s.enable_hyperspace()
    assert "deptIndex" in q.explain_string()
    hs.deleteIndex("deptIndex")
    assert hs.index("deptIndex")["state"] == "DELETED" and "deptIndex" not in q.explain_string()
    with pytest.raises(HyperspaceException):
        hs.refreshIndex("deptIndex", "full")
    hs.restoreIndex("deptIndex")
    assert hs.index("deptIndex")["state"] == "ACTIVE" and "deptIndex" in q.explain_string()
    hs.deleteIndex("deptIndex")
    hs.vacuumIndex("deptIndex")
    assert hs.index("deptIndex")["state"] == "DOESNOTEXIST"
    assert not os.path.exists(tmp / "indexes" / "deptIndex" / "v__=0")
    hs.createIndex(dept, IndexConfig("deptIndex", ["deptId"], ["deptName"]))   # the name is free again
    assert q.collect()["deptName"].to_pylist() == ["Sales"]
