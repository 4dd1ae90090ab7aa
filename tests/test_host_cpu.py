"""Host-side logic that needs no GPU: IndexConfig validation (mirrors HT/index/IndexConfigTest), the log entry
JSON (field set of the golden example HT/index/IndexLogEntryTest.scala:83-185), the operation log's optimistic
concurrency (HT/index/IndexLogManagerImplTest), path/bucket-id conventions, and the plan rewrites of the index
rules over an index whose files are produced here with the ORACLE (so that no GPU is involved)."""
import json
import os

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

from fixtures import departments_table, employees_table


def test_index_config_validation():
    from hyperspace_1_b200 import IllegalArgumentException, IndexConfig
    with pytest.raises(IllegalArgumentException):
        IndexConfig("", ["a"], [])
    with pytest.raises(IllegalArgumentException):
        IndexConfig("x", [], ["a"])
    with pytest.raises(IllegalArgumentException):
        IndexConfig("x", ["a", "A"], [])
    with pytest.raises(IllegalArgumentException):
        IndexConfig("x", ["a"], ["b", "B"])
    with pytest.raises(IllegalArgumentException):
        IndexConfig("x", ["a"], ["A"])
    assert IndexConfig("idx", ["A", "b"], ["c", "D"]) == IndexConfig("IDX", ["a", "B"], ["d", "C"])
    assert IndexConfig("idx", ["a", "b"], []) != IndexConfig("idx", ["b", "a"], [])
    assert "indexedColumns: a, b" in repr(IndexConfig("idx", ["A", "b"], ["c"]))


def test_log_entry_has_the_reference_field_set(tmp_path):
    from hyperspace_1_b200 import log_entry as le
    src = tmp_path / "src"
    src.mkdir()
    pq.write_table(departments_table(), src / "part-0.parquet")
    from hyperspace_1_b200.dataframe import Relation
    rel = Relation(str(src))
    tracker = le.FileIdTracker()
    e = le.make_entry("deptIndex", ["deptId"], ["deptName"], pa.schema([rel.schema.field("deptId"), rel.schema.field("deptName")]),
                      200, [], [rel.root], rel.file_infos(tracker), rel.schema)
    # the golden example's structure (IndexLogEntryTest.scala:83-185)
    assert list(e.keys()) == ["name", "derivedDataset", "content", "source", "properties", "version", "id", "state",
                              "timestamp", "enabled"]
    assert e["derivedDataset"]["kind"] == "CoveringIndex"
    assert list(e["derivedDataset"]["properties"].keys()) == ["columns", "schemaString", "numBuckets", "properties"]
    assert e["content"]["fingerprint"] == {"kind": "NoOp", "properties": {}}
    plan = e["source"]["plan"]
    assert plan["kind"] == "Spark" and plan["properties"]["rawPlan"] is None and plan["properties"]["sql"] is None
    r = plan["properties"]["relations"][0]
    assert list(r.keys()) == ["rootPaths", "data", "dataSchemaJson", "fileFormat", "options"]
    assert r["data"]["kind"] == "HDFS" and r["rootPaths"][0].startswith("file:/")
    assert plan["properties"]["fingerprint"]["kind"] == "LogicalPlan"
    sig = plan["properties"]["fingerprint"]["properties"]["signatures"][0]
    assert sig["provider"] == "com.microsoft.hyperspace.index.IndexSignatureProvider" and len(sig["value"]) == 32
    assert e["version"] == "0.1" and e["enabled"] is True
    schema = json.loads(e["derivedDataset"]["properties"]["schemaString"])
    assert schema == {"type": "struct", "fields": [
        {"name": "deptId", "type": "integer", "nullable": True, "metadata": {}},
        {"name": "deptName", "type": "string", "nullable": True, "metadata": {}}]}
    # Jackson-style pretty printing and a lossless round trip
    txt = le.to_json(e)
    assert '"name" : "deptIndex"' in txt and json.loads(txt) == e
    # content tree from the filesystem root, flattened back to full paths
    files = le.content_files(r["data"]["properties"]["content"])
    assert [f["name"] for f in files] == [le.to_uri(str(src / "part-0.parquet"))]
    assert files[0]["id"] == 0
    assert r["data"]["properties"]["content"]["root"]["name"] == "file:/"


def test_signature_changes_when_a_file_changes(tmp_path):
    from hyperspace_1_b200 import log_entry as le
    a = [{"name": "file:/d/f1", "size": 10, "modifiedTime": 5, "id": 0}, {"name": "file:/d/f2", "size": 7, "modifiedTime": 6, "id": 1}]
    assert le.source_signature(a) == le.source_signature(list(reversed(a)))   # sorted by path
    b = [dict(a[0]), dict(a[1], size=8)]
    assert le.source_signature(a) != le.source_signature(b)
    assert le.md5_hex("") == "d41d8cd98f00b204e9800998ecf8427e"   # HashingUtilsTest


def test_log_manager_optimistic_concurrency(tmp_path):
    from hyperspace_1_b200 import log_entry as le
    lm = le.IndexLogManager(str(tmp_path / "idx"))
    assert lm.get_latest_id() is None and lm.get_latest_stable_log() is None
    e = {"version": "0.1", "id": 0, "state": le.States.CREATING, "name": "x"}
    assert lm.write_log(0, e)
    assert not lm.write_log(0, e)          # the id exists: the second writer loses
    assert lm.get_latest_stable_log() is None
    assert lm.write_log(1, dict(e, id=1, state=le.States.ACTIVE))
    assert lm.get_latest_id() == 1
    assert lm.get_latest_stable_log()["id"] == 1     # found by walking back, no latestStable yet
    assert lm.create_latest_stable_log(1) and not lm.create_latest_stable_log(0)
    assert os.path.exists(tmp_path / "idx" / "_hyperspace_log" / "latestStable")
    assert lm.write_log(2, dict(e, id=2, state=le.States.REFRESHING))
    assert lm.get_latest_log()["state"] == le.States.REFRESHING and lm.get_latest_stable_log()["id"] == 1
    with open(tmp_path / "idx" / "_hyperspace_log" / "3", "w") as f:
        json.dump(dict(e, version="9.9"), f)
    with pytest.raises(ValueError):
        lm.get_log(3)


def test_bucket_id_from_file_name_and_version_dirs(tmp_path):
    from hyperspace_1_b200 import log_entry as le
    assert le.bucket_id_of("part-00053-7c1b3e52-2d6f-4f6d-9c1e-aaaaaaaaaaaa_00053.c000.snappy.parquet") == 53
    assert le.bucket_id_of("/x/y/part-00000-uuid_00199.c000.snappy.parquet") == 199
    assert le.bucket_id_of("part-00000-uuid.parquet") is None
    dm = le.IndexDataManager(str(tmp_path))
    assert dm.get_latest_version_id() is None
    os.makedirs(dm.get_path(0))
    os.makedirs(dm.get_path(3))
    assert dm.get_latest_version_id() == 3 and dm.get_path(4).endswith("v__=4")


def _fabricate_index(oracle, session, name, rel, indexed, included, nb):
    """An index written WITHOUT the GPU: oracle build + the product's file writer + log entry."""
    from hyperspace_1_b200 import index_writer, log_entry as le
    t = rel.read(columns=indexed + included)
    cols = [t[c].combine_chunks() for c in indexed + included]
    perm, off = oracle.build_index(cols, len(indexed), nb)
    arrays = oracle.take(cols, perm)
    path = os.path.join(session.system_path(), name)
    files, _ = index_writer.write_buckets(arrays, indexed + included, off, os.path.join(path, "v__=0"))
    tracker = le.FileIdTracker()
    e = le.make_entry(name, indexed, included, pa.schema([rel.schema.field(c) for c in indexed + included]), nb,
                      [le.file_info(f, tracker) for f in files], [rel.root], rel.file_infos(tracker), rel.schema, entry_id=1)
    lm = le.IndexLogManager(path)
    assert lm.write_log(0, dict(e, id=0, state=le.States.CREATING)) and lm.write_log(1, e)
    lm.create_latest_stable_log(1)
    return files


def test_rules_rewrite_plans_to_index_scans(tmp_path, oracle):
    from hyperspace_1_b200 import Hyperspace, Session
    from hyperspace_1_b200.dataframe import Filter, IndexScan, Join, Project
    os.makedirs(tmp_path / "departments")
    os.makedirs(tmp_path / "employees")
    pq.write_table(departments_table(), tmp_path / "departments" / "part-0.parquet")
    pq.write_table(employees_table(), tmp_path / "employees" / "part-0.parquet")
    s = Session(system_path=str(tmp_path / "indexes"))
    dept, emp = s.read_parquet(str(tmp_path / "departments")), s.read_parquet(str(tmp_path / "employees"))
    files = _fabricate_index(oracle, s, "deptIndex", dept.plan.relation, ["deptId"], ["deptName"], 200)
    _fabricate_index(oracle, s, "empIndex", emp.plan.relation, ["deptId"], ["empName"], 200)
    # quick-start layout: three files, buckets 53 / 122 / 181, task id == bucket id
    assert sorted(os.path.basename(f)[:10] for f in files) == ["part-00053", "part-00122", "part-00181"]
    assert sorted(int(os.path.basename(f).split("_")[1][:5]) for f in files) == [53, 122, 181]
    hs = Hyperspace(s)
    assert sorted(hs.indexes()["name"].to_pylist()) == ["deptIndex", "empIndex"]
    q = dept.filter("deptId = 20").select("deptName")
    assert "IndexScan" not in q.explain_string()          # disabled: untouched
    s.enable_hyperspace()
    plan = q.optimized_plan()
    assert isinstance(plan, Project) and isinstance(plan.child, Filter) and isinstance(plan.child.child, IndexScan)
    assert plan.child.child.use_bucket_spec is False       # FilterIndexRule.scala:59-65
    assert "Hyperspace(Type: CI, Name: deptIndex, LogVersion: 1)" in q.explain_string()
    # a filter that does not touch the first indexed column is not rewritten
    assert "IndexScan" not in dept.filter("deptName = 'Sales'").select("deptName").explain_string()
    # a query needing a column the index does not cover is not rewritten
    assert "IndexScan" not in dept.filter("deptId = 20").select("location").explain_string()
    j = emp.join(dept, emp["deptId"] == dept["deptId"]).select(emp["empName"], dept["deptName"])
    jp = j.optimized_plan()
    assert isinstance(jp, Project) and isinstance(jp.child, Join) and isinstance(jp.child.left, IndexScan)
    assert jp.child.left.use_bucket_spec and jp.child.right.use_bucket_spec
    assert jp.child.left.entry["name"] == "empIndex" and jp.child.right.entry["name"] == "deptIndex"
    # changing the source invalidates the signature: the rule stops firing
    pq.write_table(departments_table(), tmp_path / "departments" / "part-1.parquet")
    dept2 = s.read_parquet(str(tmp_path / "departments"))
    assert "IndexScan" not in dept2.filter("deptId = 20").select("deptName").explain_string()
    # un-indexed execution (the host engine) gives the reference answers
    s.disable_hyperspace()
    assert q.collect()["deptName"].to_pylist() == ["Research"]
    assert j.collect().num_rows == 14


def test_predicate_parser_and_postfix():
    from hyperspace_1_b200.dataframe import Col, expr_to_postfix, parse_predicate
    e = parse_predicate("deptId = 20 AND name <> 'x y' and score >= 1.5")
    assert expr_to_postfix(e, {"deptid": 0, "name": 1, "score": 2}) == [
        ("eq", 0, 20), ("ne", 1, "x y"), ("and",), ("ge", 2, 1.5), ("and",)]
    e = ~((Col("a") > 1) | Col("b").isNull())
    assert expr_to_postfix(e, {"a": 0, "b": 1}) == [("gt", 0, 1), ("is_null", 1), ("or",), ("not",)]
