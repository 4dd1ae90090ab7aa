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


def test_signature_is_the_three_step_formula_of_the_reference():
    """md5Hex(md5Hex(fold over files) + fold over plan nodes): FileBasedSignatureProvider.scala:39-40 applies
    md5Hex once more to the fold of DefaultFileBasedSource.scala:188-210 before IndexSignatureProvider.scala:44-50
    joins it with PlanSignatureProvider.scala:36-43. Values below were worked out step by step with hashlib."""
    import hashlib
    from hyperspace_1_b200 import log_entry as le
    md5 = lambda t: hashlib.md5(t.encode()).hexdigest()
    a = [{"name": "file:/d/f2", "size": 7, "modifiedTime": 6, "id": 1}, {"name": "file:/d/f1", "size": 10, "modifiedTime": 5, "id": 0}]
    fold = md5(md5("" + "10" + "5" + "file:/d/f1") + "7" + "6" + "file:/d/f2")
    assert fold == "e9837d664321adf6da7a1cba5b74975e"
    f = md5(fold)
    assert f == "414c807ad09f419b7dc4d849b3da5a53"
    p = md5("" + "LogicalRelation")
    assert p == "9afbc93584b4d85d5bda90a51959f4c8"
    assert le.source_signature(a) == md5(f + p) == "17c677285177d3645e7a50190eb312c7"
    # a plan of two nodes folds bottom-up; no files -> no signature at all
    assert le.source_signature(a, ("LogicalRelation", "Project")) == md5(f + md5(md5("LogicalRelation") + "Project"))
    assert le.source_signature([]) is None


def test_golden_log_entry_of_the_reference_is_read_and_reproduced(tmp_path):
    """The 'IndexLogEntry spec example' (HT/index/IndexLogEntryTest.scala:83-185) as a fixture: the log manager
    reads it, the helpers recover what the reference's test asserts, and an entry made here has the same shape."""
    from hyperspace_1_b200 import log_entry as le
    import pyarrow as pa
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "index_log_entry_spec.json")))["entry"]
    # written the way an index directory holds it, read back through the log manager
    lm = le.IndexLogManager(str(tmp_path / "idx"))
    os.makedirs(tmp_path / "idx" / "_hyperspace_log")
    with open(tmp_path / "idx" / "_hyperspace_log" / "0", "w") as fh:
        fh.write(le.to_json(gold))
    got = lm.get_log(0)
    assert got == gold and lm.get_latest_stable_log()["state"] == "ACTIVE"
    rel = got["source"]["plan"]["properties"]["relations"][0]
    files = le.content_files(rel["data"]["properties"]["content"])
    assert [f["name"] for f in files] == ["test/f1", "test/f2"] and [f["id"] for f in files] == [0, 1]
    assert sum(f["size"] for f in files) == 200                      # sourceFilesSizeInBytes (:184)
    deleted = le.content_files(rel["data"]["properties"]["update"]["deletedFiles"])
    assert [(f["size"], f["modifiedTime"], f["id"]) for f in deleted] == [(10, 10, 2)]
    assert rel["data"]["properties"]["update"]["appendedFiles"] is None
    assert json.loads(got["derivedDataset"]["properties"]["schemaString"])["fields"][0]["name"] == "RGUID"

    # an entry made here: the same keys, in the same order, with the same JSON types, all the way down
    schema = pa.schema([("RGUID", pa.string()), ("Date", pa.string())])
    src = [{"name": "file:/test/f1", "size": 100, "modifiedTime": 100, "id": 0},
           {"name": "file:/test/f2", "size": 100, "modifiedTime": 200, "id": 1}]
    mine = le.make_entry("indexName", ["col1"], ["col2", "col3"], schema, 200, [], ["/rootpath"], src, schema)
    mine["source"]["plan"]["properties"]["relations"][0]["data"]["properties"]["update"] = {
        "deletedFiles": le.content_of([{"name": "file:/f1", "size": 10, "modifiedTime": 10, "id": 2}]), "appendedFiles": None}

    def shape(x, path=""):
        if isinstance(x, dict):
            return {k: shape(v, path + "/" + k) for k, v in x.items()}
        if isinstance(x, list):
            return [shape(x[0], path + "[]")] if x else []
        return type(x).__name__
    gs, ms = shape(gold), shape(mine)
    # directory trees differ in depth (ours start at the filesystem root): compare them apart from subDirs
    def strip(d):
        if isinstance(d, dict):
            return {k: strip(v) for k, v in d.items() if k != "subDirs"}
        if isinstance(d, list):
            return [strip(v) for v in d]
        return d
    gs, ms = strip(gs), strip(ms)
    ms["content"]["root"]["files"] = gs["content"]["root"]["files"]
    ms["source"]["plan"]["properties"]["relations"][0]["data"]["properties"]["content"]["root"]["files"] = \
        gs["source"]["plan"]["properties"]["relations"][0]["data"]["properties"]["content"]["root"]["files"]
    ms["source"]["plan"]["properties"]["relations"][0]["data"]["properties"]["update"]["deletedFiles"]["root"]["files"] = \
        gs["source"]["plan"]["properties"]["relations"][0]["data"]["properties"]["update"]["deletedFiles"]["root"]["files"]
    assert list(gs.keys()) == list(ms.keys())
    assert json.dumps(gs) == json.dumps(ms)
    assert mine["derivedDataset"]["properties"]["schemaString"] == gold["derivedDataset"]["properties"]["schemaString"]


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


def test_hybrid_scan_candidates_and_quick_refresh_metadata(tmp_path, oracle):
    """Plan-level behaviour of hybrid scan (RuleUtils.scala:84-147, 236-262) without touching a GPU: an index
    whose source gained and lost files is a candidate only with the hybrid-scan switch, only below the
    appended/deleted ratios, and deleted files need lineage; refreshIndex(name, "quick") records the update in
    the log entry and makes the index applicable again."""
    from hyperspace_1_b200 import Hyperspace, Session
    from hyperspace_1_b200.dataframe import IndexScan
    from hyperspace_1_b200 import log_entry as le
    os.makedirs(tmp_path / "t")
    base = pa.table({"k": pa.array(np.arange(4000)), "v": pa.array(np.arange(4000) * 2.0)})
    for i in range(8):
        pq.write_table(base.slice(i * 500, 500), tmp_path / "t" / f"part-{i}.parquet")
    s = Session(system_path=str(tmp_path / "indexes"))
    df = s.read_parquet(str(tmp_path / "t"))
    _fabricate_index(oracle, s, "kIndex", df.plan.relation, ["k"], ["v"], 8)
    s.enable_hyperspace()
    q = lambda d: d.filter("k >= 10").select("k", "v")
    assert "kIndex" in q(df).explain_string()
    pq.write_table(base.slice(0, 300), tmp_path / "t" / "part-new.parquet")
    df2 = s.read_parquet(str(tmp_path / "t"))
    assert "kIndex" not in q(df2).explain_string()                        # signature mismatch
    s.conf["spark.hyperspace.index.hybridscan.enabled"] = "true"
    plan = q(df2).optimized_plan()
    scan = plan.child.child
    assert isinstance(scan, IndexScan) and [os.path.basename(f) for f in scan.hybrid["appended"]] == ["part-new.parquet"]
    assert scan.hybrid["deleted_ids"] == [] and "HybridScan: +1 files -0 files" in q(df2).explain_string()
    s.conf["spark.hyperspace.index.hybridscan.maxAppendedRatio"] = "0.01"
    assert "kIndex" not in q(df2).explain_string()                        # too much appended data
    s.conf["spark.hyperspace.index.hybridscan.maxAppendedRatio"] = "0.3"
    os.remove(tmp_path / "t" / "part-3.parquet")
    df3 = s.read_parquet(str(tmp_path / "t"))
    assert "kIndex" not in q(df3).explain_string()                        # deleted file, index without lineage
    hs = Hyperspace(s)
    with pytest.raises(Exception):
        hs.refreshIndex("kIndex", "quick")                                # same rule for the quick refresh
    # an index WITH lineage: deletes are fine, and a quick refresh stores the update lists
    e = hs.index("kIndex")
    e = dict(e, id=2)
    e["derivedDataset"]["properties"]["properties"]["lineage"] = "true"
    lm = le.IndexLogManager(str(tmp_path / "indexes" / "kIndex"))
    assert lm.write_log(2, e)
    lm.delete_latest_stable_log()
    lm.create_latest_stable_log(2)
    plan = q(df3).optimized_plan()
    assert len(plan.child.child.hybrid["deleted_ids"]) == 1
    s.conf["spark.hyperspace.index.hybridscan.enabled"] = "false"
    assert "kIndex" not in q(df3).explain_string()
    hs.refreshIndex("kIndex", "quick")
    e2 = hs.index("kIndex")
    assert e2["state"] == "ACTIVE" and e2["id"] == 4
    upd = e2["source"]["plan"]["properties"]["relations"][0]["data"]["properties"]["update"]
    assert [os.path.basename(f["name"]) for f in le.content_files(upd["appendedFiles"])] == ["part-new.parquet"]
    assert [os.path.basename(f["name"]) for f in le.content_files(upd["deletedFiles"])] == ["part-3.parquet"]
    assert e2["content"] == e["content"]                                   # metadata only: same index files
    assert "HybridScan: +1 files -1 files" in q(df3).explain_string()      # applicable again, without the switch


def test_parquet_framing_reads_back_with_pyarrow(tmp_path):
    """parquet_plain.py's Thrift framing with page bodies encoded by hand (numpy) instead of by the device:
    PLAIN int64 / double / byte-array values and RLE or bit-packed definition levels, two pages per column.
    pyarrow must read back the rows; this pins the container format on the CPU side."""
    import struct
    from hyperspace_1_b200 import log_entry as le, parquet_plain as pp

    def def_levels(valid):
        n = len(valid)
        if valid.all():
            body = bytes([n << 1, 1]) if n < 64 else None
            if body is None:
                t = pp._Thrift()
                t._varint(n << 1)
                body = bytes(t.b) + b"\x01"
        else:
            groups = (n + 7) // 8
            t = pp._Thrift()
            t._varint((groups << 1) | 1)
            body = bytes(t.b) + np.packbits(np.pad(valid, (0, groups * 8 - n)), bitorder="little").tobytes()
        return struct.pack("<I", len(body)) + body

    cols = {"a": pa.array([1, None, 3, 4, 5, None, 7], type=pa.int64()),
            "b": pa.array([0.5, 1.5, 2.5, 3.5, 4.5, 5.5, 6.5]),
            "s": pa.array(["x", "", None, "hello", "wörld", "z", None])}
    names, types = list(cols), [c.type for c in cols.values()]
    pages = [(0, 4), (4, 7)]
    path = tmp_path / "t.parquet"
    with open(path, "wb") as f:
        f.write(pp.MAGIC)
        pos, chunks = 4, []
        for name, arr in cols.items():
            start = pos
            for a, e in pages:
                part = arr.slice(a, e - a)
                valid = np.array([v is not None for v in part.to_pylist()])
                vals = [v for v in part.to_pylist() if v is not None]
                if pa.types.is_int64(arr.type):
                    values = np.array(vals, dtype="<i8").tobytes()
                elif pa.types.is_float64(arr.type):
                    values = np.array(vals, dtype="<f8").tobytes()
                else:
                    values = b"".join(struct.pack("<I", len(v.encode())) + v.encode() for v in vals)
                body = def_levels(valid) + values
                hdr = pp.page_header(e - a, len(body))
                f.write(hdr + body)
                pos += len(hdr) + len(body)
            chunks.append((start, pos - start, len(arr), None))
        foot = pp.footer(names, types, 7, chunks, le.schema_json(pa.schema([pa.field(n, t) for n, t in zip(names, types)])))
        f.write(foot + struct.pack("<I", len(foot)) + pp.MAGIC)
    got = pq.ParquetFile(path).read()
    assert got.column_names == names
    for n in names:
        assert got[n].to_pylist() == cols[n].to_pylist()


@pytest.mark.parametrize("snappy", [False, True])
def test_dictionary_page_framing_is_read_by_pyarrow(tmp_path, snappy):
    """The layout hs_parquet_encode_pages produces for a dictionary-encoded chunk, assembled here by hand: a
    PLAIN dictionary page, then data pages = definition levels ++ bit width ++ ONE bit-packed run of the indices
    (last group zero-padded). pyarrow must read the rows back; this pins the framing (headers, footer offsets)
    without a GPU. snappy: the same bodies compressed (here by pyarrow's codec, on the GPU by k_snappy_frag), with
    the uncompressed sizes in the page headers and the codec in the footer, as codec = 1 of
    hs_parquet_encode_pages is framed."""
    import struct
    import pyarrow.parquet as pq
    from hyperspace_1_b200 import log_entry as le, parquet_plain as pp

    def varint(v):
        out = bytearray()
        while v >= 0x80:
            out.append((v & 0x7F) | 0x80)
            v >>= 7
        out.append(v)
        return bytes(out)

    def def_levels(valid):
        groups = (len(valid) + 7) // 8
        body = varint((groups << 1) | 1) + np.packbits(np.asarray(valid, dtype=np.uint8), bitorder="little").tobytes()
        return struct.pack("<I", len(body)) + body

    def packed(idx, bw):
        bits = np.zeros(((len(idx) + 7) // 8) * 8 * bw, dtype=np.uint8)
        for j, x in enumerate(idx):
            for k in range(bw):
                bits[j * bw + k] = (x >> k) & 1
        return bytes([bw]) + varint((((len(idx) + 7) // 8) << 1) | 1) + np.packbits(bits, bitorder="little").tobytes()

    rng = np.random.default_rng(5)
    dictionary = np.array([-7, 3, 2**40, 11, 0], dtype="<i8")
    idx = rng.integers(0, len(dictionary), size=29)
    valid = rng.random(29) < 0.8
    values = [int(dictionary[i]) if v else None for i, v in zip(idx, valid)]
    pages = [(0, 13), (13, 29)]
    path = tmp_path / "d.parquet"
    squeeze = (lambda b: pa.Codec("snappy").compress(b, asbytes=True)) if snappy else (lambda b: b)
    with open(path, "wb") as f:
        f.write(pp.MAGIC)
        dbody = dictionary.tobytes()
        stored = squeeze(dbody)
        hdr = pp.dictionary_page_header(len(dictionary), len(stored), len(dbody))
        f.write(hdr + stored)
        pos = 4 + len(hdr) + len(stored)
        upos = len(hdr) + len(dbody)
        data_off = pos
        for a, e in pages:
            body = def_levels(valid[a:e]) + packed([int(i) for i, v in zip(idx[a:e], valid[a:e]) if v], 3)
            stored = squeeze(body)
            hdr = pp.page_header(e - a, len(stored), pp.ENC_PLAIN_DICTIONARY, len(body))
            f.write(hdr + stored)
            pos += len(hdr) + len(stored)
            upos += len(hdr) + len(body)
        names, types = ["k"], [pa.int64()]
        foot = pp.footer(names, types, 29, [(4, pos - 4, 29, data_off, upos)],
                         le.schema_json(pa.schema([pa.field("k", pa.int64())])), 1 if snappy else 0)
        f.write(foot + struct.pack("<I", len(foot)) + pp.MAGIC)
    pf = pq.ParquetFile(path)
    col = pf.metadata.row_group(0).column(0)
    assert col.compression == ("SNAPPY" if snappy else "UNCOMPRESSED")
    assert col.total_compressed_size == pos - 4 and col.total_uncompressed_size == upos
    assert col.has_dictionary_page and col.dictionary_page_offset == 4 and col.data_page_offset == data_off
    assert pf.read()["k"].to_pylist() == values


def test_refresh_and_optimize_decisions_without_a_gpu(tmp_path, oracle):
    """validate() of the refresh and optimize actions (the parts that decide WHAT to compute) on an index
    fabricated by the oracle: which source files count as appended / deleted (RefreshActionBase.scala:112-147),
    no-change refreshes are logged no-ops (NoChangesException), deletes need lineage, and optimize picks the
    files of buckets holding more than one file, below the size threshold in quick mode
    (OptimizeAction.scala:115-133)."""
    import shutil
    from hyperspace_1_b200 import Hyperspace, HyperspaceException, Session, hyperspace as hsmod, index_writer
    from hyperspace_1_b200 import log_entry as le
    os.makedirs(tmp_path / "t")
    base = pa.table({"k": pa.array(np.arange(3000)), "v": pa.array(np.arange(3000) * 1.5)})
    for i in range(3):
        pq.write_table(base.slice(i * 1000, 1000), tmp_path / "t" / f"part-{i}.parquet")
    s = Session(system_path=str(tmp_path / "indexes"))
    df = s.read_parquet(str(tmp_path / "t"))
    _fabricate_index(oracle, s, "kIndex", df.plan.relation, ["k"], ["v"], 4)
    hs = Hyperspace(s)
    path = hs._index_path("kIndex")
    # nothing changed: every mode is a no-op that leaves the log where it is
    for mode in ("full", "incremental", "quick"):
        a = hsmod.RefreshAction(s, path, mode)
        with pytest.raises(hsmod.NoChangesException):
            a.validate()
        assert hs.refreshIndex("kIndex", mode) is not None and hs.index("kIndex")["id"] == 1
    with pytest.raises(HyperspaceException):
        hs.refreshIndex("kIndex", "sideways")
    # one file appended, one deleted
    pq.write_table(base.slice(0, 10), tmp_path / "t" / "part-9.parquet")
    os.remove(tmp_path / "t" / "part-1.parquet")
    a = hsmod.RefreshAction(s, path, "full")
    a.validate()
    assert [os.path.basename(f) for f in a.appended] == ["part-9.parquet"]
    assert [os.path.basename(f) for f in a.deleted] == ["part-1.parquet"]
    assert a.num_buckets == 4 and a.indexed == ["k"] and a.included == ["v"] and not a.lineage
    for mode in ("incremental", "quick"):
        with pytest.raises(HyperspaceException, match="lineage"):
            hsmod.RefreshAction(s, path, mode).validate()
    # optimize: a second file in buckets 0 and 2 (copies of the first), nothing extra in 1 and 3
    e = hs.index("kIndex")
    files = sorted(le.from_uri(f["name"]) for f in le.content_files(e["content"]))
    by_bucket = {le.bucket_id_of(f): f for f in files}
    extra = []
    os.makedirs(os.path.join(path, "v__=1"))
    for b in (0, 2):
        dst = os.path.join(path, "v__=1", os.path.basename(by_bucket[b]).replace("part-", "part-x"))
        shutil.copy(by_bucket[b], dst)
        extra.append(dst)
    tracker = le.FileIdTracker()
    e2 = dict(e, id=2)
    e2["content"] = le.content_of([le.file_info(f, tracker) for f in files + extra])
    lm = le.IndexLogManager(path)
    assert lm.write_log(2, e2)
    lm.delete_latest_stable_log()
    lm.create_latest_stable_log(2)
    o = hsmod.OptimizeAction(s, path, "full")
    o.validate()
    assert sorted(le.bucket_id_of(f["name"]) for f in o.to_optimize) == [0, 0, 2, 2]
    assert sorted(le.bucket_id_of(f["name"]) for f in o.to_ignore) == sorted(set(by_bucket) - {0, 2})
    s.conf["spark.hyperspace.index.optimize.fileSizeThreshold"] = "1"      # every file is "large": nothing to do
    with pytest.raises(hsmod.NoChangesException):
        hsmod.OptimizeAction(s, path, "quick").validate()
    with pytest.raises(HyperspaceException):
        hsmod.OptimizeAction(s, path, "sometimes").validate()
    # the files of the fabricated index group into one run; with the copies, two runs for buckets 0 and 2
    runs, names, _ = index_writer.read_index(e2)
    assert names == ["k", "v"] and len(runs) == 2
    assert (np.diff(runs[1][1]) > 0).tolist() == [b in (0, 2) for b in range(4)]
