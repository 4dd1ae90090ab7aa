"""The metadata that makes a directory of bucketed Parquet files an index: the IndexLogEntry JSON
(HS/index/IndexLogEntry.scala:433-603, golden example HT/index/IndexLogEntryTest.scala:83-185), the
operation log with optimistic concurrency (HS/index/IndexLogManager.scala:33-165) and the versioned data
directories (HS/index/IndexDataManager.scala:38-71). Field names and nesting are the reference's, so that
Jackson on the reference side can read entries written here."""
import hashlib
import json
import os
import re
import time

HYPERSPACE_LOG = "_hyperspace_log"           # IndexConstants.scala:61
INDEX_VERSION_DIRECTORY_PREFIX = "v__"       # IndexConstants.scala:62
LATEST_STABLE = "latestStable"
HYPERSPACE_VERSION = "0.1"                   # IndexLogEntry.scala:609
DATA_FILE_NAME_ID = "_data_file_id"
UNKNOWN_FILE_ID = -1


class States:  # HS/actions/Constants.scala:20-33
    ACTIVE, CREATING, DELETING, DELETED = "ACTIVE", "CREATING", "DELETING", "DELETED"
    REFRESHING, VACUUMING, RESTORING, OPTIMIZING = "REFRESHING", "VACUUMING", "RESTORING", "OPTIMIZING"
    DOESNOTEXIST, CANCELLING = "DOESNOTEXIST", "CANCELLING"
    STABLE = {"ACTIVE", "DELETED", "DOESNOTEXIST"}


def md5_hex(s):
    return hashlib.md5(str(s).encode("utf-8")).hexdigest()  # HS/util/HashingUtils.scala:32-34


def to_uri(path):
    """Hadoop-style path string of a local file (file:/abs/path)."""
    return "file:" + os.path.abspath(path)


def from_uri(uri):
    return uri[5:] if uri.startswith("file:") else uri


class FileIdTracker:
    """Unique int64 id per (path, size, mtime), starting at 0 (IndexLogEntry.scala:617-686)."""

    def __init__(self):
        self.max_id = -1
        self.ids = {}

    def add_file(self, path, size, mtime):
        key = (path, size, mtime)
        if key not in self.ids:
            self.max_id += 1
            self.ids[key] = self.max_id
        return self.ids[key]

    def add_known(self, path, size, mtime, fid):
        self.ids[(path, size, mtime)] = fid
        self.max_id = max(self.max_id, fid)


def file_info(path, tracker):
    st = os.stat(path)
    uri = to_uri(path)
    size, mtime = st.st_size, int(st.st_mtime * 1000)
    fid = tracker.add_file(uri, size, mtime) if tracker is not None else UNKNOWN_FILE_ID
    return {"name": uri, "size": size, "modifiedTime": mtime, "id": fid}


def directory_tree(file_infos):
    """Content.root: a directory tree from the filesystem root down to the files
    (IndexLogEntry.scala:235-291; root name like 'file:/'). file_infos carry full 'file:/...' names."""
    root = {"name": "file:/", "files": [], "subDirs": []}
    for fi in file_infos:
        parts = [p for p in from_uri(fi["name"]).split("/") if p]
        node = root
        for d in parts[:-1]:
            nxt = next((s for s in node["subDirs"] if s["name"] == d), None)
            if nxt is None:
                nxt = {"name": d, "files": [], "subDirs": []}
                node["subDirs"].append(nxt)
            node = nxt
        node["files"].append({"name": parts[-1], "size": fi["size"], "modifiedTime": fi["modifiedTime"], "id": fi["id"]})
    return root


def content_of(file_infos):
    return {"root": directory_tree(file_infos), "fingerprint": {"kind": "NoOp", "properties": {}}}


def content_files(content):
    """Content.fileInfos: flatten the tree back to full-path FileInfos (IndexLogEntry.scala:43-113)."""
    out = []

    def walk(node, prefix):
        here = prefix + node["name"] if prefix == "" else (prefix.rstrip("/") + "/" + node["name"])
        for f in node["files"]:
            out.append({"name": here.rstrip("/") + "/" + f["name"], "size": f["size"],
                        "modifiedTime": f["modifiedTime"], "id": f["id"]})
        for s in node["subDirs"]:
            walk(s, here)
    walk(content["root"], "")
    return out


def arrow_to_spark_type(t):
    import pyarrow as pa
    if pa.types.is_boolean(t):
        return "boolean"
    if pa.types.is_int8(t):
        return "byte"
    if pa.types.is_int16(t):
        return "short"
    if pa.types.is_int32(t):
        return "integer"
    if pa.types.is_int64(t):
        return "long"
    if pa.types.is_float32(t):
        return "float"
    if pa.types.is_float64(t):
        return "double"
    if pa.types.is_string(t) or pa.types.is_large_string(t):
        return "string"
    if pa.types.is_binary(t) or pa.types.is_large_binary(t):
        return "binary"
    if pa.types.is_date32(t):
        return "date"
    if pa.types.is_timestamp(t):
        return "timestamp"
    raise TypeError(f"no Spark type for {t}")


def schema_json(schema):
    """Spark StructType.json of an Arrow schema (what CoveringIndex.Properties.schemaString holds)."""
    return json.dumps({"type": "struct", "fields": [
        {"name": f.name, "type": arrow_to_spark_type(f.type), "nullable": True, "metadata": {}} for f in schema]},
        separators=(",", ":"))


def source_signature(source_file_infos, plan_node_names=("LogicalRelation",)):
    """The value RuleUtils.getCandidateIndexes compares (HS/index/rules/RuleUtils.scala:61-76):
        fold  = foldLeft("") over the files sorted by path of md5Hex(acc + len + mtime + path)
                                                      (sources/default/DefaultFileBasedSource.scala:188-210)
        f     = md5Hex(fold)                          (FileBasedSignatureProvider.scala:39-40; None when fold is "")
        p     = fold over the plan's node names bottom-up of md5Hex(acc + nodeName)
                                                      (PlanSignatureProvider.scala:36-43)
        value = md5Hex(f + p)                         (IndexSignatureProvider.scala:44-50)
    Returns None when the relation has no files (the reference then provides no signature)."""
    fold = ""
    for fi in sorted(source_file_infos, key=lambda f: f["name"]):
        fold = md5_hex(fold + str(fi["size"]) + str(fi["modifiedTime"]) + fi["name"])
    if fold == "":
        return None
    fsig = md5_hex(fold)
    psig = ""
    for n in plan_node_names:
        psig = md5_hex(psig + n)
    if psig == "":
        return None
    return md5_hex(fsig + psig)


SIGNATURE_PROVIDER = "com.microsoft.hyperspace.index.IndexSignatureProvider"


def make_entry(name, indexed, included, index_schema, num_buckets, index_files, source_root_paths,
               source_files, source_schema, file_format="parquet", properties=None, state=States.ACTIVE, entry_id=0):
    """IndexLogEntry as a JSON-ready dict, field for field as the reference's case classes."""
    relation = {
        "rootPaths": [to_uri(p) for p in source_root_paths],
        "data": {"properties": {"content": content_of(source_files), "update": None}, "kind": "HDFS"},
        "dataSchemaJson": schema_json(source_schema),
        "fileFormat": file_format,
        "options": {},
    }
    return {
        "name": name,
        "derivedDataset": {
            "properties": {"columns": {"indexed": list(indexed), "included": list(included)},
                           "schemaString": schema_json(index_schema), "numBuckets": int(num_buckets),
                           "properties": dict(properties or {})},
            "kind": "CoveringIndex"},
        "content": content_of(index_files),
        "source": {"plan": {"properties": {
            "relations": [relation], "rawPlan": None, "sql": None,
            "fingerprint": {"properties": {"signatures": [
                {"provider": SIGNATURE_PROVIDER, "value": source_signature(source_files)}]}, "kind": "LogicalPlan"}},
            "kind": "Spark"}},
        "properties": {},
        "version": HYPERSPACE_VERSION,
        "id": entry_id,
        "state": state,
        "timestamp": int(time.time() * 1000),
        "enabled": True,
    }


def to_json(entry):
    return json.dumps(entry, indent=2, separators=(",", " : "))  # Jackson's pretty printer uses ' : '


_BUCKET_RE = re.compile(r".*_(\d+)(?:\..*)?$")


def bucket_id_of(file_name):
    """Spark's BucketingUtils.getBucketId; used by OptimizeAction.scala:129."""
    m = _BUCKET_RE.match(os.path.basename(file_name))
    return int(m.group(1)) if m else None


class IndexLogManager:
    """Numbered immutable JSON entries + a latestStable copy; a writer loses if the id already exists
    (temp file + atomic no-replace link), HS/index/IndexLogManager.scala:149-165."""

    def __init__(self, index_path):
        self.index_path = index_path
        self.log_path = os.path.join(index_path, HYPERSPACE_LOG)

    def _path(self, entry_id):
        return os.path.join(self.log_path, str(entry_id))

    def get_log(self, entry_id):
        p = self._path(entry_id)
        if not os.path.exists(p):
            return None
        with open(p) as f:
            e = json.load(f)
        if e.get("version") != HYPERSPACE_VERSION:   # LogEntry.scala:40-45
            raise ValueError(f"Unsupported log entry version {e.get('version')}")
        return e

    def get_latest_id(self):
        if not os.path.isdir(self.log_path):
            return None
        ids = [int(n) for n in os.listdir(self.log_path) if n.isdigit()]
        return max(ids) if ids else None

    def get_latest_log(self):
        i = self.get_latest_id()
        return None if i is None else self.get_log(i)

    def get_latest_stable_log(self):
        p = os.path.join(self.log_path, LATEST_STABLE)
        if os.path.exists(p):
            with open(p) as f:
                e = json.load(f)
            assert e["state"] in States.STABLE
            return e
        i = self.get_latest_id()
        while i is not None and i >= 0:
            e = self.get_log(i)
            if e is not None and e["state"] in States.STABLE:
                return e
            i -= 1
        return None

    def create_latest_stable_log(self, entry_id):
        e = self.get_log(entry_id)
        if e is None or e["state"] not in States.STABLE:
            return False
        with open(self._path(entry_id), "rb") as src, open(os.path.join(self.log_path, LATEST_STABLE), "wb") as dst:
            dst.write(src.read())
        return True

    def delete_latest_stable_log(self):
        p = os.path.join(self.log_path, LATEST_STABLE)
        if os.path.exists(p):
            os.remove(p)
        return True

    def write_log(self, entry_id, entry):
        os.makedirs(self.log_path, exist_ok=True)
        final = self._path(entry_id)
        if os.path.exists(final):
            return False
        tmp = os.path.join(self.log_path, f".temp{os.getpid()}_{time.time_ns()}")
        with open(tmp, "w") as f:
            f.write(to_json(entry))
        try:
            os.link(tmp, final)        # atomic, fails if `final` exists: the optimistic-concurrency check
            return True
        except FileExistsError:
            return False
        finally:
            os.remove(tmp)


class IndexDataManager:
    """<index>/v__=<n>/ directories (HS/index/IndexDataManager.scala:56-71)."""

    def __init__(self, index_path):
        self.index_path = index_path

    def get_latest_version_id(self):
        if not os.path.isdir(self.index_path):
            return None
        pre = INDEX_VERSION_DIRECTORY_PREFIX + "="
        ids = [int(n[len(pre):]) for n in os.listdir(self.index_path) if n.startswith(pre)]
        return max(ids) if ids else None

    def get_path(self, version):
        return os.path.join(self.index_path, f"{INDEX_VERSION_DIRECTORY_PREFIX}={version}")

    def delete(self, version):
        import shutil
        shutil.rmtree(self.get_path(version), ignore_errors=True)
