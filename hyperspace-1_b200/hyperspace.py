"""Hyperspace — the user-facing API of the reference, kept verbatim in names and meaning
(HS/Hyperspace.scala:26-166; python/hyperspace/hyperspace.py:9-192), with the index build routed to the GPU.

Every mutating call is an Action with the reference's two-phase log protocol (HS/actions/Action.scala:83-104):
validate -> write entry id+1 in a transient state -> op() -> write entry id+2 in the final state and refresh
latestStable. The compute inside op() is GpuIndexWriter (hs_build_index / hs_merge_runs)."""
import os
import shutil

import numpy as np
import pyarrow as pa

from . import index_writer, log_entry as le
from .index_writer import LINEAGE_COL
from .dataframe import DataFrame, Relation, Scan, resolve_name
from .index_config import IndexConfig
from .native import HyperspaceException

REFRESH_MODES = ("full", "incremental", "quick")   # IndexConstants.scala:78-80
OPTIMIZE_MODES = ("quick", "full")                 # IndexConstants.scala:88-90


class NoChangesException(Exception):
    """HS/actions/NoChangesException.scala — turns an action into a logged no-op."""


class _Action:
    transient = final = None

    def __init__(self, session, index_path):
        self.session = session
        self.index_path = index_path
        self.log = le.IndexLogManager(index_path)
        self.data = le.IndexDataManager(index_path)
        latest = self.log.get_latest_id()
        self.base_id = -1 if latest is None else latest
        self.stats = {}

    def validate(self):
        pass

    def op(self):
        pass

    def log_entry(self):
        raise NotImplementedError

    def _save(self, entry_id, state):
        e = self.log_entry()
        e["state"], e["id"] = state, entry_id
        if not self.log.write_log(entry_id, e):
            raise HyperspaceException("Could not acquire proper state")

    def run(self):
        try:
            self.validate()
            self._save(self.base_id + 1, self.transient)       # begin()
            self.op()
            if not self.log.delete_latest_stable_log():         # end()
                raise HyperspaceException("Could not delete latest stable log")
            self._save(self.base_id + 2, self.final)
            self.log.create_latest_stable_log(self.base_id + 2)
        except NoChangesException:
            return False
        return True


def _next_version_dir(data):
    v = data.get_latest_version_id()
    return data.get_path(0 if v is None else v + 1)   # CreateActionBase.scala:33-38


class _CreateBase(_Action):
    """Shared by create and refresh: the index writer and the log entry built from what was written."""

    def _resolve(self, relation, config):
        names = relation.schema.names
        indexed = [resolve_name(c, names) for c in config.indexedColumns]
        included = [resolve_name(c, names) for c in config.includedColumns]
        if None in indexed or None in included:
            raise HyperspaceException("Index config is not applicable to dataframe schema.")   # CreateAction.scala:52-54
        return indexed, included

    def _write(self, relation, files, indexed, included, num_buckets, out_dir, lineage_ids=None):
        """lineage_ids: {source file path: file id} — every row then carries the id of the file it came from in
        a trailing _data_file_id column (CreateActionBase.scala:171-203: the reference joins input_file_name()
        with a broadcast (path, id) table; here the id is attached while each file is read)."""
        if lineage_ids is None:
            table = relation.read(columns=indexed + included, files=files)
            inc = list(included)
        else:
            parts = []
            for f in (relation.files if files is None else files):
                t = relation.read(columns=indexed + included, files=[f])
                parts.append(t.append_column(LINEAGE_COL, pa.array(np.full(t.num_rows, lineage_ids[f], dtype=np.int64))))
            table = pa.concat_tables(parts) if parts else relation.read(columns=indexed + included, files=[]).append_column(
                LINEAGE_COL, pa.array([], type=pa.int64()))
            inc = list(included) + [LINEAGE_COL]
        written, stats = index_writer.build_and_write(self.session, table, indexed, inc, num_buckets, out_dir)
        self.stats = stats
        return written

    def _entry(self, name, indexed, included, relation, num_buckets, index_files, tracker, properties=None):
        properties = dict(properties or {})
        fields = [relation.schema.field(c) for c in indexed + included]
        if str(properties.get("lineage", "false")).lower() == "true":
            fields.append(pa.field(LINEAGE_COL, pa.int64(), nullable=False))
        index_schema = pa.schema(fields)
        src = relation.file_infos(tracker)
        idx = [le.file_info(f, tracker) for f in index_files]
        return le.make_entry(name, indexed, included, index_schema, num_buckets, idx, [relation.root], src,
                             relation.schema, properties=properties)


def _lineage_enabled(session):
    return str(session.conf.get("spark.hyperspace.index.lineage.enabled", "false")).lower() == "true"


class CreateAction(_CreateBase):
    transient, final = le.States.CREATING, le.States.ACTIVE

    def __init__(self, session, df, config, index_path):
        super().__init__(session, index_path)
        self.df, self.config = df, config
        self.files = []
        self.out_dir = _next_version_dir(self.data)

    def validate(self):
        if not isinstance(self.df.plan, Scan):     # CreateAction.scala:46-49
            raise HyperspaceException("Only creating index over HDFS file based scan nodes is supported.")
        self.relation = self.df.plan.relation
        self.indexed, self.included = self._resolve(self.relation, self.config)
        prev = self.log.get_latest_log()
        if prev is not None and prev["state"] != le.States.DOESNOTEXIST:   # :57-63
            raise HyperspaceException(f"Another Index with name {self.config.indexName} already exists")

    def op(self):
        self.tracker = le.FileIdTracker()
        lineage_ids = None
        if _lineage_enabled(self.session):
            # source file ids start at 0 and are the ids the log entry records (IndexLogEntry.scala:617-686)
            lineage_ids = {le.from_uri(f["name"]): f["id"] for f in self.relation.file_infos(self.tracker)}
        self.files = self._write(self.relation, None, self.indexed, self.included, self.session.num_buckets(),
                                 self.out_dir, lineage_ids)

    def log_entry(self):
        tracker = getattr(self, "tracker", None) or le.FileIdTracker()
        props = {"lineage": "true"} if _lineage_enabled(self.session) else {}
        return self._entry(self.config.indexName, self.indexed, self.included, self.relation,
                           self.session.num_buckets(), self.files, tracker, props)


def drop_deleted_rows(session, arrays, n_indexed, boff, names, deleted_ids):
    """Rows of an index run minus those whose _data_file_id is one of deleted_ids: hs_filter_scan with
    NOT (_data_file_id IN (...)) — rows keep their order, so every bucket stays sorted and only its boundaries
    move (RefreshIncrementalAction.scala:73-95 writes the result without re-partitioning, for the same reason)."""
    lin = names.index(LINEAGE_COL)
    out = session.ctx.filter_scan(arrays, n_indexed, boff, [("in", lin, list(deleted_ids)), ("not",)],
                                  list(range(len(names))))
    nb = len(boff) - 1
    if len(out[0]) == 0:
        return out, np.zeros(nb + 1, dtype=np.int64)
    ids = session.ctx.hash_bucket(out[:n_indexed], nb)
    new_off = np.zeros(nb + 1, dtype=np.int64)
    np.cumsum(np.bincount(ids, minlength=nb), out=new_off[1:])
    return out, new_off


class RefreshAction(_CreateBase):
    """mode full: rebuild over the relation as it is now (RefreshAction.scala:41). mode incremental
    (RefreshIncrementalAction.scala:53-96, :129-144): appended source files are indexed into a new version
    directory and the old files are kept, so that buckets end up with several sorted files; if source files were
    deleted (needs the lineage column) the old index files are rewritten without their rows. mode quick
    (RefreshQuickAction.scala): nothing is computed — the entry records which files were appended and deleted and
    queries bridge the gap with hybrid scan."""
    transient, final = le.States.REFRESHING, le.States.ACTIVE

    def __init__(self, session, index_path, mode):
        super().__init__(session, index_path)
        self.mode = mode
        self.prev = self.log.get_latest_log()
        self.out_dir = _next_version_dir(self.data)
        self.new_files = []
        self.rewritten = False

    def validate(self):
        if self.prev is None or self.prev["state"] != le.States.ACTIVE:   # RefreshActionBase.scala:97-103
            raise HyperspaceException(f"Refresh is only supported in {le.States.ACTIVE} state.")
        rel = self.prev["source"]["plan"]["properties"]["relations"][0]
        self.relation = Relation(le.from_uri(rel["rootPaths"][0]))          # df rebuilt from the logged relation
        props = self.prev["derivedDataset"]["properties"]
        self.indexed, self.included = list(props["columns"]["indexed"]), list(props["columns"]["included"])
        self.num_buckets = props["numBuckets"]                               # refresh keeps the original count
        self.lineage = index_writer.has_lineage(self.prev)                   # ... and the lineage flag (:57-65)
        self.logged = le.content_files(rel["data"]["properties"]["content"])
        logged = {(f["name"], f["size"], f["modifiedTime"]) for f in self.logged}
        self.current = self.relation.file_infos()
        current = {(f["name"], f["size"], f["modifiedTime"]) for f in self.current}
        self.appended = sorted(le.from_uri(n) for n, _, _ in current - logged)   # RefreshActionBase.scala:142-147
        self.deleted = sorted(le.from_uri(n) for n, _, _ in logged - current)    # :112-117
        self.deleted_infos = [f for f in self.logged if (f["name"], f["size"], f["modifiedTime"]) not in current]
        self.appended_infos = [f for f in self.current if (f["name"], f["size"], f["modifiedTime"]) not in logged]
        if not self.appended and not self.deleted:
            raise NoChangesException(f"Refresh {self.mode} aborted as no source data change found.")
        if self.mode in ("incremental", "quick") and self.deleted and not self.lineage:
            raise HyperspaceException("Index refresh (to handle deleted source data) is only supported on an "
                                      "index with lineage.")            # RefreshIncrementalAction.scala:102-115

    def _tracker(self):
        """File ids survive a refresh: the lineage column refers to them (IndexLogEntry.scala:617-686)."""
        t = le.FileIdTracker()
        for f in self.logged:
            t.add_known(f["name"], f["size"], f["modifiedTime"], f["id"])
        for f in le.content_files(self.prev["content"]):
            t.add_known(f["name"], f["size"], f["modifiedTime"], f["id"])
        return t

    def op(self):
        self.tracker = self._tracker()
        if self.mode == "quick":
            return
        lineage_ids = None
        if self.lineage:
            lineage_ids = {le.from_uri(f["name"]): f["id"] for f in self.relation.file_infos(self.tracker)}
        if self.mode == "full":
            self.new_files = self._write(self.relation, None, self.indexed, self.included, self.num_buckets,
                                         self.out_dir, lineage_ids)
            return
        if self.deleted:
            runs, names, _ = index_writer.read_index(self.prev, session=self.session)
            ids = [f["id"] for f in self.deleted_infos]
            for arrays, boff in runs:
                kept, koff = drop_deleted_rows(self.session, arrays, len(self.indexed), boff, names, ids)
                files, _ = index_writer.write_index_files(self.session, kept, names, koff, self.out_dir)
                self.new_files += files
            self.rewritten = True
        if self.appended:
            self.new_files += self._write(self.relation, self.appended, self.indexed, self.included, self.num_buckets,
                                          self.out_dir, lineage_ids)

    def log_entry(self):
        tracker = getattr(self, "tracker", None) or self._tracker()
        props = self.prev["derivedDataset"]["properties"]["properties"]
        if self.mode == "quick":
            # same index files; the source relation keeps its logged content and gains the update lists, the
            # fingerprint becomes that of the files as they are now (RefreshQuickAction.scala: logEntry)
            e = self._entry(self.prev["name"], self.indexed, self.included, self.relation, self.num_buckets, [],
                            tracker, props)
            e["content"] = self.prev["content"]
            rel = e["source"]["plan"]["properties"]["relations"][0]
            rel["data"]["properties"]["content"] = self.prev["source"]["plan"]["properties"]["relations"][0]["data"][
                "properties"]["content"]
            rel["data"]["properties"]["update"] = {
                "appendedFiles": le.content_of([dict(f, id=tracker.add_file(f["name"], f["size"], f["modifiedTime"]))
                                                for f in self.appended_infos]) if self.appended_infos else None,
                "deletedFiles": le.content_of(self.deleted_infos) if self.deleted_infos else None}
            return e
        keep = []
        if self.mode == "incremental" and not self.rewritten:
            keep = [le.from_uri(f["name"]) for f in le.content_files(self.prev["content"])]
        return self._entry(self.prev["name"], self.indexed, self.included, self.relation, self.num_buckets,
                           keep + self.new_files, tracker, props)


class OptimizeAction(_CreateBase):
    """Compaction (OptimizeAction.scala:85-155): files of buckets that hold more than one file (quick: only files
    below the size threshold) are merged into one sorted file per bucket by hs_merge_runs."""
    transient, final = le.States.OPTIMIZING, le.States.ACTIVE

    def __init__(self, session, index_path, mode):
        super().__init__(session, index_path)
        self.mode = mode
        self.prev = self.log.get_latest_log()
        self.out_dir = _next_version_dir(self.data)
        self.new_files = []

    def validate(self):
        if self.prev is None or self.prev["state"] != le.States.ACTIVE:
            raise HyperspaceException(f"Optimize is only supported in {le.States.ACTIVE} state.")
        if self.mode.lower() not in OPTIMIZE_MODES:
            raise HyperspaceException(f"Unsupported optimize mode '{self.mode}' found.")
        infos = le.content_files(self.prev["content"])
        if self.mode.lower() == "quick":
            thr = int(self.session.conf["spark.hyperspace.index.optimize.fileSizeThreshold"])
            cand = [f for f in infos if f["size"] < thr]
            large = [f for f in infos if f["size"] >= thr]
        else:
            cand, large = infos, []
        per_bucket = {}
        for f in cand:
            per_bucket.setdefault(le.bucket_id_of(f["name"]), []).append(f)
        self.to_optimize = [f for fl in per_bucket.values() if len(fl) > 1 for f in fl]
        self.to_ignore = [f for fl in per_bucket.values() if len(fl) <= 1 for f in fl] + large
        if not self.to_optimize:
            raise NoChangesException("Optimize aborted as no optimizable index files smaller than "
                                     f"{self.session.conf['spark.hyperspace.index.optimize.fileSizeThreshold']} found.")

    def op(self):
        props = self.prev["derivedDataset"]["properties"]
        indexed = list(props["columns"]["indexed"])
        runs, names, _ = index_writer.read_index(self.prev, files=self.to_optimize, session=self.session)
        arrays, boff = self.session.ctx.merge_runs([(a, len(indexed), off) for a, off in runs])
        self.new_files, _ = index_writer.write_index_files(self.session, arrays, names, boff, self.out_dir)

    def log_entry(self):
        tracker = le.FileIdTracker()
        for f in self.to_ignore:
            tracker.add_known(f["name"], f["size"], f["modifiedTime"], f["id"])
        e = dict(self.prev)
        files = [dict(f) for f in self.to_ignore] + [le.file_info(f, tracker) for f in self.new_files]
        e["content"] = le.content_of(files)
        return e


class _StateOnly(_Action):
    """delete / restore / vacuum / cancel: metadata-only lifecycle steps (HS/actions/DeleteAction.scala etc.)."""

    def __init__(self, session, index_path, required, transient, final, after=None):
        super().__init__(session, index_path)
        self.required, self.transient, self.final, self.after = required, transient, final, after
        self.prev = self.log.get_latest_log()

    def validate(self):
        if self.prev is None or (self.required and self.prev["state"] not in self.required):
            raise HyperspaceException(f"This operation is only supported in {self.required} state(s).")

    def op(self):
        if self.after:
            self.after()

    def log_entry(self):
        return dict(self.prev)


class Hyperspace:
    def __init__(self, session):
        self.session = session

    # ---- path resolution (HS/index/PathResolver.scala:39-69): names match case-insensitively
    def _index_path(self, name):
        root = self.session.system_path()
        if os.path.isdir(root):
            for d in os.listdir(root):
                if d.lower() == name.lower():
                    return os.path.join(root, d)
        return os.path.join(root, name)

    def _entries(self):
        root = self.session.system_path()
        out = []
        if os.path.isdir(root):
            for d in sorted(os.listdir(root)):
                e = le.IndexLogManager(os.path.join(root, d)).get_latest_stable_log()
                if e is not None:
                    out.append(e)
        return out

    # ---- API (names as in the reference)
    def createIndex(self, dataFrame, indexConfig):
        a = CreateAction(self.session, dataFrame, indexConfig, self._index_path(indexConfig.indexName))
        a.run()
        return a.stats

    def refreshIndex(self, indexName, mode="full"):
        if mode.lower() not in REFRESH_MODES:
            raise HyperspaceException(f"Unsupported refresh mode '{mode}' found.")
        a = RefreshAction(self.session, self._index_path(indexName), mode.lower())
        a.run()
        return a.stats

    def optimizeIndex(self, indexName, mode="quick"):
        a = OptimizeAction(self.session, self._index_path(indexName), mode)
        a.run()
        return a.stats

    def deleteIndex(self, indexName):
        _StateOnly(self.session, self._index_path(indexName), {le.States.ACTIVE}, le.States.DELETING,
                   le.States.DELETED).run()

    def restoreIndex(self, indexName):
        _StateOnly(self.session, self._index_path(indexName), {le.States.DELETED}, le.States.RESTORING,
                   le.States.ACTIVE).run()

    def vacuumIndex(self, indexName):
        path = self._index_path(indexName)

        def wipe():
            dm = le.IndexDataManager(path)
            v = dm.get_latest_version_id()
            while v is not None and v >= 0:
                dm.delete(v)
                v -= 1
        _StateOnly(self.session, path, {le.States.DELETED}, le.States.VACUUMING, le.States.DOESNOTEXIST, wipe).run()

    def cancel(self, indexName):
        path = self._index_path(indexName)
        log = le.IndexLogManager(path)
        latest = log.get_latest_log()
        if latest is None or latest["state"] in le.States.STABLE:
            raise HyperspaceException("Cancel is not supported in stable states.")
        stable = log.get_latest_stable_log()
        a = _StateOnly(self.session, path, None, le.States.CANCELLING,
                       stable["state"] if stable else le.States.DOESNOTEXIST)
        a.prev = stable if stable is not None else latest
        a.run()

    def indexes(self):
        """Summary rows like hs.indexes (HS/index/IndexStatistics.scala:64-71)."""
        rows = []
        for e in self._entries():
            p = e["derivedDataset"]["properties"]
            files = le.content_files(e["content"])
            rows.append({"name": e["name"], "indexedColumns": p["columns"]["indexed"],
                         "includedColumns": p["columns"]["included"], "numBuckets": p["numBuckets"],
                         "schema": p["schemaString"], "indexLocation": os.path.dirname(le.from_uri(files[0]["name"])) if files else "",
                         "state": e["state"]})
        return pa.Table.from_pylist(rows) if rows else pa.table({})

    def index(self, indexName):
        e = le.IndexLogManager(self._index_path(indexName)).get_latest_stable_log()
        if e is None:
            raise HyperspaceException(f"Index with name {indexName} could not be found.")
        return e

    def explain(self, df, verbose=False, redirectFunc=print):
        on = df.session.is_hyperspace_enabled()
        df.session.enable_hyperspace()
        with_hs = df.explain_string()
        df.session.disable_hyperspace()
        without = df.explain_string()
        if on:
            df.session.enable_hyperspace()
        redirectFunc("=============================================================\nPlan with indexes:\n"
                     "=============================================================\n" + with_hs +
                     "\n=============================================================\nPlan without indexes:\n"
                     "=============================================================\n" + without)

    # python/hyperspace/hyperspace.py:161-192
    @staticmethod
    def enable(session):
        return session.enable_hyperspace()

    @staticmethod
    def disable(session):
        return session.disable_hyperspace()

    @staticmethod
    def isEnabled(session):
        return session.is_hyperspace_enabled()
