"""The optimizer rules that route queries to the index, and the physical operators they lead to.

 * FilterIndexRule (HS/index/rules/FilterIndexRule.scala:43-85): [Project ->] Filter -> Scan becomes a scan of an
   index whose first indexed column is in the predicate and which covers every referenced column; the index
   relation carries NO bucket spec (:59-65). Executed by hs_filter_scan.
 * JoinIndexRule (HS/index/rules/JoinIndexRule.scala:57-90): an equi-join of two linear sub-plans becomes a join
   of two compatible indexes WITH bucket spec, ranked as JoinIndexRanker does (equal bucket counts first, then
   more buckets). Executed by hs_join_bucketed, bucket by bucket, no exchange.
Any failure inside a rule falls back to the original plan, as the reference does (JoinIndexRule.scala:85-89)."""
import numpy as np
import pyarrow as pa

from . import log_entry as le
from .dataframe import (Filter, IndexScan, Join, Project, Scan, expr_columns, expr_to_arrow, expr_to_postfix,
                        resolve_name)


# ------------------------------------------------------------------ candidate selection
def _active_indexes(session):
    from .hyperspace import Hyperspace
    return [e for e in Hyperspace(session)._entries() if e["state"] == le.States.ACTIVE]


def _signature_matches(entry, relation):
    """RuleUtils.getCandidateIndexes (HS/index/rules/RuleUtils.scala:61-76): the signature recorded at build
    time must equal the signature of the relation as it is now."""
    sigs = entry["source"]["plan"]["properties"]["fingerprint"]["properties"]["signatures"]
    now = le.source_signature(relation.file_infos())
    return any(s["provider"] == le.SIGNATURE_PROVIDER and s["value"] == now for s in sigs)


def _applicable(session, entry, relation):
    """None if the index cannot serve the relation as it is now, else the hybrid-scan work it needs
    ({} = none): {"appended": [paths], "deleted_ids": [file ids]}.
    * the signature matches (RuleUtils.scala:61-76) -> {}, or, after a quick refresh, the update lists the
      entry recorded (IndexLogEntry.hasSourceUpdate -> transformPlanToUseHybridScan, RuleUtils.scala:236-262);
    * otherwise, with spark.hyperspace.index.hybridscan.enabled, the index is a candidate if it shares files
      with the relation, deleted files are covered by lineage, and the appended / deleted bytes stay below the
      configured ratios (RuleUtils.scala:84-147)."""
    rel = entry["source"]["plan"]["properties"]["relations"][0]["data"]["properties"]
    if _signature_matches(entry, relation):
        upd = rel.get("update") or {}
        hybrid = {}
        if upd.get("appendedFiles"):
            hybrid["appended"] = [le.from_uri(f["name"]) for f in le.content_files(upd["appendedFiles"])]
        if upd.get("deletedFiles"):
            hybrid["deleted_ids"] = [f["id"] for f in le.content_files(upd["deletedFiles"])]
        return hybrid
    if str(session.conf.get("spark.hyperspace.index.hybridscan.enabled", "false")).lower() != "true":
        return None
    from .index_writer import has_lineage
    logged = le.content_files(rel["content"])
    key = lambda f: (f["name"], f["size"], f["modifiedTime"])
    now = relation.file_infos()
    lset, nset = {key(f) for f in logged}, {key(f) for f in now}
    common = lset & nset
    if not common:
        return None
    deleted = [f for f in logged if key(f) not in nset]
    appended = [f for f in now if key(f) not in lset]
    if deleted and not has_lineage(entry):
        return None
    common_bytes = sum(f["size"] for f in logged if key(f) in common)
    app_bytes, del_bytes = sum(f["size"] for f in appended), sum(f["size"] for f in deleted)
    max_app = float(session.conf["spark.hyperspace.index.hybridscan.maxAppendedRatio"])
    max_del = float(session.conf["spark.hyperspace.index.hybridscan.maxDeletedRatio"])
    if app_bytes > max_app * (common_bytes + app_bytes) or del_bytes > max_del * (common_bytes + del_bytes):
        return None
    return {"appended": [le.from_uri(f["name"]) for f in appended], "deleted_ids": [f["id"] for f in deleted]}


def _index_columns(entry):
    c = entry["derivedDataset"]["properties"]["columns"]
    return list(c["indexed"]), list(c["included"])


def _linear_scan(plan):
    """(scan, filters, projected) if plan is Project/Filter nodes over exactly one Scan, else None."""
    filters, proj = [], None
    node = plan
    while True:
        if isinstance(node, Project):
            if proj is None:
                proj = node.cols
            node = node.child
        elif isinstance(node, Filter):
            filters.append(node.pred)
            node = node.child
        elif isinstance(node, Scan):
            return node, filters, proj
        else:
            return None


# ------------------------------------------------------------------ FilterIndexRule
def _filter_rule(session, plan):
    if isinstance(plan, Project) and isinstance(plan.child, Filter) and isinstance(plan.child.child, Scan):
        proj, flt, scan = plan.cols, plan.child, plan.child.child
    elif isinstance(plan, Filter) and isinstance(plan.child, Scan):
        proj, flt, scan = None, plan, plan.child
    else:
        return None
    out_cols = proj if proj is not None else scan.output()
    fcols = expr_columns(flt.pred)
    needed = {c.lower() for c in list(out_cols) + fcols}
    cands = []
    for e in _active_indexes(session):
        indexed, included = _index_columns(e)
        allc = {c.lower() for c in indexed + included}
        if indexed[0].lower() not in {c.lower() for c in fcols}:   # FilterIndexRule.scala:141-152
            continue
        if not needed <= allc:
            continue
        hybrid = _applicable(session, e, scan.relation)
        if hybrid is None:
            continue
        cands.append((e, hybrid))
    if not cands:
        return None
    best, hybrid = cands[0]                                         # FilterIndexRanker.rank(...).head
    if hybrid:
        hybrid = dict(hybrid, relation=scan.relation)
    node = Filter(IndexScan(best, use_bucket_spec=False, hybrid=hybrid), flt.pred)
    return Project(node, out_cols) if proj is not None else Project(node, scan.output())


# ------------------------------------------------------------------ JoinIndexRule
def _join_rule(session, plan):
    if not isinstance(plan, Join):
        return None
    l, r = _linear_scan(plan.left), _linear_scan(plan.right)
    if l is None or r is None:
        return None
    (lscan, lf, lproj), (rscan, rf, rproj) = l, r
    if lf or rf:
        return None   # filters under the join are kept on the un-indexed path in this build
    lout = lproj if lproj is not None else lscan.output()
    rout = rproj if rproj is not None else rscan.output()

    def usable(scan, keys, out):
        res = []
        need = {c.lower() for c in list(out) + list(keys)}
        for e in _active_indexes(session):
            indexed, included = _index_columns(e)
            if {c.lower() for c in indexed} != {k.lower() for k in keys}:     # JoinIndexRule.scala:451-463
                continue
            if not need <= {c.lower() for c in indexed + included}:
                continue
            hybrid = _applicable(session, e, scan.relation)
            if hybrid is None:
                continue
            e = dict(e)
            e["__hybrid"] = dict(hybrid, relation=scan.relation) if hybrid else {}
            res.append(e)
        return res
    lc, rc = usable(lscan, plan.lkeys, lout), usable(rscan, plan.rkeys, rout)
    pairs = []
    for a in lc:
        for b in rc:
            ai, _ = _index_columns(a)
            bi, _ = _index_columns(b)
            # compatible: the i-th indexed column of one side joins the i-th of the other (:524-533)
            lmap = {k.lower(): v.lower() for k, v in zip(plan.lkeys, plan.rkeys)}
            if [lmap.get(c.lower()) for c in ai] == [c.lower() for c in bi]:
                pairs.append((a, b))
    if not pairs:
        return None

    def rank(p):                                                     # JoinIndexRanker.scala:52-90
        na = p[0]["derivedDataset"]["properties"]["numBuckets"]
        nb = p[1]["derivedDataset"]["properties"]["numBuckets"]
        return (0 if na == nb else 1, -max(na, nb))
    a, b = sorted(pairs, key=rank)[0]
    # different bucket counts are fine: hs_join_bucketed re-buckets the smaller side (Spark re-shuffles one)
    ai, _ = _index_columns(a)
    lmap = {k.lower(): v for k, v in zip(plan.lkeys, plan.rkeys)}
    j = Join(IndexScan(a, True, a.get("__hybrid")), IndexScan(b, True, b.get("__hybrid")), ai,
             [resolve_name(lmap[c.lower()], _index_columns(b)[0]) for c in ai], plan.how)
    j.left_out, j.right_out = list(lout), ([] if plan.how in ("left_semi", "left_anti") else list(rout))
    return j


def prune_columns(plan):
    """Spark prunes columns before user rules run; the same here for Project over Join: every side keeps only
    the columns the projection and the join condition need (a name found on both sides is kept on both)."""
    if isinstance(plan, Project) and isinstance(plan.child, Join):
        j = plan.child
        sides = []
        for child, keys in ((j.left, j.lkeys), (j.right, j.rkeys)):
            have = child.output()
            need = [c for c in have if any(c.lower() == p.lower() for p in plan.cols) or
                    any(c.lower() == k.lower() for k in keys)]
            sides.append(Project(child, need) if len(need) < len(have) and not isinstance(child, Project) else child)
        return Project(Join(sides[0], sides[1], j.lkeys, j.rkeys, j.how), plan.cols)
    return plan


def apply_rules(session, plan):
    """JoinIndexRule first (transformUp), then FilterIndexRule (transformDown), as enableHyperspace installs
    them (HS/package.scala:47-54)."""
    plan = prune_columns(plan)
    def rewrite(node, rule):
        try:
            new = rule(session, node)
        except Exception:
            new = None        # rules swallow errors and keep the original plan
        if new is not None:
            return new
        if isinstance(node, Project):
            return Project(rewrite(node.child, rule), node.cols)
        if isinstance(node, Filter):
            return Filter(rewrite(node.child, rule), node.pred)
        if isinstance(node, Join) and not isinstance(node.left, IndexScan):
            j = Join(rewrite(node.left, rule), rewrite(node.right, rule), node.lkeys, node.rkeys, node.how)
            return j
        return node
    plan = rewrite(plan, _join_rule)
    plan = rewrite(plan, _filter_rule)
    return plan


# ------------------------------------------------------------------ execution
def plan_string(plan, depth=0):
    pad = "  " * depth
    if isinstance(plan, Scan):
        return f"{pad}Scan parquet {plan.relation.root}\n"
    if isinstance(plan, IndexScan):
        return f"{pad}IndexScan {plan.describe()} bucketed={plan.use_bucket_spec}\n"
    if isinstance(plan, Filter):
        op = "GpuFilteredIndexScan" if isinstance(plan.child, IndexScan) else "Filter"
        return f"{pad}{op}\n" + plan_string(plan.child, depth + 1)
    if isinstance(plan, Project):
        return f"{pad}Project {plan.cols}\n" + plan_string(plan.child, depth + 1)
    if isinstance(plan, Join):
        op = "GpuBucketedSortMergeJoin" if isinstance(plan.left, IndexScan) else "Join"
        return f"{pad}{op} {plan.lkeys} = {plan.rkeys}\n" + plan_string(plan.left, depth + 1) + plan_string(plan.right, depth + 1)
    return f"{pad}{plan}\n"


def _single_run(session, entry, columns=None, hybrid=None):
    """All files of an index as ONE sorted run per bucket; several files per bucket are merged on the GPU
    (what Spark does with a per-side SortExec, E2EHyperspaceRulesTest.scala:474-478). With hybrid scan
    (RuleUtils.scala:307-450): rows of deleted source files leave through NOT (_data_file_id IN ids), appended
    source files are indexed on the fly with the index's bucket count (the reference injects a shuffle,
    RuleUtils.scala:519-578) and merged bucket by bucket (BucketUnionExec.scala:52-121). The lineage column is
    not part of the result."""
    from .hyperspace import drop_deleted_rows
    from .index_writer import LINEAGE_COL, read_index
    indexed, included = _index_columns(entry)
    runs, names, _ = read_index(entry, session=session)
    nb = entry["derivedDataset"]["properties"]["numBuckets"]
    nk = len(indexed)
    hybrid = hybrid or {}
    if len(runs) > 1:
        arrays, boff = session.ctx.merge_runs([(a, nk, off) for a, off in runs if a is not None])
    elif runs:
        arrays, boff = runs[0]
    else:
        arrays, boff = None, np.zeros(nb + 1, dtype=np.int64)
    if arrays is not None and hybrid.get("deleted_ids"):
        arrays, boff = drop_deleted_rows(session, arrays, nk, boff, names, hybrid["deleted_ids"])
    if LINEAGE_COL in names:
        if arrays is not None:
            arrays = [a for a, n in zip(arrays, names) if n != LINEAGE_COL]
        names = [n for n in names if n != LINEAGE_COL]
    if hybrid.get("appended"):
        t = hybrid["relation"].read(columns=names, files=hybrid["appended"])
        cols = [t[c].combine_chunks() for c in names]
        new, noff = session.ctx.build_index(cols, nk, nb)
        if arrays is None:
            arrays, boff = new, noff
        else:
            arrays, boff = session.ctx.merge_runs([(arrays, nk, boff), (new, nk, noff)])
    return arrays, names, boff


class _DeviceRun:
    """The files of an index decoded into device memory: one Table per column + the bucket offsets."""

    def __init__(self, tables, names, boff, n_indexed):
        from . import abi
        from .native import IndexView
        self.tables, self.names, self.boff = tables, names, boff
        spec = []
        for t in tables:
            c = t.column(0)
            spec.append({"type": c.type, "length": c.length, "null_count": c.null_count, "validity": c.validity,
                         "offsets": c.offsets, "data": c.data})
        self.view = IndexView(abi.DeviceColumnSet(spec), n_indexed, boff)

    def free(self):
        for t in self.tables:
            t.free()
        self.tables = []


def _device_run(session, scan):
    """An index scan whose files can go from disk to the join / filter kernels without a stop in host arrays: one
    file per bucket, no hybrid-scan work, no lineage column, files the device reader covers. None otherwise (the
    caller takes the general path through _single_run)."""
    from . import parquet_read
    from .index_writer import LINEAGE_COL, index_column_names
    from .native import HyperspaceException
    if session.conf.get("hsgpu.index.fileReader", "gpu") != "gpu" or scan.hybrid:
        return None
    entry = scan.entry
    names = index_column_names(entry)
    if LINEAGE_COL in names:
        return None
    nb = entry["derivedDataset"]["properties"]["numBuckets"]
    per_bucket = {}
    for fi in le.content_files(entry["content"]):
        per_bucket.setdefault(le.bucket_id_of(fi["name"]), []).append(le.from_uri(fi["name"]))
    if not per_bucket or any(len(v) != 1 for v in per_bucket.values()):
        return None
    buckets = sorted(per_bucket)
    try:
        tables, got_names, rows = parquet_read.decode_files(session.ctx, [per_bucket[b][0] for b in buckets], names, device=True)
    except HyperspaceException:
        return None
    off = np.zeros(nb + 1, dtype=np.int64)
    for b, r in zip(buckets, rows):
        off[b + 1] = r
    tables = [tables[got_names.index(c)] for c in names]
    import pyarrow.parquet as pq
    schema = pq.read_schema(per_bucket[buckets[0]][0])
    run = _DeviceRun(tables, names, np.cumsum(off), len(_index_columns(entry)[0]))
    run.types = [schema.field(c).type for c in names]
    return run


def execute(session, plan):
    if isinstance(plan, Scan):
        return plan.relation.read()
    if isinstance(plan, Project):
        t = execute(session, plan.child)
        return t.select([resolve_name(c, t.column_names) for c in plan.cols]).rename_columns(list(plan.cols))
    if isinstance(plan, Filter):
        if isinstance(plan.child, IndexScan):
            entry = plan.child.entry
            arrays, names, boff = _single_run(session, entry, hybrid=plan.child.hybrid)
            if arrays is None:
                return pa.table({n: pa.array([]) for n in names})
            idx = {n.lower(): i for i, n in enumerate(names)}
            terms = expr_to_postfix(plan.pred, idx)
            n_indexed = len(_index_columns(entry)[0])
            out = session.ctx.filter_scan(arrays, n_indexed, boff, terms, list(range(len(names))))
            return pa.table(out, names=names)
        t = execute(session, plan.child)
        return t.filter(expr_to_arrow(plan.pred, lambda n: resolve_name(n, t.column_names)))
    if isinstance(plan, Join):
        if isinstance(plan.left, IndexScan):
            # both indexes straight from their files into device memory, joined there; only the result comes back
            ld = _device_run(session, plan.left)
            rd = _device_run(session, plan.right) if ld is not None else None
            if ld is not None and rd is not None:
                try:
                    lout = [ld.names.index(resolve_name(c, ld.names)) for c in plan.left_out]
                    rout = [rd.names.index(resolve_name(c, rd.names)) for c in plan.right_out]
                    t = session.ctx.join_table(ld.view, rd.view, lout, rout, how=plan.how)
                    try:
                        out = t.to_arrow([ld.types[i] for i in lout] + [rd.types[i] for i in rout])
                    finally:
                        t.free()
                    return pa.table(out, names=_dedupe([ld.names[i] for i in lout] + [rd.names[i] for i in rout]))
                finally:
                    ld.free()
                    rd.free()
            if ld is not None:
                ld.free()
            la, ln, lo = _single_run(session, plan.left.entry, hybrid=plan.left.hybrid)
            ra, rn, ro = _single_run(session, plan.right.entry, hybrid=plan.right.hybrid)
            lout = [ln.index(resolve_name(c, ln)) for c in plan.left_out]
            rout = [rn.index(resolve_name(c, rn)) for c in plan.right_out]
            nk = len(_index_columns(plan.left.entry)[0])
            if la is None or ra is None:
                return pa.table({})
            out, _ = session.ctx.join(la, lo, ra, ro, nk, lout, rout, how=plan.how)
            names = _dedupe([ln[i] for i in lout] + [rn[i] for i in rout])
            return pa.table(out, names=names)
        lt, rt = execute(session, plan.left), execute(session, plan.right)
        lk = [resolve_name(k, lt.column_names) for k in plan.lkeys]
        rk = [resolve_name(k, rt.column_names) for k in plan.rkeys]
        names = _dedupe(lt.column_names + rt.column_names)
        lt2 = lt.rename_columns(names[: lt.num_columns])
        rt2 = rt.rename_columns(names[lt.num_columns:])
        lk2 = [names[lt.column_names.index(k)] for k in lk]
        rk2 = [names[lt.num_columns + rt.column_names.index(k)] for k in rk]
        # equi-join on the host engine; NULL keys match nothing, both key columns are kept like Spark does
        pa_how = {"inner": "inner", "left_outer": "left outer", "right_outer": "right outer",
                  "full_outer": "full outer", "left_semi": "left semi", "left_anti": "left anti"}[plan.how]
        j = lt2.join(rt2, keys=lk2, right_keys=rk2, join_type=pa_how, coalesce_keys=False)
        if plan.how in ("left_semi", "left_anti"):
            return j.select(names[: lt.num_columns])
        return j.select(names)
    if isinstance(plan, IndexScan):
        arrays, names, _ = _single_run(session, plan.entry, hybrid=plan.hybrid)
        return pa.table(arrays, names=names)
    raise TypeError(plan)


def _dedupe(names):
    seen, out = {}, []
    for n in names:
        k = seen.get(n, 0)
        out.append(n if k == 0 else f"{n}_{k}")
        seen[n] = k + 1
    return out
