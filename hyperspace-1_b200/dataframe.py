"""A minimal relational front end that stands where Spark stands for the reference: a session with
Hyperspace's configuration keys, Parquet relations, and DataFrames whose logical plans are small enough to be
pattern-matched by the index rules (Scan, Filter, Project, Join). Un-indexed plans run on pyarrow — that is
the host engine being plugged into, not part of the GPU hot path."""
import os
import re

import pyarrow as pa
import pyarrow.compute as pc
import pyarrow.parquet as pq

from . import log_entry as le


# ------------------------------------------------------------------ expressions
class Expr:
    def __and__(self, o): return BoolOp("and", self, o)
    def __or__(self, o): return BoolOp("or", self, o)
    def __invert__(self): return NotOp(self)


class Col(Expr):
    def __init__(self, name, df=None):
        self.name, self.df = name, df

    def _cmp(self, op, v):
        return Cmp(op, self, v)

    def __eq__(self, v): return EqCols(self, v) if isinstance(v, Col) else self._cmp("eq", v)  # noqa: E704
    def __ne__(self, v): return self._cmp("ne", v)  # noqa: E704
    def __lt__(self, v): return self._cmp("lt", v)  # noqa: E704
    def __le__(self, v): return self._cmp("le", v)  # noqa: E704
    def __gt__(self, v): return self._cmp("gt", v)  # noqa: E704
    def __ge__(self, v): return self._cmp("ge", v)  # noqa: E704
    def isNull(self): return NullTest(self, True)  # noqa: E704
    def isNotNull(self): return NullTest(self, False)  # noqa: E704
    __hash__ = object.__hash__


class EqCols(Expr):
    """left_col === right_col, the only join condition JoinIndexRule accepts (JoinIndexRule.scala:118-124)."""
    def __init__(self, l, r):
        self.l, self.r = l, r


class Cmp(Expr):
    def __init__(self, op, col, lit):
        self.op, self.col, self.lit = op, col, lit


class NullTest(Expr):
    def __init__(self, col, is_null):
        self.col, self.is_null = col, is_null


class BoolOp(Expr):
    def __init__(self, op, a, b):
        self.op, self.a, self.b = op, a, b


class NotOp(Expr):
    def __init__(self, a):
        self.a = a


_SQL_CMP = re.compile(r"^\s*([A-Za-z_][A-Za-z_0-9]*)\s*(=|==|!=|<>|<=|>=|<|>)\s*(.+?)\s*$")
_SQL_OPS = {"=": "eq", "==": "eq", "!=": "ne", "<>": "ne", "<": "lt", "<=": "le", ">": "gt", ">=": "ge"}


def parse_predicate(text):
    """'deptId = 20 AND name = \\'x\\'' -> Expr (the subset the quick start uses, App.scala:91)."""
    parts = re.split(r"\s+[aA][nN][dD]\s+", text)
    out = None
    for p in parts:
        m = _SQL_CMP.match(p)
        if not m:
            raise ValueError(f"cannot parse predicate {p!r}")
        name, op, lit = m.groups()
        if lit[0] in "'\"":
            val = lit[1:-1]
        else:
            val = float(lit) if any(ch in lit for ch in ".eE") else int(lit)
        e = Cmp(_SQL_OPS[op], Col(name), val)
        out = e if out is None else BoolOp("and", out, e)
    return out


def expr_columns(e):
    if isinstance(e, Cmp) or isinstance(e, NullTest):
        return [e.col.name]
    if isinstance(e, BoolOp):
        return expr_columns(e.a) + expr_columns(e.b)
    if isinstance(e, NotOp):
        return expr_columns(e.a)
    raise TypeError(e)


def expr_to_postfix(e, col_index):
    """Expr -> the postfix tuple list of hs_predicate."""
    if isinstance(e, Cmp):
        return [(e.op, col_index[e.col.name.lower()], e.lit)]
    if isinstance(e, NullTest):
        return [("is_null" if e.is_null else "is_not_null", col_index[e.col.name.lower()])]
    if isinstance(e, BoolOp):
        return expr_to_postfix(e.a, col_index) + expr_to_postfix(e.b, col_index) + [(e.op,)]
    if isinstance(e, NotOp):
        return expr_to_postfix(e.a, col_index) + [("not",)]
    raise TypeError(e)


def expr_to_arrow(e, resolve):
    if isinstance(e, Cmp):
        f = pc.field(resolve(e.col.name))
        return {"eq": f == e.lit, "ne": f != e.lit, "lt": f < e.lit, "le": f <= e.lit, "gt": f > e.lit,
                "ge": f >= e.lit}[e.op]
    if isinstance(e, NullTest):
        f = pc.field(resolve(e.col.name))
        return f.is_null() if e.is_null else f.is_valid()
    if isinstance(e, BoolOp):
        a, b = expr_to_arrow(e.a, resolve), expr_to_arrow(e.b, resolve)
        return (a & b) if e.op == "and" else (a | b)
    if isinstance(e, NotOp):
        return ~expr_to_arrow(e.a, resolve)
    raise TypeError(e)


# ------------------------------------------------------------------ relations and plans
class Relation:
    """A Parquet file relation: root path + the data files under it (names starting with '_' or '.' are not
    data, HS/util/PathUtils.scala:34-39)."""

    def __init__(self, root, files=None):
        self.root = os.path.abspath(root)
        if files is None:
            files = []
            for dp, _, fs in os.walk(self.root):
                for f in sorted(fs):
                    if f.startswith(".") or (f.startswith("_") and "=" not in f):
                        continue
                    files.append(os.path.join(dp, f))
        self.files = sorted(os.path.abspath(f) for f in files)
        self.schema = pq.read_schema(self.files[0]) if self.files else pa.schema([])

    def file_infos(self, tracker=None):
        return [le.file_info(f, tracker) for f in self.files]

    def read(self, columns=None, files=None):
        files = self.files if files is None else files
        if not files:
            return self.schema.empty_table() if columns is None else self.schema.empty_table().select(columns)
        return pa.concat_tables([pq.ParquetFile(f).read(columns=columns) for f in files])


class Scan:
    node_name = "LogicalRelation"

    def __init__(self, relation):
        self.relation = relation
        self.children = []

    def output(self):
        return list(self.relation.schema.names)


class IndexScan:
    """A relation swapped for index files (IndexHadoopFsRelation, HS/index/plans/logical/
    IndexHadoopFsRelation.scala:29-50). use_bucket_spec mirrors RuleUtils.scala:264-292."""
    node_name = "IndexScan"

    def __init__(self, entry, use_bucket_spec, hybrid=None):
        self.entry, self.use_bucket_spec = entry, use_bucket_spec
        # hybrid scan (RuleUtils.scala:307-450): source files appended since the index was built are indexed on
        # the fly and merged in; rows of deleted source files are dropped through the lineage column
        self.hybrid = hybrid or {}
        self.children = []

    def output(self):
        props = self.entry["derivedDataset"]["properties"]["columns"]
        return list(props["indexed"]) + list(props["included"])

    def describe(self):
        h = ""
        if self.hybrid.get("appended") or self.hybrid.get("deleted_ids"):
            h = f", HybridScan: +{len(self.hybrid.get('appended', []))} files -{len(self.hybrid.get('deleted_ids', []))} files"
        return f"Hyperspace(Type: CI, Name: {self.entry['name']}, LogVersion: {self.entry['id']}{h})"


class Filter:
    node_name = "Filter"

    def __init__(self, child, pred):
        self.child, self.pred = child, pred
        self.children = [child]

    def output(self):
        return self.child.output()


class Project:
    node_name = "Project"

    def __init__(self, child, cols):
        self.child, self.cols = child, list(cols)
        self.children = [child]

    def output(self):
        return list(self.cols)


class Join:
    node_name = "Join"

    def __init__(self, left, right, lkeys, rkeys, how="inner"):
        self.left, self.right, self.lkeys, self.rkeys = left, right, list(lkeys), list(rkeys)
        self.how = normalise_join_type(how)
        self.children = [left, right]

    def output(self):
        if self.how in ("left_semi", "left_anti"):
            return self.left.output()
        return self.left.output() + self.right.output()


_JOIN_TYPES = {"inner": "inner", "left": "left_outer", "leftouter": "left_outer", "left_outer": "left_outer",
               "right": "right_outer", "rightouter": "right_outer", "right_outer": "right_outer",
               "outer": "full_outer", "full": "full_outer", "fullouter": "full_outer", "full_outer": "full_outer",
               "semi": "left_semi", "leftsemi": "left_semi", "left_semi": "left_semi",
               "anti": "left_anti", "leftanti": "left_anti", "left_anti": "left_anti"}


def normalise_join_type(how):
    """Spark's joinType spellings (Dataset.join(right, joinExprs, joinType)) -> one name per type."""
    k = how.lower().replace(" ", "")
    if k not in _JOIN_TYPES:
        raise ValueError(f"Unsupported join type '{how}'")
    return _JOIN_TYPES[k]


def resolve_name(name, available):
    """Case-insensitive column resolution returning the source spelling (HS/util/ResolverUtils.scala:52-73)."""
    for a in available:
        if a.lower() == name.lower():
            return a
    return None


class DataFrame:
    def __init__(self, session, plan):
        self.session, self.plan = session, plan

    def __getitem__(self, name):
        return Col(name, self)

    def __call__(self, name):
        return Col(name, self)

    @property
    def columns(self):
        return self.plan.output()

    def filter(self, pred):
        if isinstance(pred, str):
            pred = parse_predicate(pred)
        return DataFrame(self.session, Filter(self.plan, pred))

    where = filter

    def select(self, *cols):
        names = [c.name if isinstance(c, Col) else c for c in cols]
        return DataFrame(self.session, Project(self.plan, names))

    def join(self, other, on, how="inner"):
        """on: EqCols, a conjunction of EqCols, or a list of column names present on both sides;
        how: Spark's join type names (inner, left_outer, right_outer, full_outer, left_semi, left_anti, ...)."""
        if isinstance(on, (list, tuple)) and all(isinstance(c, str) for c in on):
            return DataFrame(self.session, Join(self.plan, other.plan, list(on), list(on), how))
        pairs = []

        def walk(e):
            if isinstance(e, EqCols):
                pairs.append((e.l, e.r))
            elif isinstance(e, BoolOp) and e.op == "and":
                walk(e.a)
                walk(e.b)
            else:
                raise ValueError("join condition must be a conjunction of column equalities")
        walk(on)
        lk, rk = [], []
        for a, b in pairs:
            if a.df is other and b.df is self:
                a, b = b, a
            lk.append(a.name)
            rk.append(b.name)
        return DataFrame(self.session, Join(self.plan, other.plan, lk, rk, how))

    def collect(self):
        """-> pyarrow.Table. With Hyperspace enabled the plan goes through the index rules first."""
        from . import rules
        plan = rules.apply_rules(self.session, self.plan) if self.session.is_hyperspace_enabled() else self.plan
        return rules.execute(self.session, plan)

    def optimized_plan(self):
        from . import rules
        return rules.apply_rules(self.session, self.plan) if self.session.is_hyperspace_enabled() else self.plan

    def explain_string(self):
        from . import rules
        return rules.plan_string(self.optimized_plan())


DEFAULT_CONF = {
    "spark.hyperspace.system.path": None,              # IndexConstants.scala:26 (default <warehouse>/indexes)
    "spark.hyperspace.index.numBuckets": "200",        # IndexConstants.scala:29-32
    "spark.hyperspace.index.lineage.enabled": "false",  # IndexConstants.scala:75-76
    "spark.hyperspace.index.optimize.fileSizeThreshold": str(256 * 1024 * 1024),  # IndexConstants.scala:86-87
    "spark.hyperspace.index.hybridscan.enabled": "false",          # IndexConstants.scala:34-35
    "spark.hyperspace.index.hybridscan.maxAppendedRatio": "0.3",   # IndexConstants.scala:43-45
    "spark.hyperspace.index.hybridscan.maxDeletedRatio": "0.2",    # IndexConstants.scala:38-40
}


class Session:
    """What the reference gets from SparkSession: configuration, Parquet reads, the enable/disable switch."""

    def __init__(self, system_path=None, device=0, conf=None):
        self.conf = dict(DEFAULT_CONF)
        if conf:
            self.conf.update(conf)
        if system_path is not None:
            self.conf["spark.hyperspace.system.path"] = system_path
        if self.conf["spark.hyperspace.system.path"] is None:
            self.conf["spark.hyperspace.system.path"] = os.path.abspath(os.path.join("spark-warehouse", "indexes"))
        self.device = device
        self._ctx = None
        self._enabled = False

    @property
    def ctx(self):
        """The thread's hs_ctx (created on first GPU use; fails loudly without a device)."""
        if self._ctx is None:
            from . import native
            self._ctx = native.Context(self.device)
        return self._ctx

    def read_parquet(self, path):
        return DataFrame(self, Scan(Relation(path)))

    # HS/package.scala:47-79
    def enable_hyperspace(self):
        self._enabled = True
        return self

    def disable_hyperspace(self):
        self._enabled = False
        return self

    def is_hyperspace_enabled(self):
        return self._enabled

    enableHyperspace, disableHyperspace, isHyperspaceEnabled = enable_hyperspace, disable_hyperspace, is_hyperspace_enabled

    def num_buckets(self):
        return int(self.conf["spark.hyperspace.index.numBuckets"])

    def system_path(self):
        return self.conf["spark.hyperspace.system.path"]
