"""IndexConfig — mirror of HS/index/IndexConfig.scala:28-77 (validation, case-insensitive equality) and of
python/hyperspace/indexconfig.py:1-14 (constructor shape)."""


class IllegalArgumentException(ValueError):
    """What the Scala case class throws from its constructor."""


class IndexConfig:
    def __init__(self, indexName, indexedColumns, includedColumns=()):
        self.indexName = indexName
        self.indexedColumns = list(indexedColumns)
        self.includedColumns = list(includedColumns)
        if not indexName or not self.indexedColumns:
            raise IllegalArgumentException("Empty index name or indexed columns are not allowed.")
        low_idx = [c.lower() for c in self.indexedColumns]
        low_inc = [c.lower() for c in self.includedColumns]
        if len(set(low_idx)) < len(low_idx):
            raise IllegalArgumentException("Duplicate indexed column names are not allowed.")
        if len(set(low_inc)) < len(low_inc):
            raise IllegalArgumentException("Duplicate included column names are not allowed.")
        if set(low_idx) & set(low_inc):
            raise IllegalArgumentException("Duplicate column names in indexed/included columns are not allowed.")

    def __eq__(self, other):
        return (isinstance(other, IndexConfig) and self.indexName.lower() == other.indexName.lower()
                and [c.lower() for c in self.indexedColumns] == [c.lower() for c in other.indexedColumns]
                and {c.lower() for c in self.includedColumns} == {c.lower() for c in other.includedColumns})

    def __hash__(self):
        return hash((tuple(c.lower() for c in self.indexedColumns), frozenset(c.lower() for c in self.includedColumns)))

    def __repr__(self):
        return (f"[indexName: {self.indexName}; indexedColumns: {', '.join(c.lower() for c in self.indexedColumns)}; "
                f"includedColumns: {', '.join(c.lower() for c in self.includedColumns)}]")
