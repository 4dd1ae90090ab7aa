"""Helpers shared by the parity tests: compare a GPU index/join/filter result with the oracle's."""
import numpy as np
import pyarrow as pa
import pyarrow.compute as pc


def bucket_column(boff):
    boff = np.asarray(boff, dtype=np.int64)
    return pa.array(np.repeat(np.arange(len(boff) - 1, dtype=np.int32), np.diff(boff)))


def _sortable(a):
    # sort_indices cannot order binary-like nulls/NaN consistently across versions: map floats to
    # their total-order integer image and leave the rest
    if pa.types.is_floating(a.type):
        v = a.to_numpy(zero_copy_only=False).astype(np.float64)
        bits = v.view(np.int64).copy()
        bits[np.isnan(v)] = np.int64(0x7ff8000000000000)
        neg = bits < 0
        bits[neg] = np.int64(-0x8000000000000000) - bits[neg] - 1  # flip magnitude of negatives
        m = pa.array(bits, mask=np.asarray(a.is_null()))
        return m
    return a


def canonical(cols, boff=None):
    """Table of (bucket?, cols...) sorted by every column: equal multisets <=> equal tables."""
    arrays = list(cols)
    names = [f"c{i}" for i in range(len(arrays))]
    if boff is not None:
        arrays = [bucket_column(boff)] + arrays
        names = ["bucket"] + names
    t = pa.table([_sortable(a if isinstance(a, pa.Array) else a.combine_chunks()) for a in arrays], names=names)
    idx = pc.sort_indices(t, sort_keys=[(n, "ascending") for n in names], null_placement="at_start")
    return t.take(idx)


def assert_same_multiset(got_cols, want_cols, got_boff=None, want_boff=None, what=""):
    if got_boff is not None:
        assert np.array_equal(np.asarray(got_boff), np.asarray(want_boff)), f"{what}: bucket offsets differ"
    g, w = canonical(got_cols, got_boff), canonical(want_cols, want_boff)
    assert g.num_rows == w.num_rows, f"{what}: {g.num_rows} rows vs {w.num_rows}"
    assert g.equals(w), f"{what}: row multisets differ"


def assert_keys_identical(got_cols, want_cols, nk, what=""):
    """The sorted key sequence must match position by position (bit-exact)."""
    for k in range(nk):
        g, w = _sortable(got_cols[k]), _sortable(want_cols[k])
        assert g.equals(w), f"{what}: key column {k} differs from the oracle's sorted sequence"


def check_index(oracle, got_cols, got_boff, in_cols, n_indexed, nb, what=""):
    """Full parity of an index build against the oracle."""
    perm, woff = oracle.build_index(in_cols, n_indexed, nb)
    want = oracle.take(in_cols, perm)
    assert np.array_equal(np.asarray(got_boff), woff), f"{what}: bucket offsets differ"
    # every row's bucket id, recomputed by the oracle on the GPU's output, must equal its bucket
    b = oracle.bucket_ids(got_cols[:n_indexed], nb)
    assert np.array_equal(b, np.repeat(np.arange(nb, dtype=np.int32), np.diff(woff))), f"{what}: a row is in the wrong bucket"
    assert_keys_identical(got_cols, want, n_indexed, what)
    assert_same_multiset(got_cols, want, got_boff, woff, what)
