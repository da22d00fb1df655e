"""Shared helpers for the parity tests."""
import numpy as np

from oracle import oracle as O

COLS = O.COLUMNS


def assert_tables_equal(got: dict, want: dict, what=""):
    """Bit-exact comparison of two sorted 20-column tables (dicts of numpy arrays)."""
    assert got["n_rows"] == want["n_rows"], f"{what}: {got['n_rows']} rows, oracle has {want['n_rows']}"
    for c in COLS:
        g, w = got[c], want[c]
        if c.startswith("sc_"):
            assert np.array_equal(got["sc_valid"], want["sc_valid"]), f"{what}: sc validity differs"
            m = want["sc_valid"]
            assert np.array_equal(g[m].view(np.uint32), w[m].view(np.uint32)), f"{what}: column {c} differs"
        elif c == "distance":
            bad = np.nonzero(g.view(np.uint32) != w.view(np.uint32))[0]
            assert len(bad) == 0, f"{what}: distance differs at {bad[:5]}: {g[bad[:5]]} vs {w[bad[:5]]}"
        else:
            bad = np.nonzero(g != w)[0]
            assert len(bad) == 0, f"{what}: column {c} differs at rows {bad[:5]}: {g[bad[:5]]} vs {w[bad[:5]]}"


def row_multiset(t: dict):
    """Hashable multiset view of a table, for order-free comparisons."""
    keys = list(zip(*[t[c].tolist() if t[c].dtype != object else t[c] for c in COLS[:17] if c != "distance"],
                    t["distance"].view(np.uint32).tolist()))
    return sorted(keys)
