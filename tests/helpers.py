"""Shared helpers for the parity tests: golden loading, order-normalised multiset comparison."""
import json
import os
from collections import Counter

import numpy as np
import pyarrow as pa

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lib_rs_scenarios.json")


def load_golden():
    with open(GOLDEN) as f:
        return json.load(f)


def table_from_dict(d):
    cols, names = [], []
    for name, vals in d.items():
        names.append(name)
        if name.startswith("id"):
            cols.append(pa.array(vals, type=pa.int32()))
        else:
            cols.append(pa.array(vals, type=pa.string()))
    return pa.Table.from_arrays(cols, names=names)


def rows_multiset(table_or_cols):
    """Order-normalised multiset of full output rows (the comparison BASELINE.json's north_star names)."""
    if isinstance(table_or_cols, pa.Table):
        cols = [c.to_pylist() for c in table_or_cols.columns]
    else:
        cols = [list(c) for c in table_or_cols]
    if not cols:
        return Counter()
    return Counter(zip(*cols))


def pairs_multiset(build_idx, probe_idx):
    b = np.asarray(build_idx, dtype=np.uint64)
    p = np.asarray(probe_idx, dtype=np.uint32)
    order = np.lexsort((b, p))
    return b[order], p[order]
