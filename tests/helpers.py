"""Shared test helpers: feed the same generated rows to the engine (C ABI) and to the oracle."""
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import druid_oracle as orc  # noqa: E402  (tests are allowed to import the oracle)
from spark_druid_olap_b200 import tpch  # noqa: E402

F64_RTOL = 1e-9  # north_star: double sums within 1e-9 relative; everything else bit-exact

TEST_COLUMNS = ["l_returnflag", "l_linestatus", "l_shipmode", "l_quantity", "l_extendedprice", "l_discount", "l_tax", "o_orderkey",
                "o_custkey", "o_orderdate", "o_shippriority", "c_mktsegment", "c_nation", "c_region", "s_nation", "s_region", "p_type",
                "p_brand", "p_size", "ps_availqty", "ps_supplycost", "count", "o_orderpriority", "l_linenumber"]


def oracle_table(flat_tables):
    """Concatenate generator shards into an oracle Table of raw values."""
    if not isinstance(flat_tables, (list, tuple)):
        flat_tables = [flat_tables]
    t0 = flat_tables[0]
    cat = (lambda c: t0.data[c]) if len(flat_tables) == 1 else (lambda c: np.concatenate([t.data[c] for t in flat_tables]))  # one shard: no copy
    dims, metrics = {}, {}
    for c, k in t0.kinds.items():
        if k == tpch.KIND_STR_DIM:
            dims[c] = ("str", cat(c), t0.dicts[c])
        elif k == tpch.KIND_INT_DIM:
            dims[c] = ("int", cat(c))
        elif k in (tpch.KIND_F64, tpch.KIND_I64):
            metrics[c] = cat(c)
    return orc.Table(cat(tpch.TIME_COLUMN), dims, metrics)


def engine_rows(res):
    """Engine result -> list of row dicts shaped like the oracle's (dimension values as strings)."""
    out = []
    for i in range(res.n_rows):
        r = {}
        for n in res.names:
            v = res.columns[n][i]
            t = res.types[n]
            if t == 2:
                r[n] = str(v)
            elif t == 3:
                r[n] = str(int(v))
            elif t == 1:
                r[n] = float(v)
            else:
                r[n] = int(v)
        out.append(r)
    return out


def _key(r, names):
    return tuple((0, v) if isinstance(v, (int, float)) and not isinstance(v, bool) else (1, str(v)) for v in (r[n] for n in names))


def assert_rows_equal(got, want, ordered=False, float_cols=(), what=""):
    """Bit-exact on ints/strings; floats within F64_RTOL relative."""
    assert len(got) == len(want), f"{what}: {len(got)} rows, expected {len(want)}"
    if not want:
        return
    names = list(want[0].keys())
    assert set(got[0].keys()) == set(names), f"{what}: columns {sorted(got[0].keys())} vs {sorted(names)}"
    exact = [n for n in names if not isinstance(want[0][n], float)]
    if not ordered:
        got = sorted(got, key=lambda r: _key(r, exact))
        want = sorted(want, key=lambda r: _key(r, exact))
    for i, (g, w) in enumerate(zip(got, want)):
        for n in names:
            if isinstance(w[n], float):
                a, b = float(g[n]), w[n]
                if math.isinf(b) or math.isinf(a):
                    assert a == b, f"{what}: row {i} col {n}: {a} vs {b}"
                else:
                    assert abs(a - b) <= F64_RTOL * max(abs(b), 1e-300) or a == b, f"{what}: row {i} col {n}: {a!r} vs {b!r} (rel {abs(a-b)/max(abs(b),1e-300):.3e})"
            else:
                assert g[n] == w[n], f"{what}: row {i} col {n}: {g[n]!r} vs {w[n]!r}"
