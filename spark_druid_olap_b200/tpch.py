"""Synthetic flattened TPC-H (orderLineItemPartSupplier) for tests and the bench.

Wraps libsdo_tpchgen.so (csrc/tpch_flat_gen.cpp).  Schema: the reference's 53-column flat table
(/root/reference/src/main/scala/org/sparklinedata/druid/tools/TpchBenchMark.scala:43-97) minus
free-text columns; index layout as docs/benchmark/druid/tpch_index.json (time column l_shipdate,
month segments).  Data is a pure function of (sf, seed, shard): every rank / the oracle / the
engine see identical rows.
"""
import ctypes as C
import os

import numpy as np

from . import (SRC_INT32, SRC_STRING, dim_codes, dim_int, metric, time_column)

_HERE = os.path.dirname(os.path.abspath(__file__))
GEN_LIB = os.path.join(_HERE, "libsdo_tpchgen.so")
SEED = 20261106

KIND_TIME, KIND_INT_DIM, KIND_STR_DIM, KIND_F64, KIND_I64 = range(5)
TIME_COLUMN = "l_shipdate_ms"

_lib = None


def _gen():
    global _lib
    if _lib is None:
        if not os.path.exists(GEN_LIB):
            raise ImportError(f"{GEN_LIB} is missing: run `python -m spark_druid_olap_b200.build`")
        L = C.CDLL(GEN_LIB)
        L.sdo_tpch_column_kind.restype = C.c_int
        L.sdo_tpch_column_kind.argtypes = [C.c_char_p]
        L.sdo_tpch_count_rows.restype = C.c_int64
        L.sdo_tpch_count_rows.argtypes = [C.c_double, C.c_uint64, C.c_int, C.c_int]
        L.sdo_tpch_fill.restype = C.c_int64
        L.sdo_tpch_fill.argtypes = [C.c_double, C.c_uint64, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_void_p)]
        L.sdo_tpch_dictionary.restype = C.c_int64
        L.sdo_tpch_dictionary.argtypes = [C.c_char_p, C.c_double, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]
        L.sdo_tpch_num_columns.restype = C.c_int
        L.sdo_tpch_column_name.restype = C.c_char_p
        L.sdo_tpch_column_name.argtypes = [C.c_int]
        _lib = L
    return _lib


def column_kind(name):
    k = _gen().sdo_tpch_column_kind(name.encode())
    if k < 0:
        raise KeyError(name)
    return k


def all_columns():
    L = _gen()
    return [L.sdo_tpch_column_name(i).decode() for i in range(L.sdo_tpch_num_columns())]


def dictionary(name, sf):
    """code -> string list of a string dimension."""
    L = _gen()
    nb = C.c_int64()
    n = L.sdo_tpch_dictionary(name.encode(), sf, None, None, C.byref(nb))
    if n < 0:
        raise KeyError(name)
    offs = np.zeros(n + 1, dtype=np.int64)
    buf = np.zeros(max(1, nb.value), dtype=np.uint8)
    L.sdo_tpch_dictionary(name.encode(), sf, offs.ctypes.data, buf.ctypes.data, C.byref(nb))
    raw = buf.tobytes()
    return [raw[offs[i]:offs[i + 1]].decode() for i in range(n)]


class FlatTable:
    """Raw (un-encoded) columns of one shard: what Spark would hold in its cached base table."""

    def __init__(self, sf, columns, shard=0, n_shards=1, seed=SEED):
        L = _gen()
        self.sf, self.shard, self.n_shards, self.seed = sf, shard, n_shards, seed
        cols = list(dict.fromkeys([TIME_COLUMN] + list(columns)))
        self.kinds = {c: column_kind(c) for c in cols}
        self.n_rows = int(L.sdo_tpch_count_rows(sf, seed, shard, n_shards))
        self.data = {}
        for c in cols:
            k = self.kinds[c]
            dt = np.float64 if k == KIND_F64 else (np.int32 if k == KIND_STR_DIM else np.int64)
            self.data[c] = np.empty(self.n_rows, dtype=dt)
        names = (C.c_char_p * len(cols))(*[c.encode() for c in cols])
        ptrs = (C.c_void_p * len(cols))(*[self.data[c].ctypes.data for c in cols])
        got = L.sdo_tpch_fill(sf, seed, shard, n_shards, len(cols), names, ptrs)
        if got != self.n_rows:
            raise RuntimeError(f"generator wrote {got} rows, expected {self.n_rows}")
        self.dicts = {c: dictionary(c, sf) for c in cols if self.kinds[c] == KIND_STR_DIM}

    def engine_columns(self):
        """Columns in the shape sdo_segment_load takes (Spark source types recorded for HLL hashing)."""
        out = []
        for c, k in self.kinds.items():
            a = self.data[c]
            if k == KIND_TIME:
                out.append(time_column("__time", a))
            elif k == KIND_INT_DIM:
                out.append(dim_int(c, a, SRC_INT32))  # IntegerType in the reference schema
            elif k == KIND_STR_DIM:
                out.append(dim_codes(c, a, self.dicts[c], SRC_STRING))
            else:
                out.append(metric(c, a))
        return out

    def strings(self, c):
        """Decoded string values of a string dimension (small tables only)."""
        return np.array(self.dicts[c], dtype=object)[self.data[c]]


def load_into(engine, datasource, sf, columns, shards=(0,), n_shards=1, seed=SEED, create=True, keep=False, **options):
    """Generate the given shards and load them; returns the FlatTables if keep=True."""
    if create:
        engine.create_datasource(datasource, **options)
    kept = []
    for sh in shards:
        t = FlatTable(sf, columns, sh, n_shards, seed)
        engine.load(datasource, t.engine_columns(), partition=sh)
        if keep:
            kept.append(t)
    return kept
