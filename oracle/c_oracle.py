"""ctypes front end of the C restatement (oracle/c/oracle_c.c) -- TEST / BENCH INFRASTRUCTURE ONLY.

Translates a Druid query dict into the C oracle's flat description: string predicates are evaluated once
per dictionary value with the numpy oracle's own predicate code (so both oracles share the predicate
semantics but not the execution), integer predicates become value ranges / tables, the time bucket becomes
a day-offset lookup table.  Grouped results come back as rows and go through druid_oracle.finalize
(post-aggregations, having, ordering, limit).  Returns None for shapes it does not cover (the caller then
uses the numpy oracle)."""
import ctypes as C
import os
import subprocess

import numpy as np

from . import druid_oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "c", "oracle_c.c")
LIB = os.path.join(HERE, "c", "liboracle_c.so")
_lib = None


def build(force=False):
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        subprocess.check_call(["gcc", "-O3", "-march=native", "-fPIC", "-shared", "-pthread", "-o", LIB, SRC, "-lm"])
    return LIB


def available():
    try:
        _load()
        return True
    except Exception:
        return False


def default_threads():
    return os.cpu_count() or 1


class _Leaf(C.Structure):
    _fields_ = [("kind", C.c_int), ("col_type", C.c_int), ("col", C.c_void_p), ("table", C.c_void_p), ("lo", C.c_int64), ("hi", C.c_int64)]


class _Dim(C.Structure):
    _fields_ = [("col_type", C.c_int), ("col", C.c_void_p), ("base", C.c_int64), ("card", C.c_int64)]


class _Agg(C.Structure):
    _fields_ = [("kind", C.c_int), ("col_is_f64", C.c_int), ("col", C.c_void_p), ("hll_col_type", C.c_int), ("hll_src", C.c_int),
                ("hll_table", C.c_void_p), ("filter_leaf", C.c_int)]


class _Query(C.Structure):
    _fields_ = [("time", C.c_void_p), ("n_rows", C.c_int64), ("n_intervals", C.c_int), ("intervals", C.c_void_p), ("n_leaves", C.c_int),
                ("n_ops", C.c_int), ("leaves", C.POINTER(_Leaf)), ("op_code", C.c_void_p), ("op_arg", C.c_void_p), ("n_dims", C.c_int),
                ("dims", C.POINTER(_Dim)), ("has_bucket", C.c_int), ("t0", C.c_int64), ("unit", C.c_int64), ("n_lut", C.c_int64),
                ("n_buckets", C.c_int64), ("lut", C.c_void_p), ("n_aggs", C.c_int), ("aggs", C.POINTER(_Agg)), ("n_threads", C.c_int)]


def _load():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        L.oc_execute.restype = C.c_void_p
        L.oc_execute.argtypes = [C.POINTER(_Query)]
        L.oc_result_groups.restype = C.c_int64
        L.oc_result_groups.argtypes = [C.c_void_p]
        L.oc_result_fetch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oc_result_free.argtypes = [C.c_void_p]
        L.oc_xxh64_int.restype = C.c_uint64
        L.oc_xxh64_int.argtypes = [C.c_int32, C.c_uint64]
        L.oc_xxh64_long.restype = C.c_uint64
        L.oc_xxh64_long.argtypes = [C.c_int64, C.c_uint64]
        _lib = L
    return _lib


class CTable:
    """Contiguous, typed copies of an oracle Table's columns (int32 codes, int64 ints, float64 / int64 metrics)."""

    def __init__(self, tbl):
        self.tbl = tbl
        self.time = np.ascontiguousarray(tbl.time, dtype=np.int64)
        self.dims = {}
        for n, d in tbl.dims.items():
            if d[0] == "str":
                self.dims[n] = ("str", np.ascontiguousarray(d[1], dtype=np.int32), list(d[2]))
            else:
                self.dims[n] = (d[0], np.ascontiguousarray(d[1], dtype=np.int64))
        self.metrics = {n: np.ascontiguousarray(v, dtype=np.float64 if np.asarray(v).dtype.kind == "f" else np.int64) for n, v in tbl.metrics.items()}
        self.day0 = int(self.time.min()) // 86400000 if len(self.time) else 0
        self.days = (int(self.time.max()) // 86400000 - self.day0 + 1) if len(self.time) else 0
        self._hll_tables = {}
        self._ranked = {}

    def ranked(self, name):
        """(codes re-ranked by the UTF-8 byte order of their strings, that order) -- cached: it is table preparation, not query work."""
        if name not in self._ranked:
            d = self.dims[name]
            order = sorted(range(len(d[2])), key=lambda i: d[2][i].encode("utf-8"))
            rank = np.empty(len(d[2]), dtype=np.int32)
            rank[order] = np.arange(len(order), dtype=np.int32)
            self._ranked[name] = (np.ascontiguousarray(rank[d[1]]), order)
        return self._ranked[name]

    def hll_table(self, name):
        if name not in self._hll_tables:
            d = self.dims[name]
            t = np.zeros(len(d[2]), dtype=np.uint16)
            for i, s in enumerate(d[2]):
                idx, rho = orc.hll_idx_rho(orc.spark_hash(s, "string"))
                t[i] = (idx << 6) | rho
            self._hll_tables[name] = t
        return self._hll_tables[name]


class _Unsupported(Exception):
    pass


_AGG_KIND = {"count": 0, "longSum": 1, "doubleSum": 2, "longMin": 3, "longMax": 4, "doubleMin": 5, "doubleMax": 6, "min": 5, "max": 6}


def execute(ct, q, threads=None, return_registers=False):
    try:
        return _execute(ct, q, threads or default_threads(), return_registers)
    except _Unsupported:
        return None


def execute_partial(ct, q, threads=None):
    """The grouped rows of `q` over this table BEFORE post-aggregations / having / ordering / limit, for merging with the
    partials of other tables (bench.py's parity gate runs one shard of the generator at a time).
    Returns (rows carrying "__key", registers, key_names) or None when the shape is not covered."""
    try:
        return _execute(ct, q, threads or default_threads(), True, partial=True)
    except _Unsupported:
        return None


def merge_partials(q, partials, return_registers=False):
    """Merge execute_partial() outputs of disjoint row sets and finish the query: count / sum -> sum (doubles with
    math.fsum over the partial sums), min -> min, max -> max, cardinality -> register-wise max then the estimator
    (the merge algebra of sources/druid/PostAggregate.scala:39-60)."""
    import math
    kinds = {}
    for a in q.get("aggregations", []):
        name = a.get("name")
        if a["type"] == "filtered":
            name = a.get("name") or a["aggregator"]["name"]
            a = a["aggregator"]
        t = a["type"]
        if t == "javascript":
            params, body = orc.js_subset.parse_function(a["fnAggregate"])
            fold, _ = orc.js_subset.split_fold(params, body)
            kinds[name] = {"sum": "fsum", "min": "min", "max": "max"}[fold]
        elif t in ("cardinality", "hyperUnique"):
            kinds[name] = "hll"
        elif t in ("count", "longSum"):
            kinds[name] = "isum"
        elif t == "doubleSum":
            kinds[name] = "fsum"
        else:
            kinds[name] = "min" if t.lower().endswith("min") else "max"
    acc, regs, key_names = {}, {}, None
    for rows, registers, kn in partials:
        key_names = kn
        for i, r in enumerate(rows):
            k = r["__key"]
            t = acc.get(k)
            if t is None:
                t = acc[k] = {n: ([v] if kinds.get(n) == "fsum" else v) for n, v in r.items()}
                for n in registers:
                    regs[(n, k)] = registers[n][i].copy()
                continue
            for n, kind in kinds.items():
                if kind == "isum":
                    t[n] += r[n]
                elif kind == "fsum":
                    t[n].append(r[n])
                elif kind == "min":
                    t[n] = min(t[n], r[n])
                elif kind == "max":
                    t[n] = max(t[n], r[n])
            for n in registers:
                np.maximum(regs[(n, k)], registers[n][i], out=regs[(n, k)])
    keys = sorted(acc)
    out = []
    for k in keys:
        t = acc[k]
        for n, kind in kinds.items():
            if kind == "fsum":
                t[n] = math.fsum(t[n])
            elif kind == "hll":
                t[n] = orc.hll_estimate(regs[(n, k)])
        out.append(t)
    reg_names = [n for n, kind in kinds.items() if kind == "hll"]
    registers = {n: np.array([regs[(n, k)] for k in keys], dtype=np.uint8).reshape(-1, 512) for n in reg_names}
    return orc.finalize(q, out, registers, key_names or [], return_registers)


def _execute(ct, q, threads, return_registers, partial=False):
    L = _load()
    keep = []  # numpy buffers that must outlive the call
    qt = q["queryType"]
    if qt not in ("groupBy", "timeseries", "topN"):
        raise _Unsupported()  # search / select: the numpy oracle only
    # intervals
    ivs = q.get("intervals")
    if isinstance(ivs, dict):
        ivs = [s["itvl"] for s in ivs["segments"]] if ivs.get("type") == "segments" else ivs["intervals"]
    iv = np.array([x for s in (ivs or []) for x in orc.parse_interval(s)], dtype=np.int64)
    keep.append(iv)
    # filter -> leaves + postfix program
    leaves, op_code, op_arg = [], [], []

    def leaf_of(f):
        dim = f["dimension"]
        if dim == "__time":
            raise _Unsupported()
        lf = _Leaf()
        if dim not in ct.dims:
            # unknown column = all-null: constant predicate on ""
            ok = bool(orc.filter_mask(orc.Table(np.zeros(1, dtype=np.int64), {}, {}), f)[0])
            arr = np.zeros(1, dtype=np.int32)
            tab = np.array([1 if ok else 0], dtype=np.uint8)
            zero = np.zeros(ct.tbl.n, dtype=np.int32)
            keep.extend([tab, zero, arr])
            lf.kind, lf.col_type, lf.col, lf.table = 0, 0, zero.ctypes.data, tab.ctypes.data
            return lf
        d = ct.dims[dim]
        if d[0] == "str":
            probe = orc.Table(np.zeros(len(d[2]), dtype=np.int64), {dim: ("str", np.arange(len(d[2])), d[2])}, {})
            tab = orc.filter_mask(probe, f).astype(np.uint8)
            keep.append(tab)
            lf.kind, lf.col_type, lf.col, lf.table = 0, 0, d[1].ctypes.data, tab.ctypes.data
        else:
            u = np.unique(d[1])
            lo = int(u.min())
            if int(u.max()) - lo > (1 << 26):
                raise _Unsupported()
            probe = orc.Table(np.zeros(len(u), dtype=np.int64), {dim: (d[0], u)}, {})
            ok = orc.filter_mask(probe, f)
            tab = np.zeros(int(u.max()) - lo + 1, dtype=np.uint8)
            tab[u[ok] - lo] = 1
            shifted = (d[1] - lo).astype(np.int64)
            keep.extend([tab, shifted])
            lf.kind, lf.col_type, lf.col, lf.table = 0, 1, shifted.ctypes.data, tab.ctypes.data
        return lf

    def emit(f):
        t = f["type"]
        if t in ("and", "or"):
            if not f["fields"]:
                raise _Unsupported()
            for i, c in enumerate(f["fields"]):
                emit(c)
                if i:
                    op_code.append(1 if t == "and" else 2)
                    op_arg.append(0)
        elif t == "not":
            emit(f["field"])
            op_code.append(3)
            op_arg.append(0)
        elif t == "noop":
            raise _Unsupported()
        else:
            leaves.append(leaf_of(f))
            op_code.append(0)
            op_arg.append(len(leaves) - 1)

    if q.get("filter"):
        f = q["filter"]
        if f.get("type") == "selector" and f.get("dimension") == "__time":
            raise _Unsupported()
        emit(f)
    # dimensions
    dims = [q["dimension"]] if qt == "topN" else (q.get("dimensions") or [] if qt == "groupBy" else [])
    dims = [{"type": "default", "dimension": d, "outputName": d} if isinstance(d, str) else d for d in dims]
    cdims, decoders = [], []
    for d in dims:
        if d.get("type", "default") != "default":
            raise _Unsupported()
        dd = ct.dims[d["dimension"]]
        cd = _Dim()
        name = d.get("outputName", d["dimension"])
        if dd[0] == "str":
            col, order = ct.ranked(d["dimension"])  # codes re-ranked by UTF-8 order so that key order is string order
            cd.col_type, cd.col, cd.base, cd.card = 0, col.ctypes.data, 0, len(dd[2])
            ss = [dd[2][i] for i in order]
            decoders.append((name, len(dd[2]), 0, (lambda ss_: lambda v: ss_[int(v)])(ss)))
        else:
            lo, hi = int(dd[1].min()), int(dd[1].max())
            cd.col_type, cd.col, cd.base, cd.card = 1, dd[1].ctypes.data, lo, hi - lo + 1
            decoders.append((name, hi - lo + 1, lo, lambda v: str(int(v))))
        cdims.append(cd)
    # time bucket
    bfn = orc.bucket_fn(q.get("granularity", "all"))
    cq = _Query()
    buckets = None
    if bfn is not None:
        day = 86400000
        starts = np.array([bfn((ct.day0 + i) * day) for i in range(ct.days)], dtype=np.int64)
        ends = np.array([bfn((ct.day0 + i) * day + day - 1) for i in range(ct.days)], dtype=np.int64)
        if not np.array_equal(starts, ends) or np.any(ct.time % day):
            raise _Unsupported()  # sub-day buckets or sub-day timestamps: numpy oracle handles them
        buckets, inv = np.unique(starts, return_inverse=True)
        lut = inv.astype(np.int32)
        keep.append(lut)
        cq.has_bucket, cq.t0, cq.unit, cq.n_lut, cq.n_buckets, cq.lut = 1, ct.day0 * day, day, len(lut), len(buckets), lut.ctypes.data
    total_card = (len(buckets) if buckets is not None else 1)
    for _, card, _, _ in decoders:
        total_card *= card
    if total_card >= 1 << 63:
        raise _Unsupported()
    # aggregations
    caggs, agg_names = [], []
    for a in q.get("aggregations", []):
        name = a.get("name")
        fl = -1
        if a["type"] == "filtered":
            leaves.append(leaf_of(a["filter"]))
            fl = len(leaves) - 1
            name = a.get("name") or a["aggregator"]["name"]
            a = a["aggregator"]
        ca = _Agg()
        ca.filter_leaf = fl
        t = a["type"]
        if t == "javascript" or "expression" in a:
            # the argument expression becomes one more float64 column (numpy, IEEE per-row arithmetic)
            if t == "javascript":
                params, body = orc.js_subset.parse_function(a["fnAggregate"])
                fold, arg = orc.js_subset.split_fold(params, body)
                cols = [None] + [np.asarray(ct.metrics[f], dtype=np.float64) for f in a["fieldNames"]]
            else:
                params, arg = orc.js_subset.parse_expression(a["expression"])
                fold = {"doubleSum": "sum", "doubleMin": "min", "doubleMax": "max"}.get(t)
                if fold is None:
                    raise _Unsupported()
                cols = [np.asarray(ct.metrics[f], dtype=np.float64) for f in params]
            col = np.ascontiguousarray(np.broadcast_to(orc.js_subset.evaluate_numeric(arg, cols), (len(ct.time),)), dtype=np.float64)
            keep.append(col)
            ca.kind = {"sum": 2, "min": 5, "max": 6}[fold]
            ca.col, ca.col_is_f64 = col.ctypes.data, 1
        elif t in _AGG_KIND:
            ca.kind = _AGG_KIND[t]
            if t != "count":
                m = ct.metrics[a["fieldName"]]
                ca.col, ca.col_is_f64 = m.ctypes.data, int(m.dtype.kind == "f")
        elif t == "cardinality" or (t == "hyperUnique" and a["fieldName"] in ct.dims):
            fld = a["fieldNames"][0] if t == "cardinality" else a["fieldName"]
            dd = ct.dims[fld]
            ca.kind = 7
            if dd[0] == "str":
                tab = ct.hll_table(fld)
                ca.col, ca.hll_col_type, ca.hll_table = dd[1].ctypes.data, 0, tab.ctypes.data
            else:
                ca.col, ca.hll_col_type, ca.hll_src = dd[1].ctypes.data, 1, 1 if dd[0] == "int" else 2
        else:
            raise _Unsupported()
        caggs.append(ca)
        agg_names.append((name, t))
    if len(caggs) > 64:
        raise _Unsupported()
    leaf_arr = (_Leaf * max(1, len(leaves)))(*leaves)
    dim_arr = (_Dim * max(1, len(cdims)))(*cdims)
    agg_arr = (_Agg * max(1, len(caggs)))(*caggs)
    oc = np.array(op_code, dtype=np.uint8)
    oa = np.array(op_arg, dtype=np.int32)
    cq.time, cq.n_rows = ct.time.ctypes.data, len(ct.time)
    cq.n_intervals, cq.intervals = (len(iv) // 2 if ivs is not None else 0), iv.ctypes.data
    if ivs is not None and len(iv) == 0:
        cq.n_intervals, cq.intervals = 1, np.array([0, 0], dtype=np.int64).ctypes.data
    cq.n_leaves, cq.n_ops, cq.leaves, cq.op_code, cq.op_arg = len(leaves), len(op_code), leaf_arr, oc.ctypes.data, oa.ctypes.data
    cq.n_dims, cq.dims, cq.n_aggs, cq.aggs, cq.n_threads = len(cdims), dim_arr, len(caggs), agg_arr, threads
    h = L.oc_execute(C.byref(cq))
    try:
        n = int(L.oc_result_groups(h))
        n_hll = sum(1 for _, t in agg_names if t in ("cardinality", "hyperUnique"))
        keys = np.zeros(n, dtype=np.uint64)
        counts = np.zeros(n, dtype=np.int64)
        vals = np.zeros((n, max(1, len(caggs))), dtype=np.float64)
        hll = np.zeros((n, max(1, n_hll), 512), dtype=np.uint8)
        L.oc_result_fetch(h, keys.ctypes.data, counts.ctypes.data, vals.ctypes.data, hll.ctypes.data if n_hll else None)
    finally:
        L.oc_result_free(h)
    # decode rows
    order = np.argsort(keys)
    out, registers = [], {name: [] for name, t in agg_names if t in ("cardinality", "hyperUnique")}
    vi = vals.view(np.int64)
    for r in order:
        k = int(keys[r])
        comps = []
        for name, card, base, dec in reversed(decoders):
            comps.append((name, k % card, base, dec))
            k //= card
        comps.reverse()
        row, key_t = {}, []
        if buckets is not None:
            row["timestamp"] = int(buckets[k])
            key_t.append(int(buckets[k]))
        for name, c, base, dec in comps:
            row[name] = dec(c + base)
            key_t.append(c + base if base else c)
        hi = 0
        for a, (name, t) in enumerate(agg_names):
            if t in ("cardinality", "hyperUnique"):
                row[name] = orc.hll_estimate(hll[r, hi])
                registers[name].append(hll[r, hi].copy())
                hi += 1
            elif t in ("count", "longSum", "longMin", "longMax"):
                row[name] = int(vi[r, a])
            else:
                row[name] = float(vals[r, a])
        row["__key"] = tuple(key_t)
        out.append(row)
    registers = {k: np.array(v, dtype=np.uint8).reshape(-1, 512) for k, v in registers.items()}
    key_names = (["timestamp"] if buckets is not None else []) + [d[0] for d in decoders]
    if partial:
        return out, registers, key_names
    return orc.finalize(q, out, registers, key_names, return_registers)
