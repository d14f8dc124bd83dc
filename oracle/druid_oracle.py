"""CPU oracle (TEST INFRASTRUCTURE ONLY -- never imported by the product path).

A numpy restatement of what the reference's parity harness treats as truth: Spark SQL executing
the *un-rewritten* query on the raw flattened table (cTest,
/root/reference/src/test/scala/org/sparklinedata/druid/client/test/AbstractTest.scala:127-143), expressed
over the same Druid query JSON the engine receives.  It works on RAW values (strings, ints, doubles,
epoch ms) -- no dictionary ids, no segments -- so it shares no data layout with the engine.

Spec -> relational meaning follows:
  filters      sources/druid/ProjectFilterTransfom.scala:198-416 (selector/bound/in/and/or/not, null == "")
  intervals    QueryIntervals.scala:96-130 (half-open [start, end) in epoch ms)
  dims / aggs  sources/druid/AggregateTransform.scala:48-96, 331-518
  post-aggs    DruidQuerySpec.scala:379-430
  limit / topN DruidQuerySpec.scala:437-506, query/QuerySpecTransforms.scala:281-298
  merge        sources/druid/PostAggregate.scala:39-60 (merge_partials, used by the multi-rank tests)
Value semantics are Spark's: SUM(long) exact int64, SUM(double) exact (math.fsum), string order =
UTF-8 byte order, approx_count_distinct = HyperLogLog++ p = 9 over XxHash64(seed 42) of the typed value.

PARITY PINNING: the reference holds no golden result values for this path (its fixture data is
stripped and every value check is differential against a live Spark); Druid 0.9.1 and Spark 2.0.2
are un-vendored dependencies (build.sbt:16-17).  What IS pinned here: XXH64 against the public
XXH64 known-answer vectors, spec shapes against the reference's JSON fixtures
(docs/benchmark/druid/queries/*.json) and FilterPlanningTest's expected filter trees
(tests/golden/).  Aggregate values: parity unpinned (no reference-run outputs exist offline).
"""
import datetime as _dt
import math
import re

import numpy as np

try:
    from . import js_subset
except ImportError:  # imported as a top-level module
    import js_subset

M64 = (1 << 64) - 1
P1, P2, P3, P4, P5 = 0x9E3779B185EBCA87, 0xC2B2AE3D27D4EB4F, 0x165667B19E3779F9, 0x85EBCA77C2B2AE63, 0x27D4EB2F165667C5
HLL_P, HLL_M, HLL_SEED = 9, 512, 42


# ---- XXH64 (published algorithm; Spark's XxHash64 = XXH64 of the value's little-endian bytes) -----
def _rotl(x, r):
    return ((x << r) | (x >> (64 - r))) & M64


def _round(acc, inp):
    return (_rotl((acc + inp * P2) & M64, 31) * P1) & M64


def _merge(h, v):
    return ((h ^ _round(0, v)) * P1 + P4) & M64


def xxh64(data: bytes, seed: int = 0) -> int:
    n = len(data)
    i = 0
    if n >= 32:
        v1, v2, v3, v4 = (seed + P1 + P2) & M64, (seed + P2) & M64, seed & M64, (seed - P1) & M64
        while i <= n - 32:
            v1 = _round(v1, int.from_bytes(data[i:i + 8], "little"))
            v2 = _round(v2, int.from_bytes(data[i + 8:i + 16], "little"))
            v3 = _round(v3, int.from_bytes(data[i + 16:i + 24], "little"))
            v4 = _round(v4, int.from_bytes(data[i + 24:i + 32], "little"))
            i += 32
        h = (_rotl(v1, 1) + _rotl(v2, 7) + _rotl(v3, 12) + _rotl(v4, 18)) & M64
        for v in (v1, v2, v3, v4):
            h = _merge(h, v)
    else:
        h = (seed + P5) & M64
    h = (h + n) & M64
    while i + 8 <= n:
        h ^= _round(0, int.from_bytes(data[i:i + 8], "little"))
        h = (_rotl(h, 27) * P1 + P4) & M64
        i += 8
    if i + 4 <= n:
        h ^= (int.from_bytes(data[i:i + 4], "little") * P1) & M64
        h = (_rotl(h, 23) * P2 + P3) & M64
        i += 4
    while i < n:
        h ^= (data[i] * P5) & M64
        h = (_rotl(h, 11) * P1) & M64
        i += 1
    h ^= h >> 33
    h = (h * P2) & M64
    h ^= h >> 29
    h = (h * P3) & M64
    h ^= h >> 32
    return h


def spark_hash(value, src_type):
    """XxHash64(seed 42) of a typed Spark value: 'int' -> 4 LE bytes, 'long' -> 8 LE bytes, 'string' -> UTF-8."""
    if src_type == "int":
        return xxh64(int(value).to_bytes(4, "little", signed=True), HLL_SEED)
    if src_type == "long":
        return xxh64(int(value).to_bytes(8, "little", signed=True), HLL_SEED)
    return xxh64(value.encode("utf-8") if isinstance(value, str) else bytes(value), HLL_SEED)


def hll_idx_rho(h):
    idx = h >> (64 - HLL_P)
    w = ((h << HLL_P) & M64) | (1 << (HLL_P - 1))
    rho = (64 - w.bit_length()) + 1
    return idx, rho


def hll_estimate(regs):
    """HyperLogLogPlusPlusHelper.query for p = 9 without the empirical bias tables (see engine.cpp note)."""
    m = float(HLL_M)
    zinv = 0.0
    V = 0
    for r in regs:
        zinv += 1.0 / float(1 << int(r))
        if r == 0:
            V += 1
    alpha_m2 = (0.7213 / (1.0 + 1.079 / m)) * m * m
    est = alpha_m2 / zinv
    if V > 0:
        H = m * math.log(m / V)
        if H <= 400.0:
            est = H
    return float(math.floor(est + 0.5))


# ---- time -------------------------------------------------------------------------------------------
_EPOCH = _dt.datetime(1970, 1, 1)
_ISO = re.compile(r"^(-?\d+)(?:-(\d\d)(?:-(\d\d))?)?(?:[T ](\d\d)(?::(\d\d)(?::(\d\d)(?:[.,](\d+))?)?)?)?(Z|[+-]\d\d:?\d\d)?$")


def parse_instant(s):
    m = _ISO.match(s.strip())
    if not m:
        raise ValueError(f"bad ISO instant {s!r}")
    y, mo, d, h, mi, se, frac, tz = m.groups()
    dt = _dt.datetime(int(y), int(mo or 1), int(d or 1)) + _dt.timedelta(hours=int(h or 0), minutes=int(mi or 0), seconds=int(se or 0))
    ms = int((dt - _EPOCH).total_seconds()) * 1000
    if frac:
        ms += int((frac + "000")[:3])
    if tz and tz != "Z":
        sign = -1 if tz[0] == "-" else 1
        t = tz[1:].replace(":", "")
        ms -= sign * (int(t[:2]) * 3600000 + int(t[2:]) * 60000)
    return ms


def parse_interval(s):
    a, b = s.split("/")
    return parse_instant(a), parse_instant(b)


def _tz_ms(tz):
    if not tz or tz in ("UTC", "Z", "GMT", "Etc/UTC"):
        return 0
    m = re.match(r"^(?:UTC|GMT)?([+-])(\d\d):?(\d\d)$", tz)
    if not m:
        raise NotImplementedError(f"time zone {tz}")
    return (1 if m.group(1) == "+" else -1) * (int(m.group(2)) * 3600000 + int(m.group(3)) * 60000)


def _zone(tz):
    """A fixed offset in ms, or a zoneinfo.ZoneInfo for a region id (the system's zone database, read by Python)."""
    try:
        return _tz_ms(tz)
    except NotImplementedError:
        import zoneinfo
        return zoneinfo.ZoneInfo(tz)


def _zoff(z, ms):
    """Offset of zone z (ms east of UTC) at the instant ms."""
    if isinstance(z, int):
        return z
    d = _dt.datetime(1970, 1, 1, tzinfo=_dt.timezone.utc) + _dt.timedelta(milliseconds=int(ms))
    return int(d.astimezone(z).utcoffset().total_seconds() * 1000)


def _local_to_utc(z, local_ms):
    """Wall-clock ms in zone z -> UTC ms; ambiguous times mean their first occurrence, skipped ones the offset before
    the change (PEP 495 fold = 0, as java.time / Joda's lenient conversion)."""
    if isinstance(z, int):
        return local_ms - z
    naive = _EPOCH + _dt.timedelta(milliseconds=int(local_ms))
    return local_ms - int(naive.replace(tzinfo=z).utcoffset().total_seconds() * 1000)


def _dt_of(ms, tz_ms=0):
    return _EPOCH + _dt.timedelta(milliseconds=int(ms) + tz_ms)


_MON = ["January", "February", "March", "April", "May", "June", "July", "August", "September", "October", "November", "December"]
_DOW = ["Monday", "Tuesday", "Wednesday", "Thursday", "Friday", "Saturday", "Sunday"]


def joda_format(fmt, ms, tz_ms=0):
    """Joda DateTimeFormat subset (DateTimeExtractor.scala:358-372 uses YYYY, MM, dd, DD, ww, HH, mm, ss, ...)."""
    d = _dt_of(ms, tz_ms)
    iso = d.isocalendar()
    out = []
    i = 0
    while i < len(fmt):
        ch = fmt[i]
        if ch == "'":
            j = fmt.index("'", i + 1)
            out.append("'" if j == i + 1 else fmt[i + 1:j])
            i = j + 1
            continue
        if not ch.isalpha():
            out.append(ch)
            i += 1
            continue
        j = i
        while j < len(fmt) and fmt[j] == ch:
            j += 1
        n = j - i
        z = lambda v: str(v).zfill(n)
        if ch in "yY":
            out.append(str(d.year % 100).zfill(2) if n == 2 else z(d.year))
        elif ch == "x":
            out.append(str(iso[0] % 100).zfill(2) if n == 2 else z(iso[0]))
        elif ch == "M":
            out.append(_MON[d.month - 1] if n >= 4 else (_MON[d.month - 1][:3] if n == 3 else z(d.month)))
        elif ch == "d":
            out.append(z(d.day))
        elif ch == "D":
            out.append(z(d.timetuple().tm_yday))
        elif ch == "w":
            out.append(z(iso[1]))
        elif ch == "e":
            out.append(z(iso[2]))
        elif ch == "E":
            out.append(_DOW[iso[2] - 1] if n >= 4 else _DOW[iso[2] - 1][:3])
        elif ch == "H":
            out.append(z(d.hour))
        elif ch == "m":
            out.append(z(d.minute))
        elif ch == "s":
            out.append(z(d.second))
        elif ch == "S":
            out.append(str(d.microsecond // 1000).zfill(3)[:n].ljust(n, "0"))
        else:
            raise NotImplementedError(f"pattern letter {ch}")
        i = j
    return "".join(out)


_SIMPLE = {"second": 1000, "minute": 60000, "fifteen_minute": 900000, "thirty_minute": 1800000, "hour": 3600000, "day": 86400000,
           "week": 7 * 86400000}


def _period(p):
    m = re.match(r"^P(?:(\d+)Y)?(?:(\d+)M)?(?:(\d+)W)?(?:(\d+)D)?(?:T(?:(\d+)H)?(?:(\d+)M)?(?:(\d+)S)?)?$", p)
    if not m:
        raise ValueError(p)
    y, mo, w, d, h, mi, s = [int(x or 0) for x in m.groups()]
    return y * 12 + mo, ((w * 7 + d) * 86400 + h * 3600 + mi * 60 + s) * 1000


def bucket_fn(gran):
    """granularity JSON -> function(ms) -> bucket start ms (None for 'all')."""
    if gran is None:
        return None
    if isinstance(gran, str):
        g = gran.lower()
        if g == "all":
            return None
        if g == "none":
            return lambda t: t
        if g in _SIMPLE:
            per = _SIMPLE[g]
            org = -3 * 86400000 if g == "week" else 0
            return lambda t: t - (t - org) % per
        months = {"month": 1, "quarter": 3, "year": 12}[g]
        return _month_bucket(months, 0)
    t = gran.get("type")
    if t == "all":
        return None
    if t == "none":
        return lambda x: x
    if t == "duration":
        dur = int(gran["duration"])
        org = gran.get("origin", 0)
        org = parse_instant(org) if isinstance(org, str) else int(org)
        return lambda x: x - (x - org) % dur
    if t == "period":
        months, fixed = _period(gran["period"])
        tz = _zone(gran.get("timeZone"))
        if months:
            return _month_bucket(months, tz)
        if gran.get("origin"):
            org = parse_instant(gran["origin"])
            return lambda x: x - (x - org) % fixed
        if fixed < 86400000:
            return lambda x: x - (x + _zoff(tz, x)) % fixed
        lorg = -3 * 86400000 if fixed % (7 * 86400000) == 0 else 0

        def local_bucket(x):
            loc = x + _zoff(tz, x)
            return _local_to_utc(tz, loc - (loc - lorg) % fixed)
        return local_bucket
    raise ValueError(gran)


def _month_bucket(months, tz):
    def f(t):
        d = _dt_of(t, _zoff(tz, t))
        idx = d.year * 12 + d.month - 1
        idx -= idx % months
        s = _dt.datetime(idx // 12, idx % 12 + 1, 1)
        return _local_to_utc(tz, int((s - _EPOCH).total_seconds()) * 1000)
    return f


# ---- table ------------------------------------------------------------------------------------------
class Table:
    """Raw columns.  dims[name] = ('str', codes, dictionary) or ('int'|'long', values); metrics[name] = values."""

    def __init__(self, time_ms, dims, metrics):
        self.time = np.asarray(time_ms, dtype=np.int64)
        self.dims = dims
        self.metrics = metrics
        self.n = len(self.time)

    def dim_strings_of_unique(self, name):
        """(unique string values list, inverse index per row).  Unknown column = all null ("")."""
        if name not in self.dims:
            return [""], np.zeros(self.n, dtype=np.int64)
        d = self.dims[name]
        if d[0] == "str":
            return list(d[2]), np.asarray(d[1], dtype=np.int64)
        u, inv = np.unique(d[1], return_inverse=True)
        return [str(int(x)) for x in u], inv


def _cmp_bytes(a, b):
    x, y = a.encode("utf-8"), b.encode("utf-8")
    return (x > y) - (x < y)


def _numeric(s):
    try:
        return float(s)
    except ValueError:
        return None


def _bound_ok(f, s):
    alpha = bool(f.get("alphaNumeric", False))

    def cmp(a, b):
        if alpha:
            x, y = _numeric(a), _numeric(b)
            if x is not None and y is not None:
                return (x > y) - (x < y)
        return _cmp_bytes(a, b)
    if f.get("lower") is not None:
        c = cmp(s, str(f["lower"]))
        if c < 0 or (c == 0 and f.get("lowerStrict", False)):
            return False
    if f.get("upper") is not None:
        c = cmp(s, str(f["upper"]))
        if c > 0 or (c == 0 and f.get("upperStrict", False)):
            return False
    return True


def filter_mask(tbl, f):
    t = f["type"]
    if t == "and":
        m = np.ones(tbl.n, dtype=bool)
        for c in f["fields"]:
            m &= filter_mask(tbl, c)
        return m
    if t == "or":
        m = np.zeros(tbl.n, dtype=bool)
        for c in f["fields"]:
            m |= filter_mask(tbl, c)
        return m
    if t == "not":
        return ~filter_mask(tbl, f["field"])
    if t == "noop":
        return np.ones(tbl.n, dtype=bool)
    dim = f["dimension"]
    if dim == "__time":
        if t == "selector" and (f.get("value") in (None, "")):
            return np.zeros(tbl.n, dtype=bool)
        raise NotImplementedError("__time filter")
    vals, inv = tbl.dim_strings_of_unique(dim)
    if t == "selector":
        want = f.get("value") or ""
        ok = [v == want for v in vals]
    elif t == "bound":
        ok = [_bound_ok(f, v) for v in vals]
    elif t == "extraction":
        mp = f["extractionFn"]["lookup"]["map"]
        want = f.get("value", "true")
        keys = {k for k, v in mp.items() if str(v).lower() == want.lower() or v == want}
        ok = [v in keys for v in vals]
    elif t == "in":
        keys = {"" if v is None else str(v) for v in f["values"]}
        ok = [v in keys for v in vals]
    elif t == "regex":
        rx = re.compile(f["pattern"])
        ok = [rx.search(v) is not None for v in vals]
    elif t == "contains":
        ok = [f["value"] in v for v in vals]
    elif t == "javascript":
        # JavascriptFilterSpec (DruidQuerySpec.scala:218-236): Druid calls the function with the dimension value, null
        # for the empty value, and keeps the row when the result is truthy
        params, body = js_subset.parse_function(f["function"])
        ok = [js_subset.truthy(js_subset.evaluate(body, [v if v != "" else None])) for v in vals]
    else:
        raise NotImplementedError(t)
    return np.asarray(ok, dtype=bool)[inv]


_SDF_FIELDS = "yMdHmsS"


def _parse_by_pattern(pat, v):
    """SimpleDateFormat-style numeric pattern (TimeParsingExtractionFunctionSpec.timeFormat) -> epoch ms or None."""
    rx, fields, i = "", [], 0
    while i < len(pat):
        c = pat[i]
        if c == "'":
            e = pat.index("'", i + 1)
            rx += re.escape(pat[i + 1:e] or "'")
            i = e + 1
        elif c.isalpha():
            run = len(pat[i:]) - len(pat[i:].lstrip(c))
            if c not in _SDF_FIELDS:
                raise NotImplementedError("pattern letter " + c)
            nxt_field = i + run < len(pat) and pat[i + run].isalpha()
            rx += "(\\d{%d})" % run if nxt_field else "(\\d+)"
            fields.append((c, run))
            i += run
        else:
            rx += re.escape(c)
            i += 1
    m = re.fullmatch(rx, v)
    if not m:
        return None
    f = dict(y=1970, M=1, d=1, H=0, m=0, s=0, S=0)
    for (c, run), g in zip(fields, m.groups()):
        x = int(g)
        if c == "y" and run == 2:
            x += 2000 if x < 70 else 1900
        f[c] = x
    if not (1 <= f["M"] <= 12 and 1 <= f["d"] <= 31 and f["H"] <= 23 and f["m"] <= 59 and f["s"] <= 59):
        return None
    days = (_dt.date(f["y"], f["M"], 1) - _dt.date(1970, 1, 1)).days + f["d"] - 1
    return days * 86400000 + f["H"] * 3600000 + f["m"] * 60000 + f["s"] * 1000 + f["S"]


def extract_value(fn, v):
    """Druid dimension extraction functions (DruidQuerySpec.scala:45-106) on one value; '' is null."""
    t = fn["type"]
    if t == "timeFormat":
        return "" if v == "" else joda_format(fn.get("format", "yyyy-MM-dd'T'HH:mm:ss.SSS'Z'"), parse_instant(v), _zoff(_zone(fn.get("timeZone")), parse_instant(v)))
    if t == "time":
        if v == "":
            return ""
        ms = _parse_by_pattern(fn["timeFormat"], v)
        return v if ms is None else joda_format(fn["resultFormat"], ms, 0)
    if t == "javascript":
        params, body = js_subset.parse_function(fn["function"])
        r = js_subset.evaluate(body, [v if v != "" else None])
        return "" if r is None else js_subset.to_string(r)
    if t == "lookup":
        mp = fn["lookup"]["map"]
        if v in mp:
            return mp[v]
        if fn.get("retainMissingValue"):
            return v
        return fn.get("replaceMissingValueWith") or ""
    if t == "regex":
        m = re.search(fn["expr"], v)
        if m:
            return m.group(1) if m.groups() else m.group(0)
        return (fn.get("replaceMissingValueWith") or "") if fn.get("replaceMissingValue") else v
    if t == "substring":
        b = v.encode("utf-8")
        i, n = int(fn.get("index", 0)), fn.get("length")
        return (b[i:] if n is None else b[i:i + int(n)]).decode("utf-8", "replace")
    if t == "upper":
        return "".join(c.upper() if "a" <= c <= "z" else c for c in v)
    if t == "lower":
        return "".join(c.lower() if "A" <= c <= "Z" else c for c in v)
    if t == "strlen":
        return str(len(v.encode("utf-8")))
    raise NotImplementedError(t)


def _fsum_groups(values, gid, ng):
    order = np.argsort(gid, kind="stable")
    sv = values[order]
    bounds = np.searchsorted(gid[order], np.arange(ng + 1))
    return [math.fsum(sv[bounds[g]:bounds[g + 1]].tolist()) for g in range(ng)]


def _src_type(tbl, name):
    d = tbl.dims[name]
    return {"str": "string", "int": "int", "long": "long"}[d[0]]


def hll_registers(tbl, name, rows, gid, ng):
    """Per-group HLL++ registers of dimension `name` over the selected rows."""
    d = tbl.dims[name]
    st = _src_type(tbl, name)
    if d[0] == "str":
        codes = np.asarray(d[1])[rows]
        u, inv = np.unique(codes, return_inverse=True)
        packed = [hll_idx_rho(spark_hash(d[2][int(c)], st)) for c in u]
    else:
        vals = np.asarray(d[1])[rows]
        u, inv = np.unique(vals, return_inverse=True)
        packed = [hll_idx_rho(spark_hash(int(v), st)) for v in u]
    idx = np.array([p[0] for p in packed], dtype=np.int64)[inv]
    rho = np.array([p[1] for p in packed], dtype=np.uint8)[inv]
    regs = np.zeros((ng, HLL_M), dtype=np.uint8)
    np.maximum.at(regs, (gid, idx), rho)
    return regs


def _post(p, row):
    t = p["type"]
    if t == "constant":
        return float(p["value"])
    if t in ("fieldAccess", "hyperUniqueCardinality"):
        return float(row[p["fieldName"]])
    if t == "arithmetic":
        acc = _post(p["fields"][0], row)
        for q in p["fields"][1:]:
            x = _post(q, row)
            fn = p["fn"]
            if fn == "+":
                acc += x
            elif fn == "-":
                acc -= x
            elif fn == "*":
                acc *= x
            elif fn == "/":
                acc = 0.0 if x == 0 else acc / x
            else:
                acc = acc / x
        return acc
    raise NotImplementedError(t)


def _having(h, row):
    t = h["type"]
    if t == "and":
        return all(_having(c, row) for c in h["havingSpecs"])
    if t == "or":
        return any(_having(c, row) for c in h["havingSpecs"])
    if t == "not":
        return not _having(h["havingSpec"], row)
    v = float(row[h["aggregation"]])
    return v > h["value"] if t == "greaterThan" else (v < h["value"] if t == "lessThan" else v == h["value"])


def _select(tbl, q):
    """SelectSpec (DruidQuerySpec.scala:980-1098) without the paging: every row that passes interval and filter, as
    {timestamp, dimension values (strings), metric values}; order unspecified (the engine pages through them in segment
    order, which this table does not know)."""
    mask = np.zeros(tbl.n, dtype=bool)
    ivs = q.get("intervals")
    if ivs is None:
        mask[:] = True
    else:
        if isinstance(ivs, dict):
            ivs = [s["itvl"] for s in ivs["segments"]] if ivs.get("type") == "segments" else ivs["intervals"]
        for s in ivs:
            a, b = parse_interval(s)
            mask |= (tbl.time >= a) & (tbl.time < b)
    if q.get("filter"):
        mask &= filter_mask(tbl, q["filter"])
    rows = np.nonzero(mask)[0]
    dims, mets = list(q.get("dimensions") or []), list(q.get("metrics") or [])
    if not dims and not mets:
        dims, mets = sorted(tbl.dims), sorted(tbl.metrics)
    cols = {"timestamp": tbl.time[rows].tolist()}
    for d in dims:
        vals, inv = tbl.dim_strings_of_unique(d)
        cols[d] = [vals[i] for i in inv[rows].tolist()]
    for m in mets:
        v = np.asarray(tbl.metrics[m])[rows]
        cols[m] = v.tolist()
    names = list(cols)
    return [dict(zip(names, vals)) for vals in zip(*(cols[n] for n in names))]


def _search(tbl, q):
    """SearchQuerySpec (DruidQuerySpec.scala:870-927): the values of the search dimensions that occur in the rows passing
    interval and filter and match the search query, with their row counts; ordered by value then dimension; null value
    not reported."""
    mask = np.zeros(tbl.n, dtype=bool)
    ivs = q.get("intervals")
    if ivs is None:
        mask[:] = True
    else:
        if isinstance(ivs, dict):
            ivs = [s["itvl"] for s in ivs["segments"]] if ivs.get("type") == "segments" else ivs["intervals"]
        for s in ivs:
            a, b = parse_interval(s)
            mask |= (tbl.time >= a) & (tbl.time < b)
    if q.get("filter"):
        mask &= filter_mask(tbl, q["filter"])
    sq = q["query"]
    if sq["type"] == "insensitive_contains":
        terms, fold = [sq.get("value") or ""], True
    elif sq["type"] == "contains":
        terms, fold = [sq.get("value") or ""], not sq.get("caseSensitive", True)
    elif sq["type"] == "fragment":
        terms, fold = list(sq["values"]), not sq.get("caseSensitive", False)
    else:
        raise NotImplementedError(sq["type"])
    low = lambda s: "".join(c.lower() if "A" <= c <= "Z" else c for c in s)  # noqa: E731  ASCII case folding
    hits = []
    dims = [d if isinstance(d, str) else d["dimension"] for d in (q.get("searchDimensions") or sorted(tbl.dims))]
    for d in dims:
        if d not in tbl.dims:
            continue
        vals, inv = tbl.dim_strings_of_unique(d)
        counts = np.bincount(inv[mask], minlength=len(vals))
        for v, c in zip(vals, counts.tolist()):
            if c and v != "" and all((low(t) in low(v)) if fold else (t in v) for t in terms):
                hits.append({"dimension": d, "value": v, "count": int(c)})
    strlen = (q.get("sort") or {}).get("type") == "strlen" if isinstance(q.get("sort"), dict) else q.get("sort") == "strlen"
    hits.sort(key=lambda h: ((len(h["value"].encode("utf-8")),) if strlen else ()) + (h["value"].encode("utf-8"), h["dimension"]))
    return hits[:int(q.get("limit", 1000))]


def execute(tbl, q, return_registers=False):
    """Run a groupBy / timeseries / topN query dict; returns a list of row dicts in result order.
    Dimension values are strings (integer dimensions as decimal strings); 'timestamp' holds the bucket start."""
    qt = q["queryType"]
    if qt == "search":
        return _search(tbl, q)
    if qt == "select":
        return _select(tbl, q)
    mask = np.zeros(tbl.n, dtype=bool)
    ivs = q.get("intervals")
    if ivs is None:
        mask[:] = True
    else:
        if isinstance(ivs, dict):
            ivs = [s["itvl"] for s in ivs["segments"]] if ivs.get("type") == "segments" else ivs["intervals"]
        for s in ivs:
            a, b = parse_interval(s)
            mask |= (tbl.time >= a) & (tbl.time < b)
    if q.get("filter"):
        mask &= filter_mask(tbl, q["filter"])
    rows = np.nonzero(mask)[0]
    dims = [q["dimension"]] if qt == "topN" else (q.get("dimensions") or [] if qt == "groupBy" else [])
    dims = [{"type": "default", "dimension": d, "outputName": d} if isinstance(d, str) else d for d in dims]
    bfn = bucket_fn(q.get("granularity", "all"))
    tsel = tbl.time[rows]
    ut, tinv = np.unique(tsel, return_inverse=True)
    key_cols, decoders = [], []
    if bfn is not None:
        b = np.array([bfn(int(t)) for t in ut], dtype=np.int64)
        key_cols.append(b[tinv])
        decoders.append(("timestamp", lambda v: int(v)))
    for d in dims:
        name = d.get("outputName", d["dimension"])
        if d.get("type", "default") == "extraction" and d["dimension"] != "__time":
            # extraction function on a dimension: a string -> string map of the distinct values (docs q8.json groups on
            # timeFormat "yyyy" of o_orderdate); rows group on the mapped value, null = ""
            vals, inv = tbl.dim_strings_of_unique(d["dimension"])
            mapped = [extract_value(d["extractionFn"], v) for v in vals]
            us = sorted(set(mapped), key=lambda s: s.encode("utf-8"))
            rank = {s: i for i, s in enumerate(us)}
            key_cols.append(np.array([rank[s] for s in mapped], dtype=np.int64)[inv][rows])
            decoders.append((name, (lambda us_: lambda v: us_[int(v)])(us)))
        elif d.get("type", "default") == "extraction":
            fn = d["extractionFn"]
            assert fn["type"] == "timeFormat"
            tz = _zone(fn.get("timeZone"))
            strs = [joda_format(fn["format"], int(t), _zoff(tz, int(t))) for t in ut]
            us = sorted(set(strs), key=lambda s: s.encode("utf-8"))
            rank = {s: i for i, s in enumerate(us)}
            key_cols.append(np.array([rank[s] for s in strs], dtype=np.int64)[tinv])
            decoders.append((name, (lambda us_: lambda v: us_[int(v)])(us)))
        else:
            dd = tbl.dims[d["dimension"]]
            if dd[0] == "str":
                # rank codes by UTF-8 byte order of their strings so sort order is string order
                order = sorted(range(len(dd[2])), key=lambda i: dd[2][i].encode("utf-8"))
                rank = np.empty(len(dd[2]), dtype=np.int64)
                rank[order] = np.arange(len(order))
                sorted_strs = [dd[2][i] for i in order]
                key_cols.append(rank[np.asarray(dd[1])[rows]])
                decoders.append((name, (lambda ss: lambda v: ss[int(v)])(sorted_strs)))
            else:
                key_cols.append(np.asarray(dd[1], dtype=np.int64)[rows])
                decoders.append((name, lambda v: str(int(v))))
    if key_cols:
        km = np.stack(key_cols, axis=1)
        uk, gid = np.unique(km, axis=0, return_inverse=True)
        gid = gid.reshape(-1)
        ng = len(uk)
    else:
        uk = np.zeros((1, 0), dtype=np.int64)
        gid = np.zeros(len(rows), dtype=np.int64)
        ng = 1 if len(rows) else 0
    out = [dict() for _ in range(ng)]
    for g in range(ng):
        for c, (name, dec) in enumerate(decoders):
            out[g][name] = dec(uk[g, c])
        out[g]["__key"] = tuple(int(x) for x in uk[g])
    registers = {}
    for a in q.get("aggregations", []):
        sub_rows, sub_gid = rows, gid
        name = a.get("name")
        if a["type"] == "filtered":
            fm = filter_mask(tbl, a["filter"])[rows]
            sub_rows, sub_gid = rows[fm], gid[fm]
            name = a.get("name") or a["aggregator"]["name"]
            a = a["aggregator"]
        t = a["type"]
        if t == "count":
            vals = np.bincount(sub_gid, minlength=ng).tolist()
        elif t == "cardinality" or (t == "hyperUnique" and a["fieldName"] in tbl.dims):
            fld = a["fieldNames"][0] if t == "cardinality" else a["fieldName"]
            regs = hll_registers(tbl, fld, sub_rows, sub_gid, ng)
            registers[name] = regs
            vals = [hll_estimate(regs[g]) for g in range(ng)]
        elif t == "javascript" or "expression" in a:
            # sum / min / max of an arithmetic expression over metrics, row value in IEEE double arithmetic
            # (JavascriptAggregationSpec, DruidQuerySpec.scala:350-360; JSAggGenerator.scala:36-62)
            if t == "javascript":
                params, body = js_subset.parse_function(a["fnAggregate"])
                fold, arg = js_subset.split_fold(params, body)
                cols = [None] + [np.asarray(tbl.metrics[f])[sub_rows].astype(np.float64) for f in a["fieldNames"]]
            else:
                params, arg = js_subset.parse_expression(a["expression"])
                fold = {"doubleSum": "sum", "doubleMin": "min", "doubleMax": "max"}[t]
                cols = [np.asarray(tbl.metrics[f])[sub_rows].astype(np.float64) for f in params]
            v = np.broadcast_to(js_subset.evaluate_numeric(arg, cols), (len(sub_rows),)).astype(np.float64)
            if fold == "sum":
                vals = _fsum_groups(v, sub_gid, ng)
            else:
                acc = np.full(ng, np.inf if fold == "min" else -np.inf)
                (np.minimum if fold == "min" else np.maximum).at(acc, sub_gid, v)
                vals = acc.tolist()
        else:
            col = np.asarray(tbl.metrics[a["fieldName"]])[sub_rows]
            if t == "longSum":
                acc = np.zeros(ng, dtype=np.int64)
                np.add.at(acc, sub_gid, col.astype(np.int64))
                vals = acc.tolist()
            elif t == "doubleSum":
                vals = _fsum_groups(col.astype(np.float64), sub_gid, ng)
            elif t in ("longMin", "longMax", "doubleMin", "doubleMax", "min", "max"):
                is_min = t.lower().endswith("min")
                is_long = t.startswith("long")
                c = col.astype(np.int64) if is_long else col.astype(np.float64)
                init = (np.iinfo(np.int64).max if is_min else np.iinfo(np.int64).min) if is_long else (np.inf if is_min else -np.inf)
                acc = np.full(ng, init, dtype=c.dtype)
                (np.minimum if is_min else np.maximum).at(acc, sub_gid, c)
                vals = acc.tolist()
            else:
                raise NotImplementedError(t)
        for g in range(ng):
            out[g][name] = vals[g]
    return finalize(q, out, registers, [name for name, _ in decoders], return_registers)


def finalize(q, out, registers, key_names, return_registers=False):
    """Post-aggregations, having, ordering and limit over grouped rows (each carrying its sortable "__key")."""
    qt = q["queryType"]
    for p in q.get("postAggregations") or []:
        for r in out:
            r[p["name"]] = _post(p, r)
    if q.get("having"):
        keep = [i for i, r in enumerate(out) if _having(q["having"], r)]
        out = [out[i] for i in keep]
        registers = {k: v[keep] for k, v in registers.items()}
    idx = list(range(len(out)))
    if qt == "topN":
        m = q["metric"]
        inverted = False
        while isinstance(m, dict) and m.get("type") == "inverted":
            inverted = not inverted
            m = m["metric"]
        if isinstance(m, str) or m.get("type") == "numeric":
            col = m if isinstance(m, str) else m["metric"]
            idx.sort(key=lambda i: ((out[i][col] if inverted else -out[i][col]), out[i]["__key"]))
        else:
            d0 = q["dimension"]
            dn = d0 if isinstance(d0, str) else d0.get("outputName", d0["dimension"])
            ps = m.get("previousStop") or ""
            if ps:
                idx = [i for i in idx if (_cmp_bytes(out[i][dn], ps) < 0 if inverted else _cmp_bytes(out[i][dn], ps) > 0)]
            idx.sort(key=lambda i: out[i]["__key"], reverse=inverted)
        idx = idx[:q["threshold"]]
    elif qt == "groupBy" and q.get("limitSpec"):
        ls = q["limitSpec"]
        dim_pos = {name: c for c, name in enumerate(key_names)}
        import functools

        def cmp(i, j):
            for c in ls.get("columns", []):
                cn = c if isinstance(c, str) else c["dimension"]
                asc = True if isinstance(c, str) else c.get("direction", "ascending").lower().startswith("asc")
                if cn in dim_pos:
                    a, b = out[i]["__key"][dim_pos[cn]], out[j]["__key"][dim_pos[cn]]
                else:
                    a, b = out[i][cn], out[j][cn]
                if a != b:
                    return (-1 if a < b else 1) * (1 if asc else -1)
            a, b = out[i]["__key"], out[j]["__key"]
            return (a > b) - (a < b)
        idx.sort(key=functools.cmp_to_key(cmp))
        idx = idx[:ls.get("limit", len(idx))]
    else:
        idx.sort(key=lambda i: out[i]["__key"])
    res = [out[i] for i in idx]
    for r in res:
        r.pop("__key", None)
    if return_registers:
        return res, {k: v[idx] for k, v in registers.items()}
    return res


# ---- partial merge (multi-rank tests): PostAggregate.scala:39-60 --------------------------------------
def merge_partials(parts, key_cols, ops):
    """parts: list of row-dict lists; ops[name] in {'sum','min','max'}; groups matched on key_cols."""
    acc = {}
    for rows in parts:
        for r in rows:
            k = tuple(r[c] for c in key_cols)
            if k not in acc:
                acc[k] = dict(r)
                continue
            t = acc[k]
            for name, op in ops.items():
                t[name] = t[name] + r[name] if op == "sum" else (min(t[name], r[name]) if op == "min" else max(t[name], r[name]))
    return [acc[k] for k in sorted(acc)]
