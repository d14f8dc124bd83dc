"""The oracle against known answers: public XXH64 vectors, hand-computed aggregates, calendar cases."""
import json
import math

import numpy as np
import pytest

import helpers as H
import queries as Q
from oracle import druid_oracle as orc


def test_xxh64_public_vectors():
    # Known-answer vectors of the XXH64 reference implementation (xxhash.com / xxhsum sanity set)
    assert orc.xxh64(b"", 0) == 0xEF46DB3751D8E999
    assert orc.xxh64(b"a", 0) == 0xD24EC4F1A98C6E5B
    assert orc.xxh64(b"abc", 0) == 0x44BC2CF5AD770999
    # 32+ byte input exercises the 4-lane loop
    assert orc.xxh64(b"Nobody inspects the spammish repetition", 0) == 0xFBCEA83C8A378BF1


def test_spark_hash_int_long_equal_bytes_hash():
    # Spark's XXH64.hashInt / hashLong are XXH64 of the little-endian bytes (seed 42)
    for v in (0, 1, -1, 42, 2147483647, -2147483648):
        assert orc.spark_hash(v, "int") == orc.xxh64(int(v).to_bytes(4, "little", signed=True), 42)
    for v in (0, 1, -1, 1 << 40, -(1 << 62)):
        assert orc.spark_hash(v, "long") == orc.xxh64(int(v).to_bytes(8, "little", signed=True), 42)


def test_hll_register_mapping_and_estimator():
    idx, rho = orc.hll_idx_rho(0xFFFFFFFFFFFFFFFF)
    assert (idx, rho) == (511, 1)
    idx, rho = orc.hll_idx_rho(0)
    assert (idx, rho) == (0, 56)  # 55 zero bits after the index, padded by the 1 at bit p-1
    idx, rho = orc.hll_idx_rho(1 << 54)  # first bit after the 9 index bits
    assert (idx, rho) == (0, 1)
    assert orc.hll_estimate(np.zeros(512, dtype=np.uint8)) == 0.0
    regs = np.zeros(512, dtype=np.uint8)
    regs[:100] = 1
    assert orc.hll_estimate(regs) == float(math.floor(512 * math.log(512 / 412) + 0.5))  # linear counting
    # distinct-count accuracy on a large set (relativeSD 0.05 -> within ~15% w.h.p.)
    regs = np.zeros(512, dtype=np.uint8)
    for v in range(20000):
        i, r = orc.hll_idx_rho(orc.spark_hash(v, "int"))
        regs[i] = max(regs[i], r)
    assert abs(orc.hll_estimate(regs) - 20000) / 20000 < 0.15


def test_iso_instants_and_intervals():
    assert orc.parse_instant("1970-01-01T00:00:00.000Z") == 0
    assert orc.parse_instant("1993-01-01T00:00:00.000") == 725846400000
    assert orc.parse_instant("1991-12-31T16:00:00.000-08:00") == orc.parse_instant("1992-01-01T00:00:00.000Z")
    # QueryIntervals.scala:96-130: "<=" adds 1 ms to the end (FilterPlanningTest "timestamp2")
    a, b = orc.parse_interval("1993-01-01T00:00:00.000Z/1997-08-02T00:00:00.001Z")
    assert b - orc.parse_instant("1997-08-02") == 1


def test_joda_formats_used_by_the_planner():
    t = orc.parse_instant("1995-03-07T13:05:09.123Z")
    f = orc.joda_format
    assert f("yyyy", t) == "1995" and f("YYYY-MM-dd", t) == "1995-03-07" and f("YYYY-MM-dd HH:mm:ss", t) == "1995-03-07 13:05:09"
    assert f("MM", t) == "03" and f("dd", t) == "07" and f("DD", t) == "66" and f("ww", t) == "10"
    assert f("yyyy-MMM", t) == "1995-Mar" and f("EEEE", t) == "Tuesday" and f("yyyy-MM-dd'T'HH", t) == "1995-03-07T13"
    assert f("HH", t, 8 * 3600000) == "21"


def test_granularity_buckets():
    t = orc.parse_instant("1996-02-29T23:59:59.999Z")
    assert orc.bucket_fn("day")(t) == orc.parse_instant("1996-02-29")
    assert orc.bucket_fn({"type": "period", "period": "P1M"})(t) == orc.parse_instant("1996-02-01")
    assert orc.bucket_fn({"type": "period", "period": "P1Y"})(t) == orc.parse_instant("1996-01-01")
    assert orc.bucket_fn({"type": "period", "period": "P1W"})(t) == orc.parse_instant("1996-02-26")  # a Monday
    assert orc.bucket_fn({"type": "duration", "duration": 3600000})(t) == orc.parse_instant("1996-02-29T23:00:00Z")
    assert orc.bucket_fn({"type": "period", "period": "P1D", "timeZone": "+02:00"})(t) == orc.parse_instant("1996-02-29T22:00:00Z")
    assert orc.bucket_fn("all") is None


def _tiny():
    day = 86400000
    t = np.array([0, 0, 1, 1, 2, 40, 40, 41]) * day + orc.parse_instant("1994-01-30")
    dims = {"color": ("str", np.array([0, 1, 0, 2, 1, 0, 3, 1]), ["red", "blue", "", "Zebra"]),
            "k": ("int", np.array([10, 11, 10, 12, 9, 10, 100, 11]))}
    metrics = {"m": np.array([1.5, 2.25, -1.0, 4.0, 0.5, 8.0, 1e16, 3.0]), "c": np.ones(8, dtype=np.int64), "q": np.array([1, 2, 3, 4, 5, 6, 7, 8])}
    return orc.Table(t, dims, metrics)


def test_hand_computed_group_by():
    q = {"queryType": "groupBy", "dataSource": "t", "dimensions": ["color"], "granularity": "all",
         "aggregations": [{"type": "count", "name": "n"}, {"type": "longSum", "name": "q", "fieldName": "q"},
                          {"type": "doubleSum", "name": "m", "fieldName": "m"}, {"type": "doubleMax", "name": "mx", "fieldName": "m"},
                          {"type": "longMin", "name": "qmin", "fieldName": "q"}],
         "postAggregations": [{"type": "arithmetic", "name": "avg", "fn": "/", "fields": [{"type": "fieldAccess", "fieldName": "m"}, {"type": "fieldAccess", "fieldName": "n"}]}]}
    rows = orc.execute(_tiny(), q)
    # string order is UTF-8 byte order: "" < "Zebra" < "blue" < "red"
    assert [r["color"] for r in rows] == ["", "Zebra", "blue", "red"]
    by = {r["color"]: r for r in rows}
    assert by["red"] == {"color": "red", "n": 3, "q": 10, "m": 8.5, "mx": 8.0, "qmin": 1, "avg": 8.5 / 3}
    assert by["blue"]["n"] == 3 and by["blue"]["q"] == 15 and by["blue"]["m"] == 5.75
    assert by[""]["m"] == 4.0 and by["Zebra"]["m"] == 1e16


def test_hand_computed_filters_intervals_topn():
    t = _tiny()
    base = {"queryType": "groupBy", "dataSource": "t", "dimensions": ["k"], "granularity": "all", "aggregations": [{"type": "count", "name": "n"}]}
    sel = orc.execute(t, dict(base, filter={"type": "selector", "dimension": "color", "value": ""}))
    assert sel == [{"k": "12", "n": 1}]
    bnd = orc.execute(t, dict(base, filter={"type": "bound", "dimension": "color", "lower": "a", "lowerStrict": False, "alphaNumeric": False}))
    assert sorted((r["k"], r["n"]) for r in bnd) == [("10", 3), ("11", 2), ("9", 1)]
    lex = orc.execute(t, dict(base, filter={"type": "bound", "dimension": "k", "lower": "10", "upper": "11", "alphaNumeric": False}))
    assert sorted(r["k"] for r in lex) == ["10", "100", "11"]  # lexicographic: "100" lies between "10" and "11"
    num = orc.execute(t, dict(base, filter={"type": "bound", "dimension": "k", "lower": "10", "upper": "11", "alphaNumeric": True}))
    assert sorted(r["k"] for r in num) == ["10", "11"]
    iv = orc.execute(t, dict(base, intervals=["1994-01-31T00:00:00.000Z/1994-02-02T00:00:00.000Z"]))
    assert sorted((r["k"], r["n"]) for r in iv) == [("10", 1), ("12", 1), ("9", 1)]
    top = orc.execute(t, {"queryType": "topN", "dataSource": "t", "dimension": "color", "threshold": 2, "metric": "q", "granularity": "all",
                          "aggregations": [{"type": "longSum", "name": "q", "fieldName": "q"}]})
    assert [(r["color"], r["q"]) for r in top] == [("blue", 15), ("red", 10)]
    ts = orc.execute(t, {"queryType": "timeseries", "dataSource": "t", "granularity": {"type": "period", "period": "P1M"},
                         "aggregations": [{"type": "count", "name": "n"}, {"type": "cardinality", "name": "d", "fieldNames": ["k"], "byRow": True}]})
    assert [(r["timestamp"], r["n"], r["d"]) for r in ts] == [(orc.parse_instant("1994-01-01"), 4, 3.0), (orc.parse_instant("1994-02-01"), 1, 1.0),
                                                              (orc.parse_instant("1994-03-01"), 3, 3.0)]


def test_double_sum_is_exact():
    n = 200000
    rng = np.random.default_rng(1)
    m = rng.uniform(900, 105000, n).round(2)
    tbl = orc.Table(np.zeros(n, dtype=np.int64), {}, {"m": m})
    r = orc.execute(tbl, {"queryType": "timeseries", "dataSource": "t", "granularity": "all", "aggregations": [{"type": "doubleSum", "name": "s", "fieldName": "m"}]})
    assert r[0]["s"] == math.fsum(m.tolist())


def test_generator_is_deterministic_and_shard_invariant():
    from spark_druid_olap_b200 import tpch
    cols = ["o_orderkey", "l_extendedprice", "s_nation", "l_linenumber"]
    whole = tpch.FlatTable(0.01, cols)
    parts = [tpch.FlatTable(0.01, cols, s, 3) for s in range(3)]
    assert sum(p.n_rows for p in parts) == whole.n_rows

    def canon(tabs):
        k = np.concatenate([t.data["o_orderkey"] * 8 + t.data["l_linenumber"] for t in tabs])
        o = np.argsort(k)
        return [np.concatenate([t.data[c] for t in tabs])[o] for c in cols + [tpch.TIME_COLUMN]]
    for a, b in zip(canon([whole]), canon(parts)):
        assert np.array_equal(a, b)
    assert (np.diff(whole.data[tpch.TIME_COLUMN]) >= 0).all()  # sorted by ship day
    again = tpch.FlatTable(0.01, cols)
    assert all(np.array_equal(whole.data[c], again.data[c]) for c in cols)
    # TPC-H shape sanity (flag/status group shares ~ 24.7 : 24.7 : 0.65 : 50 %)
    t = H.oracle_table(tpch.FlatTable(0.05, ["l_returnflag", "l_linestatus", "count"]))
    rows = orc.execute(t, {"queryType": "groupBy", "dataSource": "t", "dimensions": ["l_returnflag", "l_linestatus"], "granularity": "all",
                           "aggregations": [{"type": "count", "name": "n"}]})
    share = {(r["l_returnflag"], r["l_linestatus"]): r["n"] / t.n for r in rows}
    assert abs(share[("A", "F")] - 0.247) < 0.01 and abs(share[("R", "F")] - 0.247) < 0.01 and abs(share[("N", "O")] - 0.50) < 0.01
    assert 0.002 < share[("N", "F")] < 0.012


def test_js_subset_known_answers():
    """The oracle's reading of the JavaScript expression subset against answers ECMAScript fixes (ToBoolean / ToNumber /
    ToString, abstract equality and relational comparison, || and && returning operands)."""
    from oracle import js_subset as J

    def run(src, *args):
        params, body = J.parse_function(src)
        return J.evaluate(body, list(args))

    assert run("function(x) { return(x < '1995-03-15T08:00:00.000Z') }", "1995-03-15") is True
    assert run("function(x) { return(x < '1995-03-15') }", "1995-03-15") is False
    assert run("function(x) { return x < '10' }", "9") is False          # both strings: lexicographic
    assert run("function(x) { return x < 10 }", "9") is True             # string vs number: numeric
    assert run("function(x) { return x == 9 }", "9.0") is True and run("function(x) { return x === 9 }", "9") is False
    assert run("function(x) { return x == null }", None) is True and run("function(x) { return x == 0 }", None) is False
    assert run("function(x) { return x < 1 }", None) is True             # null -> 0 in relational comparison
    assert run("function(x) { return (x) || (\"d\") }", None) == "d" and run("function(x) { return x && x.length }", "abc") == 3.0
    assert run("function(x) { return 'a' + 1 + 2 }", None) == "a12" and run("function(x) { return 1 + 2 + 'a' }", None) == "3a"
    assert run("function(x) { return '' + 0.1 * 3 }", None) == "0.30000000000000004" and run("function(x) { return '' + 1e21 }", None) == "1e+21"
    assert run("function(x) { return '' + 1/3 }", None) == "0.3333333333333333" and run("function(x) { return '' + 100 }", None) == "100"
    assert run("function(x) { return parseInt('42px') + parseFloat('1.5e1x') }", None) == 57.0
    assert math.isnan(run("function(x) { return Number('42px') }", None)) and run("function(x) { return Number('') + Number(' 12 ') }", None) == 12.0
    assert run("function(x) { return x.substr(1, 2) + x.substring(3, 1) + x.slice(-2) + x.charAt(0) }", "abcdef") == "bcbcefa"
    assert run("function(x) { return x.toUpperCase().concat('-', 1) }", "aé") == "Aé-1"
    assert run("function(x) { return !x ? 'empty' : 'full' }", "") == "empty"
    assert run("function(x) { return -x + 5 % 3 - 2 * 3 }", "4") == -8.0
    assert run("function(x) { return Math.max(1, '7', 3) - Math.min(2, 1) }", None) == 6.0
    assert run("function f(a, b) { ; return a.indexOf(b); ; }", "hello", "ll") == 2.0
    for bad in ("function(x) { var y = 1; return y }", "function(x) { return new Date() }", "function(x) { return x.foo() }", "function(x) { return y }"):
        with pytest.raises((J.Unsupported, TypeError)):
            params, body = J.parse_function(bad)
            J.evaluate(body, ["v"])
    # numeric trees over columns
    params, body = J.parse_function("function(current, a, b) { return current + (a *(1 - b)); }")
    fold, arg = J.split_fold(params, body)
    a, b = np.array([10.0, 0.1]), np.array([0.25, 0.3])
    assert fold == "sum" and np.array_equal(J.evaluate_numeric(arg, [None, a, b]), a * (1 - b))
    params, body = J.parse_expression("Math.abs(p - q * 2) / 4")
    assert params == ["p", "q"] and np.array_equal(J.evaluate_numeric(body, [a, b]), np.abs(a - b * 2) / 4)


def test_extraction_function_known_answers():
    ex = orc.extract_value
    assert ex({"type": "timeFormat", "format": "yyyy"}, "1995-03-15") == "1995"
    assert ex({"type": "timeFormat", "format": "yyyy-MM-dd HH", "timeZone": "-08:00"}, "1995-03-15T08:00:00.000Z") == "1995-03-15 00"
    assert ex({"type": "timeFormat", "format": "yyyy"}, "") == ""
    assert ex({"type": "time", "timeFormat": "yyyy-MM-dd", "resultFormat": "dd/MM/yyyy"}, "1998-12-01") == "01/12/1998"
    assert ex({"type": "time", "timeFormat": "yyyyMMdd", "resultFormat": "yyyy-MM"}, "19981201") == "1998-12"
    assert ex({"type": "time", "timeFormat": "yyyy-MM-dd", "resultFormat": "yyyy"}, "not a date") == "not a date"
    assert ex({"type": "javascript", "function": "function(x) { return x.length > 3 ? x.substr(0, 3) : null }"}, "abcdef") == "abc"
    assert ex({"type": "javascript", "function": "function(x) { return x.length > 3 ? x.substr(0, 3) : null }"}, "ab") == ""
    assert ex({"type": "javascript", "function": "function(x) { return Math.floor(x / 10) * 10 }"}, "37") == "30"
    lk = {"type": "lookup", "lookup": {"type": "map", "map": {"a": "1"}}}
    assert ex(lk, "a") == "1" and ex(lk, "b") == "" and ex(dict(lk, retainMissingValue=True), "b") == "b" and ex(dict(lk, replaceMissingValueWith="?"), "b") == "?"
    assert ex({"type": "regex", "expr": "^(\\w+) "}, "PROMO PLATED TIN") == "PROMO" and ex({"type": "regex", "expr": "^(\\d+)"}, "x") == "x"
    assert ex({"type": "regex", "expr": "^(\\d+)", "replaceMissingValue": True, "replaceMissingValueWith": "n/a"}, "x") == "n/a"
    assert ex({"type": "substring", "index": 2, "length": 3}, "abcdef") == "cde" and ex({"type": "substring", "index": 4}, "abcdef") == "ef"
    assert ex({"type": "substring", "index": 9}, "abc") == "" and ex({"type": "upper"}, "aBc") == "ABC" and ex({"type": "strlen"}, "abcd") == "4"


def test_hand_computed_search():
    t = _tiny()  # color: red x3, blue x3, "" x1, Zebra x1; k: 10 x3, 11 x2, 12, 9, 100
    base = {"queryType": "search", "dataSource": "t", "granularity": "all", "searchDimensions": ["color", "k"], "limit": 100}
    rows = orc.execute(t, dict(base, query={"type": "insensitive_contains", "value": ""}))
    assert rows == [{"dimension": "k", "value": "10", "count": 3}, {"dimension": "k", "value": "100", "count": 1}, {"dimension": "k", "value": "11", "count": 2},
                    {"dimension": "k", "value": "12", "count": 1}, {"dimension": "k", "value": "9", "count": 1}, {"dimension": "color", "value": "Zebra", "count": 1},
                    {"dimension": "color", "value": "blue", "count": 3}, {"dimension": "color", "value": "red", "count": 3}]  # the null value is not reported
    rows = orc.execute(t, dict(base, query={"type": "insensitive_contains", "value": "E"}))
    assert [(r["value"], r["count"]) for r in rows] == [("Zebra", 1), ("blue", 3), ("red", 3)]
    rows = orc.execute(t, dict(base, query={"type": "contains", "value": "e", "caseSensitive": True}, sort={"type": "strlen"}, limit=2))
    assert [(r["value"], r["count"]) for r in rows] == [("red", 3), ("blue", 3)]
    rows = orc.execute(t, dict(base, query={"type": "fragment", "values": ["z", "RA"]}, filter={"type": "selector", "dimension": "k", "value": "100"}))
    assert rows == [{"dimension": "color", "value": "Zebra", "count": 1}]


def test_named_time_zones_in_the_oracle():
    """Region ids go through Python's zoneinfo: daylight-saving changes move month / day bucket starts and formats."""
    la = {"type": "period", "period": "P1D", "timeZone": "America/Los_Angeles"}
    f = orc.bucket_fn(la)
    # 1996-04-07 02:00 PST -> 03:00 PDT: the day before starts at 08:00Z, the day after at 07:00Z
    assert f(orc.parse_instant("1996-04-06T20:00:00Z")) == orc.parse_instant("1996-04-06T08:00:00Z")
    assert f(orc.parse_instant("1996-04-08T06:59:59Z")) == orc.parse_instant("1996-04-07T08:00:00Z")   # still the 7th, local 23:59
    assert f(orc.parse_instant("1996-04-08T07:00:00Z")) == orc.parse_instant("1996-04-08T07:00:00Z")
    m = orc.bucket_fn({"type": "period", "period": "P1M", "timeZone": "Europe/Berlin"})
    assert m(orc.parse_instant("1997-07-15T12:00:00Z")) == orc.parse_instant("1997-06-30T22:00:00Z")   # CEST = +02:00
    assert m(orc.parse_instant("1997-01-15T12:00:00Z")) == orc.parse_instant("1996-12-31T23:00:00Z")   # CET = +01:00
    z = orc._zone("America/Los_Angeles")
    assert orc.joda_format("yyyy-MM-dd HH", orc.parse_instant("1995-07-01T06:30:00Z"), orc._zoff(z, orc.parse_instant("1995-07-01T06:30:00Z"))) == "1995-06-30 23"
    assert orc._zoff(orc._zone("Asia/Kolkata"), 0) == 19800000 and orc._zone("+05:30") == 19800000
