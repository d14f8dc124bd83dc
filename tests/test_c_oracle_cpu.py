"""The threaded C restatement (timed CPU baseline) against the numpy oracle on the query corpus: the two share the
predicate semantics but not the execution (C: per-thread hash aggregate + Neumaier sums; numpy: unique/fsum)."""
import json

import numpy as np
import pytest

import helpers as H
import queries as Q
from oracle import c_oracle
from oracle import druid_oracle as orc

SF = 0.02


@pytest.fixture(scope="module")
def tables(native_libs):
    from spark_druid_olap_b200 import tpch
    c_oracle.build()
    tbl = H.oracle_table(tpch.FlatTable(SF, H.TEST_COLUMNS + ["order_year"]))
    return tbl, c_oracle.CTable(tbl)


def test_c_xxh64_matches_python(tables):
    L = c_oracle._load()
    for v in (0, 1, -1, 42, 123456789, 2147483647, -2147483648):
        assert L.oc_xxh64_int(v, 42) == orc.spark_hash(v, "int")
    for v in (0, 1, -1, 1 << 40, -(1 << 62), 600000000):
        assert L.oc_xxh64_long(v, 42) == orc.spark_hash(v, "long")


CASES = [("q1", lambda: Q.q1_style(), False), ("basic_agg", lambda: Q.basic_agg(), False), ("ship_range", lambda: Q.ship_date_range(), False),
         ("proj_filt", lambda: Q.proj_filt_range(), False), ("q3", lambda: Q.q3(), False), ("q3_limit", lambda: Q.q3(10), True), ("q5", lambda: Q.q5(), False),
         ("q5_topn", lambda: Q.q5_topn(), True), ("q8", lambda: Q.q8(), False), ("ts_day", lambda: Q.timeseries("P1D"), False),
         ("ts_month", lambda: Q.timeseries("P1M"), False), ("by_orderkey", lambda: Q.by_orderkey(), False), ("agg_zoo", lambda: Q.agg_zoo(), False),
         ("having", lambda: Q.having_query(), True), ("js_aggs", lambda: Q.js_agg_query(), False),
         ("js_aggs_by_nation", lambda: Q.js_agg_query(("s_nation",), Q.SelectorFilterSpec("c_region", "ASIA")), False)] + [("filter_" + n, (lambda f_: lambda: Q.with_filter(f_))(f), False) for n, f in Q.filter_cases() if n != "ts_null_guard"]


@pytest.mark.parametrize("name,mk,ordered", CASES, ids=[c[0] for c in CASES])
def test_c_oracle_equals_numpy_oracle(tables, name, mk, ordered):
    tbl, ct = tables
    q = json.loads(mk().to_json())
    for threads in (1, 5):
        got = c_oracle.execute(ct, q, threads=threads)
        assert got is not None, "shape not covered by the C oracle"
        H.assert_rows_equal(got, orc.execute(tbl, q), ordered=ordered, what=f"{name}[{threads} threads]")


def test_c_oracle_hll_registers_exact(tables):
    tbl, ct = tables
    q = json.loads(Q.timeseries("P1M").to_json())
    _, r1 = c_oracle.execute(ct, q, return_registers=True)
    _, r2 = orc.execute(tbl, q, return_registers=True)
    assert np.array_equal(r1["custs"], r2["custs"])


def test_shapes_outside_c_oracle_fall_back(tables):
    tbl, ct = tables
    assert c_oracle.execute(ct, json.loads(Q.q7().to_json())) is None  # timeFormat dimension: numpy oracle only


def test_c_oracle_equals_numpy_oracle_at_sf1(native_libs):
    """BASELINE config 1's own scale (SF1, 6 M rows): the two independent CPU statements agree on the Q1-style query and on
    the reference's q1.json shape, so the oracle the GPU tests trust at SF1 / SF10 is not only checked on toy tables."""
    from spark_druid_olap_b200 import tpch
    cols = ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "ps_supplycost", "ps_availqty", "count", "o_orderkey"]
    tbl = H.oracle_table(tpch.FlatTable(1.0, cols))
    ct = c_oracle.CTable(tbl)
    for spec in (Q.q1_style(), Q.basic_agg()):
        q = json.loads(spec.to_json())
        got, gr = c_oracle.execute(ct, q, return_registers=True)
        want, wr = orc.execute(tbl, q, return_registers=True)
        H.assert_rows_equal(got, want, what=spec.queryType)
        for k in wr:
            assert np.array_equal(gr[k], wr[k])


@pytest.mark.parametrize("name,mk,ordered", [c for c in CASES if c[0] in ("q1", "q5", "q3", "ts_day", "ts_month", "agg_zoo", "js_aggs", "by_orderkey")], ids=lambda c: c if isinstance(c, str) else "")
def test_merged_shard_partials_equal_the_whole_table(native_libs, name, mk, ordered):
    """bench.py's parity gate runs the C oracle one generator shard at a time and merges the partial groups
    (PostAggregate.scala:39-60: count / sum -> sum, min -> min, max -> max, HLL -> register max): three shards merged must
    give exactly what one run over the concatenated table gives."""
    from spark_druid_olap_b200 import tpch
    cols = H.TEST_COLUMNS + ["order_year"]
    shards = [tpch.FlatTable(SF, cols, i, 3) for i in range(3)]
    whole = c_oracle.CTable(H.oracle_table(shards))
    q = json.loads(mk().to_json())
    parts = [c_oracle.execute_partial(c_oracle.CTable(H.oracle_table(t)), q) for t in shards]
    got, gr = c_oracle.merge_partials(q, parts, return_registers=True)
    want, wr = c_oracle.execute(whole, q, return_registers=True)
    H.assert_rows_equal(got, want, ordered=ordered, what=name)
    for k in wr:
        assert np.array_equal(gr[k], wr[k])
