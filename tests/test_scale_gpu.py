"""Parity at the configurations' own scale (BASELINE.json): SF1 (config 1: Q1-style and the reference's q1.json shape) and
SF10 (config 2: Q5T and Q3L; config 4's shape: group by o_orderkey, 15 M groups), the CUDA engine through the C ABI
against the threaded C oracle (cross-checked with the numpy oracle at SF1 in test_c_oracle_cpu.py) and, for the
high-cardinality shape, against numpy bincount on the raw rows.  Both scan kernels run: the ahead-of-time interpreter
and the run-time specialised one.  At these sizes every persistent block draws tickets, ticket regions run dry, 32-bit
per-thread / per-block counts carry real totals, double sums fold 10^7 addends per group and the group tables leave L2 --
none of which the SF 0.02 tests reach.  Integers, keys, HLL registers and TopN order bit-exact; double sums 1e-9 relative."""
import json
import os

import numpy as np
import pytest

import helpers as H
import queries as Q
from oracle import c_oracle

pytestmark = pytest.mark.gpu

COLS = ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "ps_supplycost", "ps_availqty", "count", "o_orderkey", "s_region", "s_nation",
        "o_orderdate", "c_mktsegment", "o_shippriority", "o_custkey"]
THREADS = os.cpu_count() or 1


def _load(gpu_engine, name, sf, shards):
    from spark_druid_olap_b200 import tpch
    c_oracle.build()
    try:
        gpu_engine.drop_datasource(name)
    except Exception:
        pass
    tabs = tpch.load_into(gpu_engine, name, sf, COLS, shards=tuple(range(shards)), n_shards=shards, keep=True)
    tbl = H.oracle_table(tabs)
    return tabs, tbl, c_oracle.CTable(tbl)


@pytest.fixture(scope="module")
def sf1(gpu_engine):
    tabs, tbl, ct = _load(gpu_engine, "tpch_sf1", 1.0, 1)
    yield gpu_engine, tbl, ct, "tpch_sf1"
    gpu_engine.drop_datasource("tpch_sf1")


@pytest.fixture(scope="module")
def sf10(gpu_engine):
    tabs, tbl, ct = _load(gpu_engine, "tpch_sf10", 10.0, 2)
    yield gpu_engine, tbl, ct, "tpch_sf10"
    gpu_engine.drop_datasource("tpch_sf10")


def _q(spec, ds, jit, **ctx):
    q = json.loads(spec.to_json()) if hasattr(spec, "to_json") else dict(spec)
    q["dataSource"] = ds
    q["context"] = dict(q.get("context") or {"queryId": "scale"}, gpuJit=jit, **ctx)
    return q


def _check(loaded, spec, jit, ordered=False, **ctx):
    eng, tbl, ct, ds = loaded
    q = _q(spec, ds, jit, **ctx)
    res = eng.execute(q)
    assert res.stats["jit"] == jit
    want = c_oracle.execute(ct, q, THREADS)
    assert want is not None
    H.assert_rows_equal(H.engine_rows(res), want, ordered=ordered)
    return res, q


@pytest.mark.parametrize("jit", [False, True])
def test_sf1_q1_style_and_q1_json_shape(sf1, jit):
    """Config 1 at SF1: the Q1-style query and docs/benchmark/druid/queries/q1.json's shape (count, doubleSum, doubleMax,
    longSum, HLL of o_orderkey, a post-aggregation), HLL registers compared too."""
    res, _ = _check(sf1, Q.q1_style(), jit)
    assert res.stats["rowsScanned"] > 5_900_000 and res.n_rows == 4
    res, q = _check(sf1, Q.basic_agg(), jit)
    _, regs = c_oracle.execute(sf1[2], q, THREADS, return_registers=True)
    order = np.lexsort((res["l_linestatus"], res["l_returnflag"]))
    assert np.array_equal(res.hll["alias-7"][order], regs["alias-7"])


@pytest.mark.parametrize("jit", [False, True])
def test_sf10_q5_topn_and_q3_limit(sf10, jit):
    """Config 2 at SF10: Q5T (TopN 10 of s_nation by revenue, selector + bound filter) and Q3L (hash group-by over 3
    dimensions + limit 10 by revenue: the count pre-pass, the hash table beyond L2 and the top-k select)."""
    _check(sf10, Q.q5_topn(), jit, ordered=True)
    res, _ = _check(sf10, Q.q3(10), jit, ordered=True)
    assert res.stats["strategy"] == "hash" and res.stats["deviceTopK"] is True


def _orderkey_expect(tbl):
    k = np.asarray(tbl.dims["o_orderkey"][1])
    cnt = np.bincount(k)
    s = np.bincount(k, weights=np.asarray(tbl.metrics["l_extendedprice"]))
    return cnt, s


@pytest.mark.parametrize("jit,strategy", [(False, None), (True, None), (True, "hash")])
def test_sf10_group_by_orderkey_15m_groups(sf10, jit, strategy):
    """Config 4's shape at SF10: 15 M groups.  The whole result (columnar path) against numpy bincount on the raw rows:
    every key present once, counts exact, sums within 1e-9."""
    eng, tbl, ct, ds = sf10
    cnt, s = _orderkey_expect(tbl)
    res = eng.execute(_q(Q.by_orderkey(strategy), ds, jit))
    keys = np.asarray(res["o_orderkey"], dtype=np.int64)
    assert res.n_rows == int((cnt > 0).sum()) and len(np.unique(keys)) == res.n_rows
    assert np.array_equal(np.asarray(res["n"]), cnt[keys])
    got, want = np.asarray(res["price"]), s[keys]
    assert np.all(np.abs(got - want) <= H.F64_RTOL * np.abs(want))
    assert int(np.asarray(res["n"]).sum()) == tbl.n


@pytest.mark.parametrize("strategy", [None, "hash"])
@pytest.mark.parametrize("order_by,direction,k", [("price", "descending", 10), ("price", "ascending", 1000), ("n", "descending", 25)])
def test_sf10_top_k_of_15m_groups(sf10, strategy, order_by, direction, k):
    """Top k of 15 M groups straight off the direct / hash table (radix select + tournament over the candidates), ties
    broken by key as the oracle breaks them.  Ordering by the row count has 10^6-way ties at the k-th value (counts are
    1..7): the candidate list overflows and the full tournament takes over -- same answer."""
    eng, tbl, ct, ds = sf10
    cnt, s = _orderkey_expect(tbl)
    spec = Q.GroupByQuerySpec(Q.DS, [Q.DefaultDimensionSpec("o_orderkey")], Q.LimitSpec(k, [Q.OrderByColumnSpec(order_by, direction)]), None, "all", None,
                              [Q.agg("count", "n", "count"), Q.agg("doubleSum", "price", "l_extendedprice")], None, Q.ALL_TIME, Q.ctx(strategy))
    res = eng.execute(_q(spec, ds, True))
    assert res.stats["deviceTopK"] is True and res.n_rows == k
    live = np.nonzero(cnt)[0]
    val = (s if order_by == "price" else cnt.astype(np.float64))[live]
    order = np.lexsort((live, -val if direction == "descending" else val))[:k]
    want_keys = live[order]
    assert np.array_equal(np.asarray(res["o_orderkey"], dtype=np.int64), want_keys)
    assert np.array_equal(np.asarray(res["n"]), cnt[want_keys])
    assert np.all(np.abs(np.asarray(res["price"]) - s[want_keys]) <= H.F64_RTOL * np.abs(s[want_keys]))
    assert res.stats["groups"] == len(live)


def test_sf10_q1_style_sums_of_ten_million_addends(sf10):
    """Double sums over ~1.5 x 10^7 addends per group, accumulated per thread, folded per block and across blocks, against
    the C oracle's compensated sums: within 1e-9 relative (hard part 1), counts exact."""
    for jit in (False, True):
        res, _ = _check(sf10, Q.q1_style(), jit)
        assert max(res["count_order"]) > 14_000_000


def test_new_shape_runs_ahead_of_time_code_while_its_kernel_compiles(sf1):
    """A query shape the process has not seen does not wait for NVRTC (about a second): it runs on the ahead-of-time kernel
    while its specialised kernel compiles on a background thread, and later queries of the shape pick that up.  Same rows
    either way.  (context.gpuJit = true is the one way to make a query wait for the compiler.)"""
    import time
    eng, tbl, ct, ds = sf1
    spec = Q.GroupByQuerySpec(Q.DS, [Q.DefaultDimensionSpec("l_linestatus"), Q.DefaultDimensionSpec("s_region")], None, None, "all",
                              Q.SelectorFilterSpec("c_mktsegment", "MACHINERY"),
                              [Q.agg("doubleMax", "mx", "l_discount"), Q.agg("longSum", "avail", "ps_availqty"), Q.agg("doubleSum", "cost", "ps_supplycost")],
                              None, Q.ALL_TIME, None)
    q = json.loads(spec.to_json())
    q["dataSource"] = ds
    want = c_oracle.execute(ct, q, THREADS)
    t0 = time.perf_counter()
    first = eng.execute(q)
    first_ms = (time.perf_counter() - t0) * 1e3
    H.assert_rows_equal(H.engine_rows(first), want)
    if first.stats["jit"]:
        pytest.skip("the shape's cubin was already in the disk cache")
    assert first_ms < 200, f"the first query of a new shape took {first_ms:.0f} ms: it waited for the compiler"
    deadline = time.perf_counter() + 60
    while time.perf_counter() < deadline:
        res = eng.execute(q)
        H.assert_rows_equal(H.engine_rows(res), want)
        if res.stats["jit"]:
            break
        time.sleep(0.05)
    assert res.stats["jit"] is True, "the background compilation never delivered the specialised kernel"
