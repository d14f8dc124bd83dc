"""Edges of the device evaluators: the largest filter / key / aggregator lists a launch accepts give the oracle's answer,
one more is refused before anything runs; ragged tables (a few rows per month segment, shards of unequal size) and the
sparse row-walk path next to the dense vector path."""
import json

import pytest

import helpers as H
import queries as Q
from oracle import druid_oracle as orc

pytestmark = pytest.mark.gpu

COLS = H.TEST_COLUMNS


@pytest.fixture(scope="module")
def loaded(gpu_engine):
    from spark_druid_olap_b200 import tpch
    name = "tpch_limits"
    try:
        gpu_engine.drop_datasource(name)
    except Exception:
        pass
    tabs = tpch.load_into(gpu_engine, name, 0.01, COLS, shards=(0, 1), n_shards=2, keep=True)
    yield gpu_engine, H.oracle_table(tabs), name
    gpu_engine.drop_datasource(name)


def _run(loaded, q, what, ordered=False):
    eng, tbl, ds = loaded
    q = dict(q, dataSource=ds)
    res = eng.execute(q)
    H.assert_rows_equal(H.engine_rows(res), orc.execute(tbl, q), ordered=ordered, what=what)
    return res


def test_eight_grouping_dimensions_and_one_more(loaded):
    import spark_druid_olap_b200 as sdo
    dims = ["l_returnflag", "l_linestatus", "l_shipmode", "c_mktsegment", "c_region", "s_region", "p_size", "l_linenumber"]
    q = json.loads(Q.GroupByQuerySpec(Q.DS, [Q.DefaultDimensionSpec(d) for d in dims], None, None, "all", None,
                                      [Q.agg("count", "n", "count"), Q.agg("doubleSum", "p", "l_extendedprice")], None, Q.ALL_TIME, None).to_json())
    res = _run(loaded, q, "8 dims")
    assert res.n_rows > 10000
    q["dimensions"].append({"type": "default", "dimension": "s_nation", "outputName": "s_nation"})
    with pytest.raises(sdo.UnsupportedQuery):
        loaded[0].execute(dict(q, dataSource=loaded[2]))


def test_sixteen_aggregations_and_one_more(loaded):
    import spark_druid_olap_b200 as sdo
    metrics = ["l_extendedprice", "l_discount", "l_tax", "l_quantity", "ps_supplycost", "ps_availqty"]
    aggs = [Q.agg(t, f"{t}_{m}", m) for m in metrics for t in ("doubleSum", "doubleMin", "doubleMax")][:16]
    q = json.loads(Q.GroupByQuerySpec(Q.DS, [Q.DefaultDimensionSpec("c_region")], None, None, "all", None, aggs, None, Q.ALL_TIME, None).to_json())
    for strategy in (None, "global", "hash"):
        qq = dict(q, context={"queryId": "x", "gpuStrategy": strategy}) if strategy else q
        _run(loaded, qq, f"16 aggs {strategy}")
    q["aggregations"].append(json.loads(Q.agg("longSum", "one_more", "ps_availqty").to_json()))
    with pytest.raises(sdo.UnsupportedQuery):
        loaded[0].execute(dict(q, dataSource=loaded[2]))


def test_twelve_filter_leaves_and_one_more(loaded):
    import spark_druid_olap_b200 as sdo
    # bitset leaves that do not fold: in-lists on different dimensions combined by or / and / not
    leaves = [Q.ExtractionFilterSpec("s_nation", ["PERU", "CHINA", "KENYA"]), Q.ExtractionFilterSpec("c_nation", ["FRANCE", "INDIA", "IRAQ"]),
              Q.ExtractionFilterSpec("l_shipmode", ["AIR", "SHIP"]), Q.ExtractionFilterSpec("p_brand", ["Brand#11", "Brand#23", "Brand#55"]),
              Q.RegexFilterSpec("p_type", "BRASS$"), Q.StringContainsSpec("p_type", "PROMO"), Q.BoundFilterSpec("o_orderdate", "1994-06-01", False, "1996-02-01", True, False),
              Q.ExtractionFilterSpec("c_mktsegment", ["BUILDING", "MACHINERY"]), Q.ExtractionFilterSpec("p_size", ["3", "17", "44"]),
              Q.ExtractionFilterSpec("l_linenumber", ["1", "5"]), Q.SelectorFilterSpec("l_returnflag", "N"), Q.ExtractionFilterSpec("c_region", ["ASIA", "EUROPE"])]
    f = Q.Or(Q.And(leaves[0], Q.Not(leaves[1]), leaves[2]), Q.And(leaves[3], Q.Or(leaves[4], leaves[5])), Q.And(leaves[6], leaves[7], Q.Not(leaves[8])),
             Q.And(leaves[9], leaves[10], leaves[11]))
    res = _run(loaded, json.loads(Q.with_filter(f).to_json()), "12 leaves")
    assert len(loaded[0].explain(dict(json.loads(Q.with_filter(f).to_json()), dataSource=loaded[2]))["filter"]["leaves"]) == 12
    g = Q.Or(f, Q.ExtractionFilterSpec("s_region", ["AFRICA", "AMERICA"]))
    with pytest.raises(sdo.UnsupportedQuery):
        loaded[0].execute(dict(json.loads(Q.with_filter(g).to_json()), dataSource=loaded[2]))


@pytest.mark.parametrize("jit", [False, True])
def test_alternating_and_or_nesting_twelve_levels(loaded, jit):
    """and(a, or(b, and(c, or(d, ...)))) over 12 leaves: emitted left to right this needs a 12-entry evaluation stack, and
    the device evaluator holds 8 masks (an overflow used to drop the oldest mask silently).  The planner emits the
    deepest child first, so the nesting below needs 2 entries; both kernels must give the oracle's rows."""
    leaves = [Q.ExtractionFilterSpec("s_nation", ["PERU", "CHINA", "KENYA", "JAPAN", "INDIA"]), Q.ExtractionFilterSpec("c_nation", ["FRANCE", "INDIA", "IRAQ"]),
              Q.ExtractionFilterSpec("l_shipmode", ["AIR", "SHIP", "RAIL", "MAIL"]), Q.ExtractionFilterSpec("p_brand", ["Brand#11", "Brand#23", "Brand#55"]),
              Q.Not(Q.RegexFilterSpec("p_type", "BRASS$")), Q.StringContainsSpec("p_type", "PROMO"),
              Q.BoundFilterSpec("o_orderdate", "1993-06-01", False, "1997-02-01", True, False), Q.ExtractionFilterSpec("c_mktsegment", ["BUILDING", "MACHINERY"]),
              Q.Not(Q.ExtractionFilterSpec("p_size", ["3", "17", "44"])), Q.ExtractionFilterSpec("l_linenumber", ["1", "5"]),
              Q.Not(Q.SelectorFilterSpec("l_returnflag", "N")), Q.ExtractionFilterSpec("c_region", ["ASIA", "EUROPE"])]
    f = leaves[-1]
    for depth, leaf in enumerate(reversed(leaves[:-1])):
        f = (Q.And if depth % 2 else Q.Or)(leaf, f)
    q = json.loads(Q.with_filter(f).to_json())
    q["context"] = dict(q.get("context") or {}, gpuJit=jit)
    res = _run(loaded, q, f"12 alternating levels jit={jit}")
    assert 0 < res.n_rows
    ex = loaded[0].explain(dict(q, dataSource=loaded[2]))
    assert len(ex["filter"]["leaves"]) == 12


@pytest.mark.parametrize("sf,world,rank", [(0.05, 8, 7), (0.003, 1, 0), (0.0007, 1, 0), (0.0001, 1, 0)])
def test_ragged_tiny_tables(gpu_engine, sf, world, rank):
    """Few rows per month segment: every span ends inside its first unit, most tickets fall past the end of their span."""
    from spark_druid_olap_b200 import tpch
    eng = gpu_engine
    name = "tpch_tiny"
    try:
        eng.drop_datasource(name)
    except Exception:
        pass
    eng.create_datasource(name)
    tab = tpch.FlatTable(sf, COLS, rank, world)
    eng.load(name, tab.engine_columns(), partition=rank)
    tbl = H.oracle_table([tab])
    try:
        for qn, spec, ordered in [("q1", Q.q1_style(), False), ("okey_global", Q.by_orderkey("global"), False), ("okey_hash", Q.by_orderkey("hash"), False),
                                  ("ts_day", Q.timeseries("P1D"), False), ("ts_block", Q.timeseries("P1M", True, "block"), False), ("q3", Q.q3(), False),
                                  ("q5t", Q.q5_topn(), True), ("q7", Q.q7(), False), ("zoo", Q.agg_zoo(), False)]:
            q = json.loads(spec.to_json())
            q["dataSource"] = name
            for jit in (False, True):
                q["context"] = dict(q.get("context") or {}, gpuJit=jit)
                H.assert_rows_equal(H.engine_rows(eng.execute(q)), orc.execute(tbl, q), ordered=ordered, what=f"{qn} sf={sf} jit={jit}")
    finally:
        eng.drop_datasource(name)
