"""The run-time specialised scan kernel (NVRTC, csrc/jit.cpp) must give exactly what the ahead-of-time interpreter and
the oracle give: same corpus, context.gpuJit forced on, for every strategy."""
import json

import pytest

import helpers as H
import queries as Q
from oracle import druid_oracle as orc

pytestmark = pytest.mark.gpu

SF = 0.02
COLS = H.TEST_COLUMNS + ["order_year"]


@pytest.fixture(scope="module")
def loaded(gpu_engine):
    from spark_druid_olap_b200 import tpch
    name = "tpch_jit"
    try:
        gpu_engine.drop_datasource(name)
    except Exception:
        pass
    tabs = tpch.load_into(gpu_engine, name, SF, COLS, shards=(0, 1), n_shards=2, keep=True)
    yield gpu_engine, H.oracle_table(tabs), name
    gpu_engine.drop_datasource(name)


CASES = [("q1_lane", lambda: Q.q1_style(), False), ("q1_block", lambda: Q.q1_style("block"), False), ("q1_global", lambda: Q.q1_style("global"), False),
         ("q1_hash", lambda: Q.q1_style("hash"), False), ("basic_agg_hll", lambda: Q.basic_agg(), False), ("proj_filt", lambda: Q.proj_filt_range(), False),
         ("q3_limit", lambda: Q.q3(10), True), ("q5_topn", lambda: Q.q5_topn(), True), ("q7_timeformat", lambda: Q.q7(), False), ("q8", lambda: Q.q8(), False),
         ("ts_day_hll", lambda: Q.timeseries("P1D"), False), ("ts_month_hll", lambda: Q.timeseries("P1M"), False), ("by_orderkey", lambda: Q.by_orderkey(), False),
         ("by_orderkey_hash", lambda: Q.by_orderkey("hash"), False), ("agg_zoo", lambda: Q.agg_zoo(), False), ("agg_zoo_hash", lambda: Q.agg_zoo("hash"), False),
         ("month_grain", lambda: Q.q1_month_grain(), True), ("js_aggs_lane", lambda: Q.js_agg_query(("l_returnflag",)), False),
         ("js_aggs_block", lambda: Q.js_agg_query(("s_nation", "l_shipmode")), False), ("js_aggs_hash", lambda: Q.js_agg_query(("o_orderkey",), None, "hash"), False),
         ("extract_year", lambda: Q.with_extraction(Q.extraction_dims()[0][1]), False), ("extract_js_hash", lambda: Q.with_extraction(Q.extraction_dims()[3][1], "o_orderkey", "hash"), False),
         ("js_aggs_month", lambda: Q.js_agg_query((), Q.js_filter_cases()[0][1], None, {"type": "period", "period": "P1M"}), False)] + \
        [("filter_" + n, (lambda f_: lambda: Q.with_filter(f_))(f), False) for n, f in Q.filter_cases() if n in ("in_many", "not_in", "bound_int_lex", "regex", "or_mix", "js_date_window")]


@pytest.mark.parametrize("name,mk,ordered", CASES, ids=[c[0] for c in CASES])
def test_specialised_kernel_matches_oracle(loaded, name, mk, ordered):
    eng, tbl, ds = loaded
    q = json.loads(mk().to_json())
    q["dataSource"] = ds
    q["context"] = dict(q.get("context") or {}, gpuJit=True)
    res = eng.execute(q)
    assert res.stats["jit"] is True, "the specialised kernel did not run (NVRTC missing?)"
    H.assert_rows_equal(H.engine_rows(res), orc.execute(tbl, q), ordered=ordered, what=name)
    # a second run hits the shape cache: no compilation
    again = eng.execute(q)
    assert again.stats["jit"] is True and again.stats["jitCompileMs"] == 0.0
    q["context"]["gpuJit"] = False
    assert eng.execute(q).stats["jit"] is False
