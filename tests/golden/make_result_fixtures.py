"""Writes tests/golden/results_sf0005.json: the numpy oracle's answers for a fixed list of query shapes on the
deterministic generator's SF 0.005 table (seed 20261106, 2 shards).  Committed so that (a) a change to the oracle or to
the generator that moves an answer shows up as a diff of this file, (b) the C oracle and the CUDA engine are compared
with answers that no code in this run produced.  Re-run only on purpose:  python tests/golden/make_result_fixtures.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H  # noqa: E402
import queries as Q  # noqa: E402
from oracle import druid_oracle as orc  # noqa: E402
from spark_druid_olap_b200 import tpch  # noqa: E402

SF = 0.005
COLS = H.TEST_COLUMNS + ["order_year"]


def cases():
    return [("q1_style", Q.q1_style(), False), ("basic_agg_hll", Q.basic_agg(), False), ("ship_date_range", Q.ship_date_range(), False),
            ("q3_limit10", Q.q3(10), True), ("q5_topn", Q.q5_topn(), True), ("q7_time_format", Q.q7(), False), ("q8", Q.q8(), False),
            ("timeseries_month_hll", Q.timeseries("P1M"), False), ("agg_zoo", Q.agg_zoo(), False), ("having_ordered", Q.having_query(), True),
            ("js_aggs", Q.js_agg_query(("l_returnflag", "l_linestatus")), False), ("year_extraction", Q.with_extraction(Q.extraction_dims()[0][1]), False),
            ("search_two_dims", Q.search_cases()[2][1], True), ("filter_regex", Q.with_filter(dict(Q.filter_cases())["regex"]), False),
            ("filter_js_window", Q.with_filter(dict(Q.filter_cases())["js_date_window"]), False)]


def table():
    return H.oracle_table([tpch.FlatTable(SF, COLS, r, 2) for r in range(2)])


if __name__ == "__main__":
    tbl = table()
    out = {"sf": SF, "rows": int(tbl.n), "results": {}}
    for name, spec, ordered in cases():
        q = json.loads(spec.to_json())
        out["results"][name] = {"ordered": ordered, "rows": orc.execute(tbl, q)}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "results_sf0005.json")
    json.dump(out, open(path, "w"), indent=0, sort_keys=True)
    print(path, os.path.getsize(path), "bytes", out["rows"], "rows")
