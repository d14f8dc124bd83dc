"""Committed answers (tests/golden/results_sf0005.json, written by tests/golden/make_result_fixtures.py): the numpy
oracle must still give them on the regenerated table -- a moved answer means the generator or the oracle changed -- and
the threaded C restatement must give them for every shape it covers."""
import importlib.util
import json
import os

import pytest

import helpers as H
from oracle import c_oracle
from oracle import druid_oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_result_fixtures", os.path.join(HERE, "golden", "make_result_fixtures.py"))
fixtures = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(fixtures)
GOLDEN = json.load(open(os.path.join(HERE, "golden", "results_sf0005.json")))


@pytest.fixture(scope="module")
def tables():
    from spark_druid_olap_b200 import tpch
    tabs = [tpch.FlatTable(fixtures.SF, fixtures.COLS, r, 2) for r in range(2)]
    tbl = H.oracle_table(tabs)
    c_oracle.build()
    return tbl, c_oracle.CTable(tbl)


def test_generator_row_count_is_the_committed_one(tables):
    assert tables[0].n == GOLDEN["rows"] and GOLDEN["sf"] == fixtures.SF


@pytest.mark.parametrize("name,spec,ordered", fixtures.cases(), ids=[c[0] for c in fixtures.cases()])
def test_oracles_reproduce_the_committed_answers(tables, name, spec, ordered):
    tbl, ct = tables
    q = json.loads(spec.to_json())
    want = GOLDEN["results"][name]
    assert want["ordered"] == ordered
    H.assert_rows_equal(orc.execute(tbl, q), want["rows"], ordered=ordered, what=f"numpy oracle vs golden {name}")
    got = c_oracle.execute(ct, q, threads=3)
    if got is not None:
        H.assert_rows_equal(got, want["rows"], ordered=ordered, what=f"C oracle vs golden {name}")
    else:
        assert name in ("q7_time_format", "year_extraction", "search_two_dims", "timeseries_month_hll", "q8"), f"{name}: not covered by the C oracle"
