"""Saved datasources: the encoded segments written from HBM and read back must answer every query identically (rows,
HLL registers, metadata), including a datasource that went through an incremental load with a dictionary merge."""
import json
import os

import numpy as np
import pytest

import helpers as H
import queries as Q
from oracle import druid_oracle as orc

pytestmark = pytest.mark.gpu

SF = 0.01
COLS = H.TEST_COLUMNS + ["order_year"]


def test_save_restore_round_trip(gpu_engine, tmp_path):
    import spark_druid_olap_b200 as sdo
    from spark_druid_olap_b200 import tpch
    eng = gpu_engine
    for n in ("persist_src", "persist_back"):
        try:
            eng.drop_datasource(n)
        except Exception:
            pass
    # two loads with different dictionaries: ids are re-encoded on the device before they are saved
    tabs = tpch.load_into(eng, "persist_src", SF, COLS, shards=(0, 1), n_shards=2, keep=True)
    tbl = H.oracle_table(tabs)
    path = tmp_path / "tpch.sdoseg"
    eng.save_datasource("persist_src", path)
    assert os.path.getsize(path) > tbl.n * 20
    eng.restore_datasource("persist_back", path)
    assert eng.metadata("persist_back") == eng.metadata("persist_src") | {"id": eng.metadata("persist_back").get("id")}
    a, b = eng.segments("persist_src"), eng.segments("persist_back")
    assert [(s["interval"], s["numRows"], s["shardSpec"]["partitionNum"]) for s in a] == [(s["interval"], s["numRows"], s["shardSpec"]["partitionNum"]) for s in b]
    for name, spec, ordered in [("q1", Q.q1_style(), False), ("q3", Q.q3(10), True), ("q5t", Q.q5_topn(), True), ("q7", Q.q7(), False), ("ts", Q.timeseries("P1D"), False),
                                ("okey", Q.by_orderkey(), False), ("zoo", Q.agg_zoo(), False), ("js", Q.js_agg_query(("s_nation",)), False),
                                ("search", Q.search_cases()[2][1], True), ("filter", Q.with_filter(Q.filter_cases()[1][1]), False)]:
        q = json.loads(spec.to_json())
        got = {}
        for ds in ("persist_src", "persist_back"):
            q["dataSource"] = ds
            got[ds] = eng.execute(q)
        want = orc.execute(tbl, q)
        H.assert_rows_equal(H.engine_rows(got["persist_back"]), want, ordered=ordered, what=name)
        # same encoded bytes in HBM: keys and integers identical; double sums differ only by the order in which warps drew tickets
        H.assert_rows_equal(H.engine_rows(got["persist_back"]), H.engine_rows(got["persist_src"]), ordered=ordered, what=name + " restored vs source")
        for k in got["persist_src"].hll:
            assert np.array_equal(got["persist_src"].hll[k], got["persist_back"].hll[k])
    # a restored datasource keeps accepting loads
    more = tpch.FlatTable(SF, COLS, 0, 2)
    before = eng.metadata("persist_back")["numRows"]
    eng.load("persist_back", more.engine_columns(), partition=7)
    assert eng.metadata("persist_back")["numRows"] == before + more.n_rows
    with pytest.raises(sdo.InvalidQuery):
        eng.restore_datasource("persist_back", path)  # name taken
    bad = tmp_path / "bad.sdoseg"
    bad.write_bytes(b"not a datasource at all")
    with pytest.raises(sdo.InvalidQuery):
        eng.restore_datasource("persist_bad", bad)
    cut = tmp_path / "cut.sdoseg"
    cut.write_bytes(path.read_bytes()[: os.path.getsize(path) // 2])
    with pytest.raises(sdo.InvalidQuery):
        eng.restore_datasource("persist_cut", cut)
    with pytest.raises(sdo.DruidDataSourceException):
        eng.restore_datasource("persist_none", tmp_path / "missing.sdoseg")
    for n in ("persist_src", "persist_back"):
        eng.drop_datasource(n)
