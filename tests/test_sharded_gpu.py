"""The sharded execution path on one GPU (world size 1 over NCCL): partial -> device views -> merge -> finish must
give what sdo_query_execute gives.  The 2/4/8-rank runs use the same code (tools/check_sharded.py, bench.py)."""
import json
import os

import pytest

import helpers as H
import queries as Q
from oracle import druid_oracle as orc

pytestmark = pytest.mark.gpu

SF = 0.02
COLS = ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "o_orderkey", "o_custkey", "o_orderdate", "o_shippriority",
        "c_mktsegment", "s_region", "s_nation", "count"]


@pytest.fixture(scope="module")
def dist_engine(gpu_engine, tmp_path_factory):
    import torch
    import torch.distributed as dist
    from spark_druid_olap_b200 import tpch
    store = tmp_path_factory.mktemp("pg") / "store"
    created = False
    if not dist.is_initialized():
        dist.init_process_group("nccl", init_method=f"file://{store}", rank=0, world_size=1, device_id=torch.device("cuda", 0))
        created = True
    name = "tpch_sharded"
    try:
        gpu_engine.drop_datasource(name)
    except Exception:
        pass
    tabs = tpch.load_into(gpu_engine, name, SF, COLS, keep=True)
    yield gpu_engine, H.oracle_table(tabs), name
    gpu_engine.drop_datasource(name)
    if created:
        dist.destroy_process_group()


def _with_ds(spec, name):
    q = json.loads(spec.to_json())
    q["dataSource"] = name
    return q


@pytest.mark.parametrize("mk,ordered", [(lambda: Q.q1_style(), False), (lambda: Q.q1_style("block"), False), (lambda: Q.q1_style("global"), False),
                                        (lambda: Q.timeseries("P1M"), False), (lambda: Q.q5_topn(), True), (lambda: Q.by_orderkey("hash"), False),
                                        (lambda: Q.q3(10), True), (lambda: Q.q3(), False)])
def test_sharded_equals_oracle_world1(dist_engine, mk, ordered):
    from spark_druid_olap_b200 import sharded
    eng, tbl, name = dist_engine
    q = _with_ds(mk(), name)
    res, info = sharded.execute_sharded(eng, q)
    want = orc.execute(tbl, q)
    H.assert_rows_equal(H.engine_rows(res), want, ordered=ordered)
    assert info["dense"] == (res.stats["strategy"] != "hash")
    # same again through the exchange steps (dense: export -> all-gather -> fold; sparse: partition -> all-to-all -> merge)
    res2, info2 = sharded.execute_sharded(eng, q, force_exchange=True)
    H.assert_rows_equal(H.engine_rows(res2), want, ordered=ordered)
    if info["dense"] and res.hll:
        for k in res.hll:
            assert (res.hll[k] == res2.hll[k]).all()


def test_time_keys_are_numbered_on_the_global_grid(dist_engine):
    """With context.gpuTimeBoundary (set by the sharded layer) the time-key space comes from the datasource's time grid,
    not from the days this rank happens to hold -- so all ranks build tables of the same shape."""
    from spark_druid_olap_b200 import sharded
    eng, tbl, name = dist_engine
    day = 86400000
    lo, hi = int(tbl.time.min()) - 40 * day, int(tbl.time.max()) + 75 * day  # other ranks hold earlier and later days
    for spec in (Q.timeseries("P1D"), Q.timeseries("P1M"), Q.q7()):
        q = sharded.with_time_boundary(_with_ds(spec, name), (lo, hi))
        if spec.queryType == "groupBy":
            q["filter"] = None
            q["dimensions"] = q["dimensions"][2:]  # just the year(l_shipdate) extraction
            del q["filter"]
        plain = dict(q, context={})
        e_grid, e_local = eng.explain(q), eng.explain(plain)
        assert e_grid["timeKeys"] >= e_local["timeKeys"]
        H.assert_rows_equal(H.engine_rows(eng.execute(q)), orc.execute(tbl, plain))
    q = sharded.with_time_boundary(_with_ds(Q.timeseries("P1D"), name), (lo, hi))
    a, b = orc.parse_interval(Q.ALL_TIME[0])  # the grid is clipped by the query interval
    assert eng.explain(q)["timeKeys"] == (min(hi, b - day) - max(lo, a)) // day + 1


@pytest.mark.parametrize("mk,ordered", [(lambda: Q.by_orderkey("hash"), False), (lambda: Q.q3(), False), (lambda: Q.q3(10), True)])
def test_sparse_partition_exchange_merge_on_one_rank(dist_engine, mk, ordered):
    """The sparse path's device steps (split rows by owner, all-to-all, merge received rows) with a single owner."""
    from spark_druid_olap_b200 import sharded
    eng, tbl, name = dist_engine
    q = _with_ds(mk(), name)
    res, info = sharded.execute_sharded(eng, q, force_exchange=True)
    assert not info["dense"]
    H.assert_rows_equal(H.engine_rows(res), orc.execute(tbl, q), ordered=ordered)


def test_merge_inside_the_library_world1(dist_engine):
    """sdo_query_execute_sharded with a one-rank communicator: scan, NCCL all-gather and fold queued by one C call give
    what sdo_query_execute gives; sparse / large shapes are refused before anything is queued."""
    import ctypes as C
    import spark_druid_olap_b200 as sdo
    from spark_druid_olap_b200 import sharded
    eng, tbl, name = dist_engine
    sharded.init_comm(eng)
    sharded.init_comm(eng)  # idempotent
    L = sdo.lib()
    for mk, ordered in [(lambda: Q.q1_style(), False), (lambda: Q.q1_style("block"), False), (lambda: Q.timeseries("P1D"), False), (lambda: Q.q5_topn(), True),
                        (lambda: Q.q1_style(), "no rows")]:
        q = _with_ds(mk(), name)
        if ordered == "no rows":  # a shard without a matching row still contributes an initialised table
            q["filter"] = {"type": "selector", "dimension": "s_nation", "value": "ATLANTIS"}
            ordered = False
        qb = sharded.prepare_query(eng, q)
        for _ in range(2):
            h = C.c_void_p()
            sdo._check(L.sdo_query_execute_sharded(eng._h, qb, len(qb), 0, C.byref(h)))
            H.assert_rows_equal(H.engine_rows(sdo.Result(h)), orc.execute(tbl, q), ordered=ordered)
    qb = sharded.prepare_query(eng, _with_ds(Q.by_orderkey("hash"), name))
    h = C.c_void_p()
    assert L.sdo_query_execute_sharded(eng._h, qb, len(qb), 0, C.byref(h)) == sdo.SDO_UNSUPPORTED and not h
    assert L.sdo_query_execute_sharded(eng._h, qb, len(qb), 3, C.byref(h)) == sdo.SDO_INVALID
