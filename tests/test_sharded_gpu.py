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
