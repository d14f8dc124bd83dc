"""The sharded execution path on one GPU (world size 1 over NCCL): sdo_query_execute_sharded (scan -> gather / reduce /
hash-partitioned exchange -> finish, all inside the library) and the sdo_query_partial* entry points behind a host-side
merge (tests/host_merge.py) must give what sdo_query_execute gives.  test_sharded_two_gpus_* run the same with two ranks
when two devices are visible; the 4/8-rank runs use the same code (tools/check_sharded.py, bench.py)."""
import json
import os

import pytest

import helpers as H
import queries as Q
from oracle import druid_oracle as orc

pytestmark = pytest.mark.gpu

SF = 0.02
COLS = ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "o_orderkey", "o_custkey", "o_orderdate", "o_shippriority",
        "c_mktsegment", "s_region", "s_nation", "count", "ps_availqty", "ps_supplycost", "l_shipmode", "c_region"]


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


SHAPES = [(lambda: Q.q1_style(), False), (lambda: Q.q1_style("block"), False), (lambda: Q.q1_style("global"), False),
          (lambda: Q.timeseries("P1M"), False), (lambda: Q.timeseries("P1D"), False), (lambda: Q.q5_topn(), True), (lambda: Q.by_orderkey("hash"), False),
          (lambda: Q.by_orderkey(), False), (lambda: Q.q3(10), True), (lambda: Q.q3(), False)]


@pytest.mark.parametrize("mk,ordered", SHAPES)
def test_sharded_equals_oracle_world1(dist_engine, mk, ordered):
    """In-library merge (one rank) and the host-side merge over the partial API, against the oracle."""
    import host_merge
    from spark_druid_olap_b200 import sharded
    eng, tbl, name = dist_engine
    q = _with_ds(mk(), name)
    want = orc.execute(tbl, q)
    res, info = sharded.execute_sharded(eng, q)
    H.assert_rows_equal(H.engine_rows(res), want, ordered=ordered)
    assert info["merge"] in ("gather", "reduce", "sparse")
    res1, info1 = host_merge.execute_host_merge(eng, q)
    H.assert_rows_equal(H.engine_rows(res1), want, ordered=ordered)
    # same again through the exchange steps (dense: export -> all-gather -> fold; sparse: partition -> all-to-all -> merge)
    res2, info2 = host_merge.execute_host_merge(eng, q, force_exchange=True)
    H.assert_rows_equal(H.engine_rows(res2), want, ordered=ordered)
    if res.hll:
        for k in res.hll:
            assert (res.hll[k] == res2.hll[k]).all()


@pytest.mark.parametrize("merge", ["gather", "reduce", "sparse"])
def test_every_merge_mode_on_one_rank(dist_engine, merge):
    """The three merges of sdo_query_execute_sharded forced onto the same shapes (context.gpuShardMerge): all-gather + fold,
    per-slot ncclReduce over the transposed table (+ uint8 max over the HLL registers), hash-partitioned exchange."""
    from spark_druid_olap_b200 import sharded
    eng, tbl, name = dist_engine
    shapes = [(Q.q1_style(), False), (Q.q5_topn(), True), (Q.q5(), False), (Q.q1_style("global"), False)]
    if merge != "sparse":
        shapes += [(Q.timeseries("P1M"), False), (Q.timeseries("P1D"), False)]  # sketches have no sparse merge
    for spec, ordered in shapes:
        q = _with_ds(spec, name)
        q["context"] = dict(q.get("context") or {"queryId": "m"}, gpuShardMerge=merge)
        res, info = sharded.execute_sharded(eng, q)
        assert info["merge"] == merge
        want, regs = orc.execute(tbl, q, return_registers=True)
        H.assert_rows_equal(H.engine_rows(res), want, ordered=ordered, what=f"{spec.queryType} {merge}")
        for k in regs:
            order = sorted(range(res.n_rows), key=lambda i: res["timestamp"][i])
            assert (res.hll[k][order] == regs[k]).all()


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
    """The sparse path's device steps (split rows by owner, exchange, merge received rows) with a single owner: through the
    partial API with the host's all-to-all, and inside the library with ncclSend / ncclRecv."""
    import host_merge
    from spark_druid_olap_b200 import sharded
    eng, tbl, name = dist_engine
    q = _with_ds(mk(), name)
    want = orc.execute(tbl, q)
    res, info = host_merge.execute_host_merge(eng, q, force_exchange=True)
    assert not info["dense"]
    H.assert_rows_equal(H.engine_rows(res), want, ordered=ordered)
    res, info = sharded.execute_sharded(eng, q)
    assert info["merge"] == "sparse"
    H.assert_rows_equal(H.engine_rows(res), want, ordered=ordered)


def _zoo_without_sketches(strategy):
    q = json.loads(Q.agg_zoo(strategy).to_json())
    q["aggregations"] = [a for a in q["aggregations"] if a["type"] != "cardinality"]
    q["dimensions"] = [{"type": "default", "dimension": "o_custkey", "outputName": "o_custkey"}]
    return q


@pytest.mark.parametrize("shape", ["by_orderkey", "q3", "q3_limit10", "q5", "zoo", "q5_topn"])
def test_row_exchange_on_one_rank(dist_engine, shape):
    """The exchange of raw rows (what a sharded group-by over a sparse key space does: no local aggregation, every surviving
    row travels as a group of one to the rank that owns its key and is aggregated once there), with a single owner.
    The zoo has min / max, cross-type sums and filtered aggregators: a row's slot carries the aggregator's identity where
    its own filter fails."""
    from spark_druid_olap_b200 import sharded
    eng, tbl, name = dist_engine
    spec = {"by_orderkey": lambda: Q.by_orderkey("emit"), "q3": lambda: Q.q3(None, "emit"), "q3_limit10": lambda: Q.q3(10, "emit"), "q5": lambda: Q.q5("emit"),
            "q5_topn": lambda: Q.q5_topn(10, "emit"), "zoo": lambda: _zoo_without_sketches("emit")}[shape]()
    q = _with_ds(spec, name) if hasattr(spec, "to_json") else dict(spec, dataSource=name)
    res, info = sharded.execute_sharded(eng, q)
    assert info["merge"] == "sparse"
    H.assert_rows_equal(H.engine_rows(res), orc.execute(tbl, q), ordered=shape in ("q3_limit10", "q5_topn"), what=shape)


def test_turns_order_concurrent_sharded_queries(dist_engine):
    """sdo_query_execute_sharded_at: queries issued from several threads take the communicator in turn order, whatever
    order the threads arrive in; a turn that was already taken is refused."""
    import threading
    import spark_druid_olap_b200 as sdo
    from spark_druid_olap_b200 import sharded
    eng, tbl, name = dist_engine
    sharded.init_comm(eng)
    base = eng.sharded_next_turn()
    shapes = [Q.q1_style(), Q.q5_topn(), Q.timeseries("P1M"), Q.q3(10), Q.by_orderkey("hash")]
    wants = [orc.execute(tbl, _with_ds(s, name)) for s in shapes]
    out, errors = {}, []

    def run(i):
        try:
            res, _ = sharded.execute_sharded(eng, _with_ds(shapes[i % len(shapes)], name), turn=base + i)
            out[i] = H.engine_rows(res)
        except Exception as ex:  # noqa: BLE001
            errors.append((i, repr(ex)))

    n = 12
    ths = [threading.Thread(target=run, args=(i,)) for i in reversed(range(n))]  # the last turn's thread starts first
    for th in ths:
        th.start()
    for th in ths:
        th.join(120)
    assert not errors, errors
    for i in range(n):
        H.assert_rows_equal(out[i], wants[i % len(shapes)], ordered=shapes[i % len(shapes)].queryType == "topN" or i % len(shapes) == 3)
    assert eng.sharded_next_turn() == base + n
    with pytest.raises(sdo.InvalidQuery):
        sharded.execute_sharded(eng, _with_ds(shapes[0], name), turn=base)


def test_merge_inside_the_library_world1(dist_engine):
    """sdo_query_execute_sharded with a one-rank communicator: scan, NCCL all-gather and fold queued by one C call give
    what sdo_query_execute gives; shapes without a cross-GPU merge are refused before anything is queued."""
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
    assert L.sdo_query_execute_sharded(eng._h, qb, len(qb), 3, C.byref(h)) == sdo.SDO_INVALID and not h
    # cardinality over a sparse key space has no cross-GPU merge: refused before anything is queued
    q = _with_ds(Q.by_orderkey("hash"), name)
    q["aggregations"].append({"type": "cardinality", "name": "d", "fieldNames": ["o_custkey"], "byRow": True})
    qb = sharded.prepare_query(eng, q)
    assert L.sdo_query_execute_sharded(eng._h, qb, len(qb), 0, C.byref(h)) == sdo.SDO_UNSUPPORTED and not h


def test_sharded_two_gpus_against_the_oracle():
    """Two ranks, two GPUs (skipped on a one-GPU box): tools/check_sharded.py loads shard r of 2 on rank r, runs every shape
    through the host-side merge and through sdo_query_execute_sharded (gather, reduce and sparse merges, top-K over the
    exchange) and compares the union of what the ranks return with the oracle over the whole table."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", "29517",
                        os.path.join(root, "tools", "check_sharded.py"), "0.05"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "sharded check: PASS" in r.stdout, (r.stdout[-3000:], r.stderr[-2000:])
