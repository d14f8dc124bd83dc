"""The engine handle is shared and re-entrant (SURVEY 8b "Threading": many Spark tasks of one executor call
executeQuery at once; DruidPlanner.scala:83-90 sizes the reference's connection pool for 100 of them).  Threads run the
benchmark suite's shapes in different orders against one engine; every result must equal the single-threaded oracle
answer, including the first-use work that is shared between queries (sketch columns, kernel specialisation, plans)."""
import json
import threading

import pytest

import helpers as H
import queries as Q
from oracle import druid_oracle as orc

pytestmark = pytest.mark.gpu

SF = 0.02
COLS = H.TEST_COLUMNS + ["order_year"]
N_THREADS = 8
ROUNDS = 3


@pytest.fixture(scope="module")
def loaded(gpu_engine):
    from spark_druid_olap_b200 import tpch
    name = "tpch_mt"
    try:
        gpu_engine.drop_datasource(name)
    except Exception:
        pass
    tabs = tpch.load_into(gpu_engine, name, SF, COLS, shards=(0, 1), n_shards=2, keep=True)
    yield gpu_engine, H.oracle_table(tabs), name
    gpu_engine.drop_datasource(name)


def suite():
    # TpchBenchMark.scala:208-288 shapes + the sketch / high-cardinality / expression paths
    return [("q1", Q.q1_style(), False), ("q3", Q.q3(10), True), ("q5", Q.q5(), False), ("q5_topn", Q.q5_topn(), True), ("q7", Q.q7(), False),
            ("q8", Q.q8(), False), ("ts_day_hll", Q.timeseries("P1D"), False), ("basic_agg_hll", Q.basic_agg(), False), ("by_orderkey", Q.by_orderkey(), False),
            ("agg_zoo", Q.agg_zoo(), False), ("js_aggs", Q.js_agg_query(("s_nation",)), False), ("extract_year", Q.with_extraction(Q.extraction_dims()[0][1]), False)]


@pytest.mark.parametrize("jit", [False, True])
def test_concurrent_queries_match_oracle(loaded, jit):
    eng, tbl, ds = loaded
    cases = []
    for name, spec, ordered in suite():
        q = json.loads(spec.to_json())
        q["dataSource"] = ds
        q["context"] = dict(q.get("context") or {}, gpuJit=jit)
        cases.append((name, q, ordered, orc.execute(tbl, q)))
    errors = []
    start = threading.Barrier(N_THREADS)

    def worker(t):
        try:
            start.wait()
            for r in range(ROUNDS):
                for k in range(len(cases)):
                    name, q, ordered, want = cases[(k * (t + 1) + r + t) % len(cases)]
                    q = dict(q, context=dict(q["context"], queryId=f"{name}-{t}-{r}"))
                    res = eng.execute(q)
                    H.assert_rows_equal(H.engine_rows(res), want, ordered=ordered, what=f"{name} thread {t} round {r}")
        except BaseException as ex:  # noqa: BLE001
            errors.append((t, repr(ex)))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(N_THREADS)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(300)
    assert not any(th.is_alive() for th in threads), "a worker is stuck"
    assert not errors, errors[:3]


def test_cancel_one_query_among_many(loaded):
    """sdo_query_cancel names one queryId (DruidRDD.scala:453-481 cancels per task); the others are unaffected."""
    import spark_druid_olap_b200 as sdo
    eng, tbl, ds = loaded
    q = json.loads(Q.q1_style().to_json())
    q["dataSource"] = ds
    want = orc.execute(tbl, q)
    eng.cancel("victim")
    out = {}

    def run(qid):
        try:
            out[qid] = H.engine_rows(eng.execute(dict(q, context={"queryId": qid})))
        except sdo.QueryCancelled:
            out[qid] = "cancelled"

    threads = [threading.Thread(target=run, args=(qid,)) for qid in ("victim", "a", "b", "c")]
    for th in threads:
        th.start()
    for th in threads:
        th.join(120)
    assert out["victim"] == "cancelled"
    for qid in ("a", "b", "c"):
        H.assert_rows_equal(out[qid], want, what=qid)
