"""Single GPU: the sharded-check shapes on very small tables (few hundred rows per month segment), against the oracle."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H  # noqa: E402
import queries as Q  # noqa: E402
import spark_druid_olap_b200 as sdo  # noqa: E402
from oracle import druid_oracle as orc  # noqa: E402
from spark_druid_olap_b200 import tpch  # noqa: E402

cols = ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "o_orderkey", "o_custkey", "o_orderdate", "o_shippriority",
        "c_mktsegment", "s_region", "s_nation", "count"]
eng = sdo.Engine(0)
bad = 0
for sf, world in ((0.05, 8), (0.05, 5), (0.003, 1), (0.0007, 1)):
    for rank in sorted({0, world - 1, world // 2}):
        name = f"t_{world}_{rank}_{str(sf).replace('.', '_')}"
        eng.create_datasource(name)
        tab = tpch.FlatTable(sf, cols, rank, world)
        eng.load(name, tab.engine_columns(), partition=rank)
        tbl = H.oracle_table([tab])
        for qn, spec in [("q1", Q.q1_style()), ("okey_global", Q.by_orderkey("global")), ("okey_hash", Q.by_orderkey("hash")), ("okey", Q.by_orderkey()),
                         ("ts_day", Q.timeseries("P1D")), ("q3", Q.q3()), ("q5t", Q.q5_topn())]:
            q = json.loads(spec.to_json())
            q["dataSource"] = name
            for jit in (False, True):
                q["context"] = dict(q.get("context") or {}, gpuJit=jit)
                try:
                    H.assert_rows_equal(H.engine_rows(eng.execute(q)), orc.execute(tbl, q), ordered=(qn == "q5t"), what=f"{name} {qn} jit={jit}")
                except AssertionError as e:
                    bad += 1
                    print("FAIL", e)
        print(name, "rows", tab.n_rows, "done", flush=True)
        eng.drop_datasource(name)
print("tiny check:", "PASS" if not bad else f"{bad} FAILURES")
