"""Run one query shape a few times (for ncu / quick timing).  usage: prof_query.py <query> [sf] [iters] [strategy]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import queries as Q  # noqa: E402
import spark_druid_olap_b200 as sdo  # noqa: E402
from spark_druid_olap_b200 import tpch  # noqa: E402

name = sys.argv[1]
sf = float(sys.argv[2]) if len(sys.argv) > 2 else 10.0
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 5
strategy = sys.argv[4] if len(sys.argv) > 4 else None
hll_place = sys.argv[5] if len(sys.argv) > 5 else None
SHAPES = {
    "q5t": (lambda: Q.q5_topn(10, strategy), ["s_region", "o_orderdate", "s_nation", "l_extendedprice"]),
    "q1": (lambda: Q.q1_style(strategy), ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "count"]),
    "q3l": (lambda: Q.q3(10, strategy), ["c_mktsegment", "o_orderdate", "o_orderkey", "o_shippriority", "l_extendedprice"]),
    "tsday": (lambda: Q.timeseries("P1D", True, strategy), ["o_custkey", "l_extendedprice", "count"]),
    "tsmonth": (lambda: Q.timeseries("P1M", True, strategy), ["o_custkey", "l_extendedprice", "count"]),
    "okey": (lambda: Q.by_orderkey(strategy), ["o_orderkey", "l_extendedprice", "count"]),
    "tssum": (lambda: Q.TimeSeriesQuerySpec(Q.DS, Q.ALL_TIME, Q.PeriodGranularitySpec("P1D"), None, [Q.agg("count", "n", "count"), Q.agg("doubleSum", "rev", "l_extendedprice")],
                                            None, Q.ctx(strategy)), ["l_extendedprice", "count"]),
    "tshll": (lambda: Q.TimeSeriesQuerySpec(Q.DS, Q.ALL_TIME, Q.PeriodGranularitySpec("P1D"), None, [Q.agg("count", "n", "count"), Q.CardinalityAggregationSpec("custs", ["o_custkey"])],
                                            None, Q.ctx(strategy)), ["o_custkey", "count"]),
    "q7": (lambda: Q.q7(strategy), ["s_nation", "c_nation", "l_extendedprice"]),
    "year": (lambda: Q.with_extraction(Q.extraction_dims()[0][1], "s_nation", strategy), ["o_orderdate", "s_nation", "l_extendedprice", "o_custkey", "count"]),
    "yearsum": (lambda: Q.GroupByQuerySpec(Q.DS, [Q.extraction_dims()[0][1], Q.DefaultDimensionSpec("s_nation")], None, None, "all", None,
                                           [Q.agg("count", "n", "count"), Q.agg("doubleSum", "price", "l_extendedprice")], None, Q.ALL_TIME, Q.ctx(strategy)),
                ["o_orderdate", "s_nation", "l_extendedprice", "count"]),
    "jsagg": (lambda: Q.js_agg_query(("l_returnflag", "l_linestatus"), None, strategy),
              ["l_returnflag", "l_linestatus", "l_extendedprice", "l_discount", "l_tax", "l_quantity", "ps_availqty", "ps_supplycost", "count"]),
}
mk, cols = SHAPES[name]
eng = sdo.Engine(0)
tpch.load_into(eng, "tpch", sf, cols)
import json  # noqa: E402
q = json.loads(mk().to_json())
if hll_place:
    q["context"] = dict(q.get("context") or {}, gpuHllPlace=hll_place)
q = json.dumps(q).encode()
for i in range(iters):
    r = eng.execute(q)
    s = r.stats
    print(f"{name} sf={sf:g} {s['strategy']} rows={s['rowsScanned']} groups={s['groups']} scan={s['scanKernelMs']:.4f} ms dev={s['deviceMs']:.4f} ms total={s['totalMs']:.4f} ms "
          f"alg={s['algorithmicBytes'] / 1e6:.1f} MB -> {s['algorithmicBytes'] / s['scanKernelMs'] / 1e6:.0f} GB/s launches={s['launches']} plan={s.get('planMs', 0):.3f} assemble={s.get('assembleMs', 0):.3f}")
