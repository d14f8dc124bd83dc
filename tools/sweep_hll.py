"""A/B of the shared-memory sketch update (read first vs always atomic) on the timeseries shapes: python tools/sweep_hll.py [sf]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import spark_druid_olap_b200 as sdo  # noqa: E402
from spark_druid_olap_b200 import tpch  # noqa: E402

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 20.0
SH = bench.shapes()
eng = sdo.Engine(0)
eng.create_datasource("tpch")
pieces = max(1, int(sf // 10))
for p in range(pieces):
    t = tpch.FlatTable(sf, ["o_custkey", "l_extendedprice"], p, pieces)
    eng.load("tpch", t.engine_columns(), partition=p)
    del t
for name in ("timeseries_day_hll", "timeseries_month_hll"):
    for rf in ("1", "2", "4"):
        for mb in ("", "5", "6"):
            os.environ["SDO_JIT_NV"] = rf
            os.environ["SDO_JIT_MINBLOCKS"] = mb
            q = dict(SH[name][0])
            q["context"] = dict(q.get("context") or {"queryId": "s"}, gpuJit=True, sweep=f"{rf}/{mb}")
            for _ in range(3):
                eng.execute(q)
            ms = sorted(eng.execute(q).stats["scanKernelMs"] for _ in range(9))
            print(f"{name:24s} nv={rf} minblocks={mb or 'auto':4s} scan_ms median {ms[4]:.4f} min {ms[0]:.4f}", flush=True)
