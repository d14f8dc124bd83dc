"""Throughput of one query shape from T concurrent callers of one engine handle (raw C-ABI calls).
usage: prof_concurrency.py [sf] [threads,threads,...]"""
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import queries as Q  # noqa: E402
import spark_druid_olap_b200 as sdo  # noqa: E402
from spark_druid_olap_b200 import tpch  # noqa: E402
import bench  # noqa: E402

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
threads = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "1,4,16").split(",")]
eng = sdo.Engine(0)
tpch.load_into(eng, "tpch", sf, bench.Q5T_COLUMNS + bench.EXTRA_COLUMNS)
L = sdo.lib()
shapes = [("q1", Q.q1_style()), ("q1_expr", bench.q1_tpch()), ("q3l", Q.q3(10)), ("tsday", Q.timeseries("P1D")), ("tsmonth", Q.timeseries("P1M")),
          ("year", Q.with_extraction(Q.extraction_dims()[0][1], "s_nation")), ("q5t", Q.q5_topn())]
for name, spec in shapes:
    qb = spec.to_json().encode()
    for _ in range(3):
        L.sdo_result_free(eng.execute_raw(qb))
    out = []
    for T in threads:
        n_each = 40
        gate = threading.Barrier(T + 1)

        def run():
            gate.wait()
            for _ in range(n_each):
                L.sdo_result_free(eng.execute_raw(qb))

        ths = [threading.Thread(target=run) for _ in range(T)]
        for th in ths:
            th.start()
        gate.wait()
        t0 = time.perf_counter()
        for th in ths:
            th.join()
        dt = time.perf_counter() - t0
        out.append(f"T={T}: {dt / (T * n_each) * 1e3:.3f} ms/query")
    print(name, "  ".join(out), flush=True)
