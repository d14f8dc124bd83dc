"""Sweep the specialised scan kernel's build knobs (blocks per SM the compiler aims for, vectors per accumulate call, rows
per ticket) over bench shapes on one resident table:  python tools/sweep_jit.py [sf] [shape ...]  -> one line per variant."""
import itertools
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import spark_druid_olap_b200 as sdo  # noqa: E402
from spark_druid_olap_b200 import tpch  # noqa: E402

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 20.0
names = sys.argv[2:] or ["timeseries_day_hll", "q5_topn"]
SH = bench.shapes()
eng = sdo.Engine(0)
eng.create_datasource("tpch")
pieces = max(1, int(sf // 10))
for p in range(pieces):
    t = tpch.FlatTable(sf, bench.Q1_COLUMNS + bench.OTHER_COLUMNS, p, pieces)
    eng.load("tpch", t.engine_columns(), partition=p)
    del t
for name in names:
    base = SH[name][0]
    for mb, nv, tr in itertools.product(["", "3", "4", "5", "6", "8"], ["1", "2", "4"], [0, 1024, 2048, 4096]):
        os.environ["SDO_JIT_MINBLOCKS"] = mb
        os.environ["SDO_JIT_NV"] = nv
        q = dict(base)
        q["context"] = dict(q.get("context") or {"queryId": "sweep"}, gpuJit=True, sweep=f"{mb}/{nv}/{tr}")
        if tr:
            q["context"]["gpuTicketRows"] = tr
        try:
            for _ in range(3):
                eng.execute(q)
            ms = sorted(eng.execute(q).stats["scanKernelMs"] for _ in range(9))
            print(f"{name:24s} minblocks={mb or 'auto':4s} nv={nv} ticket={tr or 'auto':>5} scan_ms median {ms[4]:.4f} min {ms[0]:.4f}", flush=True)
        except Exception as ex:  # noqa: BLE001
            print(f"{name:24s} minblocks={mb or 'auto':4s} nv={nv} ticket={tr or 'auto':>5} failed: {str(ex)[:100]}", flush=True)
