"""CPU arm for bench.py (TEST/BENCH INFRASTRUCTURE ONLY): times the oracle's restatement of Spark
executing the un-rewritten query on the host cores.  Spark itself cannot run here (no JVM),
so `kind` is "port".  Uses the multithreaded C restatement (oracle/c/liboracle_c.so) when it is built
and supports the query shape, else the numpy oracle (1 thread)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def prepare(sf, columns, shard=0, n_shards=1):
    import helpers as H
    from spark_druid_olap_b200 import tpch
    flat = tpch.FlatTable(sf, columns, shard, n_shards)
    state = {"table": H.oracle_table(flat), "n_rows": flat.n_rows, "threads": 1, "kind": "port", "impl": "numpy"}
    try:
        from oracle import c_oracle
        if c_oracle.available():
            state["c"] = c_oracle.CTable(state["table"])
            state["threads"] = c_oracle.default_threads()
            state["impl"] = "c"
    except Exception:
        pass
    return state


def run(state, q):
    if "c" in state:
        from oracle import c_oracle
        r = c_oracle.execute(state["c"], q, state["threads"])
        if r is not None:
            return r
    from oracle import druid_oracle as orc
    state["threads"], state["impl"] = (1, "numpy") if "c" not in state else (state["threads"], state["impl"])
    return orc.execute(state["table"], q)


def time_query(q, sf, columns, threads=None, min_runs=2, budget_s=20.0):
    state = prepare(sf, columns)
    if threads:
        state["threads"] = threads
    run(state, q)
    times = []
    t_all = time.perf_counter()
    while len(times) < min_runs or (time.perf_counter() - t_all < budget_s and len(times) < 6):
        t0 = time.perf_counter()
        run(state, q)
        times.append(time.perf_counter() - t0)
    best = min(times)
    return {"value": state["n_rows"] / best, "unit": "rows/s", "cores": state["threads"], "kind": state["kind"], "impl": state["impl"],
            "ms_per_query": 1000 * best, "sample": f"SF{sf:g} shard ({state['n_rows']} rows) of the same generator, same query JSON"}
