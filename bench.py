#!/usr/bin/env python
"""Benchmark of the B200 OLAP-index engine on BASELINE.json's metric: rows scanned/s (+ query ms,
% of HBM roofline) on the TPC-H-flattened Q5-style TopN query of configs[1].

    python bench.py [--gpus N] [--steps K] [--warmup W] [--sf SF] [--impl reference]

One "step" = one execution of the Q5T query (TopN 10 of s_nation by revenue, selector s_region +
bound o_orderdate filters) over the HBM-resident segments of this rank.  N > 1 (torchrun, one
rank per GPU): weak scaling -- every rank holds one SF-sized shard of an SF x N dataset; partial
aggregates are merged over NCCL (spark_druid_olap_b200/sharded.py).  Prints ONE JSON line.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "tpch_flat_rows_scanned_per_s"
UNIT = "rows/s"
Q5T_COLUMNS = ["s_region", "o_orderdate", "s_nation", "l_extendedprice"]
Q5T_TRAFFIC_BYTES = 279396352  # ncu --set full capture of sdo_scan_jit on SF10 Q5T (profiles/r01_ncu_q5t_jit_v3_summary.txt): 275.64 MB read + 3.75 MB written
EXTRA_COLUMNS = ["c_mktsegment", "o_orderkey", "o_shippriority", "l_returnflag", "l_linestatus", "l_quantity", "l_discount", "l_tax", "o_custkey"]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 4 + i and r[4 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_baseline_q5t(sf_sample, threads=None):
    """CPU arm: the oracle's restatement of Spark running the un-rewritten Q5T on host cores, on a
    bounded sample (one SF-sized shard of the same generator).  Uses the C restatement when built."""
    import queries as Q
    from oracle import cpu_engine
    return cpu_engine.time_query(json.loads(Q.q5_topn().to_json()), sf_sample, Q5T_COLUMNS, threads)


def q1_tpch():
    """TPC-H Q1 as written (TpchBenchMark.scala:135-150): sum_disc_price and sum_charge are expressions over metrics,
    which the reference can only push down as javascript aggregators (JSAggGenerator.scala)."""
    import queries as Q
    q = json.loads(Q.q1_style().to_json())
    q["aggregations"] += [
        {"type": "javascript", "name": "sum_disc_price", "fieldNames": ["l_extendedprice", "l_discount"],
         "fnAggregate": "function(current, l_extendedprice, l_discount) { return current + ((l_extendedprice) * ((1) - (l_discount))); }",
         "fnCombine": "function(partialA, partialB) { return (partialA + partialB); }", "fnReset": "function() { return 0; }"},
        {"type": "javascript", "name": "sum_charge", "fieldNames": ["l_extendedprice", "l_discount", "l_tax"],
         "fnAggregate": "function(current, l_extendedprice, l_discount, l_tax) { return current + (((l_extendedprice) * ((1) - (l_discount))) * ((1) + (l_tax))); }",
         "fnCombine": "function(partialA, partialB) { return (partialA + partialB); }", "fnReset": "function() { return 0; }"}]

    class _Q:
        def to_json(self):
            return json.dumps(q)
    return _Q()


def run_reference(args):
    """--impl reference: the reference's CPU path.  Spark/JVM do not exist on this box, so
    the timed arm is the oracle's multithreaded C restatement of Spark's plan for the same query."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import queries as Q
    from oracle import cpu_engine
    sf = min(args.sf, args.ref_sf)
    q = json.loads(Q.q5_topn().to_json())
    state = cpu_engine.prepare(sf, Q5T_COLUMNS)
    for _ in range(min(args.warmup, 1)):
        cpu_engine.run(state, q)
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        cpu_engine.run(state, q)
        times.append(time.perf_counter() - t0)
    ms = 1000 * sum(times) / len(times)
    value = state["n_rows"] / (ms / 1000)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": min(args.warmup, 1), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"TPC-H flattened Q5T TopN10 (configs[1]); CPU sample SF{sf:g} = {state['n_rows']} rows per step"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": state["threads"], "kind": state["kind"],
                             "sample": f"SF{sf:g} shard ({state['n_rows']} rows), same generator and query"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--sf", type=float, default=10.0, help="scale factor per GPU (configs[1] = SF10)")
    ap.add_argument("--ref-sf", type=float, default=2.0, help="scale factor of the bounded CPU sample")
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--extra", action="store_true", help="also time the other config shapes (Q1, Q3L, timeseries, by-orderkey)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg (profiling runs)")
    ap.add_argument("--load-shards", type=int, default=1, help="generate and load this rank's rows in this many pieces (bounds host memory at large --sf)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import queries as Q
    import spark_druid_olap_b200 as sdo
    from spark_druid_olap_b200 import sharded, tpch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU execution path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = sdo.Engine(local)
    cols = Q5T_COLUMNS + (EXTRA_COLUMNS if args.extra else [])
    # weak scaling: dataset = SF x world, this rank holds shard `rank` of `world`
    t_load = time.perf_counter()
    eng.create_datasource("tpch")
    n_local = 0
    for piece in range(args.load_shards):
        tab = tpch.FlatTable(args.sf * world, cols, rank * args.load_shards + piece, world * args.load_shards)
        eng.load("tpch", tab.engine_columns(), partition=rank * args.load_shards + piece)
        n_local += tab.n_rows
        del tab
    t_load = time.perf_counter() - t_load

    q5t = Q.q5_topn().to_json().encode()
    plan = eng.explain(q5t)
    rows_local = plan["rowsScanned"]
    alg_bytes = plan["algorithmicBytes"]

    L = sdo.lib()
    if world > 1:
        import ctypes as C
        sharded.init_comm(eng)                       # the engine's own NCCL communicator
        q5t_all = sharded.prepare_query(eng, q5t)    # ids aligned across ranks, time keys on the global grid

    def step():
        """One Q5T query.  N = 1: the C-ABI call itself (query JSON in host memory in, columnar result in host memory out);
        N > 1: sdo_query_execute_sharded on every rank -- scan, NCCL all-gather of the partial tables and the fold on rank 0
        queued by one C call; result on rank 0.  Returns (stats JSON bytes, result rows) or None."""
        if world == 1:
            h = eng.execute_raw(q5t)
            out = (L.sdo_result_stats_json(h), int(L.sdo_result_num_rows(h)))
            L.sdo_result_free(h)
            return out
        h = C.c_void_p()
        sdo._check(L.sdo_query_execute_sharded(eng._h, q5t_all, len(q5t_all), 0, C.byref(h)))
        if not h:
            return None
        out = (L.sdo_result_stats_json(h), int(L.sdo_result_num_rows(h)))
        L.sdo_result_free(h)
        return out

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    raw_stats, result_rows = [], None
    sync()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = step()
        if r is not None:
            raw_stats.append(r[0])
            result_rows = r[1]
    ev1.record()
    sync()
    wall = time.perf_counter() - t0
    ev_ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    stats = [json.loads(x) for x in raw_stats]
    scan_ms = [x["scanKernelMs"] for x in stats]
    dev_ms = [x["deviceMs"] for x in stats]
    launches = sum(x["launches"] for x in stats)
    h2d = sum(x["h2dBytes"] for x in stats)
    d2h = sum(x["d2hBytes"] for x in stats)
    jit = bool(stats and stats[-1].get("jit"))
    if world > 1:
        t = torch.tensor([wall, ev_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        wall, ev_ms = float(t[0]), float(t[1])
        tot = torch.tensor([rows_local], dtype=torch.int64, device="cuda")
        dist.all_reduce(tot)
        rows_total = int(tot[0])
    else:
        rows_total = rows_local
    extra_sharded = None
    if args.extra and world > 1:
        # the other config shapes over all ranks (BASELINE configs[2] and [3] at --sf 12.5 x 8): wall clock per query
        # with a barrier on both sides, max over ranks; dense shapes merge inside the library, by-orderkey goes through
        # the hash-partitioned all-to-all and every rank keeps its share of the groups
        extra_sharded = {}
        top10 = Q.GroupByQuerySpec(Q.DS, [Q.DefaultDimensionSpec("o_orderkey")], Q.LimitSpec(10, [Q.OrderByColumnSpec("price", "descending")]), None, "all", None,
                                   [Q.agg("count", "n", "count"), Q.agg("doubleSum", "price", "l_extendedprice")], None, Q.ALL_TIME, None)
        for name, q in [("q1_style", Q.q1_style()), ("q1_tpch_expressions", q1_tpch()), ("timeseries_day_hll", Q.timeseries("P1D")),
                        ("timeseries_month_hll", Q.timeseries("P1M")), ("by_orderkey_top10", top10)]:
            qj = json.loads(q.to_json())
            for _ in range(3):
                r, info = sharded.execute_sharded(eng, qj)
            sync()
            t1 = time.perf_counter()
            for _ in range(6):
                r, info = sharded.execute_sharded(eng, qj)
            sync()
            t = torch.tensor([(time.perf_counter() - t1) / 6, float(eng.explain(qj)["rowsScanned"])], dtype=torch.float64, device="cuda")
            tmax = t.clone()
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(t)
            extra_sharded[name] = {"ms_per_query": 1000 * float(tmax[0]), "rows_scanned": int(t[1]), "rows_scanned_per_s": float(t[1]) / float(tmax[0]),
                                   "merge": "all-gather + fold inside the library" if info.get("native") else ("all-reduce per slot class" if info["dense"] else "hash-partitioned all-to-all")}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    e2e_ms = 1000 * wall / args.steps
    # device time of a step, same definition at every N: CUDA events on the engine's stream of the rank that finishes
    # the query (rank 0) from the plan upload to the end of the result download -- at N > 1 that span contains the wait for
    # the other ranks' partials (NCCL all-gather on the same stream) and the merge kernel
    ms_per_step = sum(dev_ms) / len(dev_ms)
    value = rows_total / (ms_per_step / 1000)
    scan = sum(scan_ms) / len(scan_ms)
    achieved = alg_bytes / (scan / 1000) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"TPC-H flattened SF{args.sf:g}/GPU, Q5-style TopN 10 of s_nation by sum(l_extendedprice), selector(s_region)+bound(o_orderdate) (BASELINE configs[1])",
                   "rows_per_gpu": n_local, "rows_scanned_per_step": rows_total, "bytes_per_row": plan["bytesPerRow"], "strategy": plan["strategy"],
                   "l2": "inputs larger than L2 (%.0f MB scanned per step per GPU)" % (alg_bytes / 1e6), "parallelism": f"segments sharded over {world} GPU(s)",
                   "value_timing": "CUDA events on the engine stream per query: plan upload + scan kernels (+ NCCL merge of partials on the same stream at N > 1) + result download",
                   "kernel_variant": "run-time specialised (NVRTC) scan kernel" if jit else "ahead-of-time scan kernel",
                   "load_s": round(t_load, 2)},
        "e2e": {"value": rows_total / (e2e_ms / 1000), "unit": UNIT, "ms_per_query": e2e_ms, "h2d_bytes_per_step": h2d // max(1, args.steps),
                "d2h_bytes_per_step": d2h // max(1, args.steps), "what": "sdo_query_execute: query JSON (host) in, columnar result (host) out"},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": Q5T_TRAFFIC_BYTES if abs(args.sf - 10.0) < 1e-9 else None,
                     "traffic_note": "dram__bytes_read+write per launch from profiles/ (ncu --set full): below the algorithmic bytes because 32-byte metric/key sectors without a surviving row are never fetched",
                     "kernel": f"scan_kernel<{plan['strategy']}>", "kernel_ms": scan, "algorithmic_bytes": alg_bytes, "peak_source": peak_src},
        "clocks": clocks,
        "result_rows": result_rows,
        "wall_ms_per_step_cuda_events": ev_ms / args.steps,
    }
    if world == 1 and not args.no_cpu:
        try:
            line["cpu_baseline"] = cpu_baseline_q5t(min(args.sf, args.ref_sf))
        except Exception as ex:  # the baseline is reporting only; never a fallback
            line["cpu_baseline"] = {"error": repr(ex)}
    if args.extra and world == 1:
        extra = {}
        shapes = [("q1_style", Q.q1_style()), ("q1_tpch_expressions", q1_tpch()), ("q3_limit10", Q.q3(10)), ("timeseries_day_hll", Q.timeseries("P1D")),
                  ("timeseries_month_hll", Q.timeseries("P1M")), ("year_of_orderdate_extraction", Q.with_extraction(Q.extraction_dims()[0][1], "s_nation")),
                  ("by_orderkey_top10", Q.GroupByQuerySpec(Q.DS, [Q.DefaultDimensionSpec("o_orderkey")], Q.LimitSpec(10, [Q.OrderByColumnSpec("price", "descending")]),
                                                           None, "all", None, [Q.agg("count", "n", "count"), Q.agg("doubleSum", "price", "l_extendedprice")], None, Q.ALL_TIME, None)),
                  ("by_orderkey_top10_hash", Q.GroupByQuerySpec(Q.DS, [Q.DefaultDimensionSpec("o_orderkey")], Q.LimitSpec(10, [Q.OrderByColumnSpec("price", "descending")]),
                                                                None, "all", None, [Q.agg("count", "n", "count"), Q.agg("doubleSum", "price", "l_extendedprice")], None, Q.ALL_TIME, Q.ctx("hash")))]
        if args.sf <= 20:  # the full 1.5 M x SF-row result goes back to the host: bounded so a large --sf cannot exhaust host memory
            shapes += [("by_orderkey_hash", Q.by_orderkey("hash")), ("by_orderkey", Q.by_orderkey())]
        for name, q in shapes:
            qb = q.to_json().encode()
            for _ in range(3):
                eng.execute(qb)
            rs = [eng.execute(qb) for _ in range(6)]  # reference protocol: 6 runs, avg/min/max (TpchBenchMark.scala:300-322)
            sc = [r.stats["scanKernelMs"] for r in rs]
            tt = [r.stats["totalMs"] for r in rs]
            st = rs[-1].stats
            extra[name] = {"strategy": st["strategy"], "rows_scanned": st["rowsScanned"], "groups": st["groups"], "scan_ms_avg": sum(sc) / 6,
                           "scan_ms_min": min(sc), "total_ms_avg": sum(tt) / 6, "total_ms_min": min(tt), "total_ms_max": max(tt),
                           "achieved_gbs": st["algorithmicBytes"] / (sum(sc) / 6 / 1000) / 1e9, "launches": st["launches"]}
        # BASELINE config 5's shape at this scale: the suite from concurrent callers on one engine handle
        # (docs/bi-benchmark/snap-sales-demo.jmx runs 4 thread groups; DruidPlanner.scala:83-90 allows 100 connections)
        suite = [(n, q.to_json().encode()) for n, q in shapes[:6]] + [("q5_topn", q5t)]  # Q1, Q1 with expressions, Q3, timeseries x2, year extraction, Q5
        suite_rows = {n: eng.execute(qb).stats["rowsScanned"] for n, qb in suite}
        for n_threads in (1, 4, 16):
            lat = {n: [] for n, _ in suite}
            rows = [0]
            lock = threading.Lock()
            gate = threading.Barrier(n_threads + 1)

            def caller(t):
                gate.wait()
                for r in range(4):
                    for k in range(len(suite)):
                        n, qb = suite[(k + t) % len(suite)]
                        # the C-ABI call itself (ctypes releases the GIL for its duration): building Python result
                        # objects under the GIL would serialise the callers and measure the interpreter
                        t0 = time.perf_counter()
                        h = eng.execute_raw(qb)
                        dt = (time.perf_counter() - t0) * 1e3
                        L.sdo_result_free(h)
                        with lock:
                            lat[n].append(dt)
                            rows[0] += suite_rows[n]

            ths = [threading.Thread(target=caller, args=(t,)) for t in range(n_threads)]
            for th in ths:
                th.start()
            gate.wait()
            t0 = time.perf_counter()
            for th in ths:
                th.join()
            wall = time.perf_counter() - t0
            extra[f"concurrent_suite_{n_threads}_callers"] = {"queries": sum(len(v) for v in lat.values()), "wall_s": wall, "rows_scanned_per_s": rows[0] / wall,
                                                             "latency_ms_mean": {n: sum(v) / len(v) for n, v in lat.items()}}
        line["extra"] = extra
    if extra_sharded is not None:
        line["extra"] = extra_sharded
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
