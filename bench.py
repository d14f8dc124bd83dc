#!/usr/bin/env python
"""Benchmark of the B200 OLAP-index engine on BASELINE.json's metric: TPC-H-flat SF100 rows scanned/s, query ms and
% of the HBM roofline at 1 - 8 B200, with the CPU restatement timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--sf SF] [--scaling strong|weak] [--impl reference]

Workload: the flattened TPC-H table (orderLineItemPartSupplier, tools/TpchBenchMark.scala:43-97) at scale factor SF
(default 100 = the scale the metric is quoted on; 600 M rows, HBM-resident on one B200).  One "step" = one execution of
the Q1-style star-join aggregate (BASELINE configs[0]'s query at the metric's scale: group by l_returnflag, l_linestatus,
ship-date bound, count + three double sums + three averages; 26 B/row, no selective filter, so the algorithmic bytes ARE
the DRAM traffic and the roofline fraction is a fraction).  The other config shapes (Q5-style TopN, Q3-style group-by +
limit, day / month timeseries + HLL(o_custkey), group by o_orderkey, SF1 Q1) are timed in the same run with the
reference's protocol (tools/TpchBenchMark.scala:300-322: several runs, avg / min / max) and reported under "queries".

N > 1 (torchrun, one rank per GPU): strong scaling by default -- the SAME SF100 table split N ways (rank r holds shards
r*L .. r*L+L-1 of N*L); every rank scans its shard, partial aggregates are merged over NCCL inside the library
(sdo_query_execute_sharded), the result is on rank 0.  `--scaling weak` gives every rank an SF-sized table instead.

Parity gate (AbstractTest.scala:127-143 is differential on every run; so is this): before anything is timed, every
reported shape's result over the WHOLE table (all shards of all ranks) is compared with the threaded C restatement of
Spark's plan (oracle/c) run shard by shard on the same generated rows and merged; group keys, counts and TopN order
bit-exact, double sums within 1e-9 relative.  A mismatch prints "parity": "failed" and no value.

Prints ONE JSON line.
"""
import argparse
import json
import math
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
PIECE_SF = 10.0  # rows are generated, loaded and checked in shards of about this scale (60 M rows, ~5 GB of raw host columns)
Q1_COLUMNS = ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount"]
OTHER_COLUMNS = ["s_region", "o_orderdate", "s_nation", "c_mktsegment", "o_orderkey", "o_shippriority", "o_custkey", "l_tax"]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(name, sf):
    """dram__bytes_read + dram__bytes_write per launch of the shape's scan kernel, from an ncu --set full capture kept
    under profiles/ (the file names the commit and scale it was taken at); None when there is no capture for this scale."""
    p = os.path.join(ROOT, "profiles", "r02_traffic.json")
    try:
        for e in json.load(open(p))["captures"]:
            if e["shape"] == name and abs(float(e["sf"]) - float(sf)) < 1e-9:
                return int(e["dram_bytes"]), f"profiles/r02_traffic.json ({e['source']}, commit {e['commit']})"
    except Exception:
        pass
    return None, None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
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


# ---- the query shapes -------------------------------------------------------------------------------------------------
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
    return q


def shapes():
    """name -> (query dict, gate kind).  Gate kinds: "merge" = C oracle per shard, partials merged, every row compared;
    "orderkey" = numpy bincount over all shards for the 1.5 M x SF-group shape (row dicts would not fit);
    "partition0" = the engine restricted to the first shard's segments against the C oracle on that shard."""
    import queries as Q
    top10 = Q.GroupByQuerySpec(Q.DS, [Q.DefaultDimensionSpec("o_orderkey")], Q.LimitSpec(10, [Q.OrderByColumnSpec("price", "descending")]), None, "all", None,
                               [Q.agg("count", "n", "count"), Q.agg("doubleSum", "price", "l_extendedprice")], None, Q.ALL_TIME, None)
    j = lambda s: json.loads(s.to_json())  # noqa: E731
    return {
        "q1_style": (j(Q.q1_style()), "merge"),                     # headline step; configs[0]'s query at SF100
        "q5_topn": (j(Q.q5_topn()), "merge"),                       # configs[1] Q5T
        "q3_limit10": (j(Q.q3(10)), "partition0"),                  # configs[1] Q3L
        "timeseries_day_hll": (j(Q.timeseries("P1D")), "merge"),    # configs[2]
        "timeseries_month_hll": (j(Q.timeseries("P1M")), "merge"),  # configs[2]
        "by_orderkey_top10": (j(top10), "orderkey"),                # configs[3]
        "q1_tpch_expressions": (q1_tpch(), "merge"),
    }


def as_groupby(q):
    """The un-limited group-by behind a topN (the oracle partials of a TopN cannot be merged, its groups can)."""
    if q["queryType"] != "topN":
        return q
    g = {k: v for k, v in q.items() if k not in ("dimension", "threshold", "metric")}
    g["queryType"] = "groupBy"
    g["dimensions"] = [q["dimension"]]
    return g


def finish_topn(q, rows):
    """Order merged group rows as the topN asks (numeric metric descending, ties by key) and cut at the threshold."""
    if q["queryType"] != "topN":
        return rows
    m = q["metric"]
    col = m if isinstance(m, str) else m["metric"]
    d = q["dimension"]
    dn = d if isinstance(d, str) else d.get("outputName", d["dimension"])
    return sorted(rows, key=lambda r: (-r[col], r[dn].encode("utf-8")))[:q["threshold"]]


def rows_equal(got, want, ordered):
    """None when equal (ints / strings bit-exact, doubles within 1e-9 relative), else a description of the first difference."""
    import helpers as H
    try:
        H.assert_rows_equal(got, want, ordered=ordered)
        return None
    except AssertionError as ex:
        return str(ex)[:300]


# ---- CPU side (gate, cpu_baseline, reference arm): the threaded C restatement, one generator shard at a time ----------
class CpuShard:
    def __init__(self, flat):
        import helpers as H
        from oracle import c_oracle
        self.n_rows = flat.n_rows
        self.table = H.oracle_table(flat)
        self.ct = c_oracle.CTable(self.table)
        self.flat = flat


def workload_config(sf_total, rows_total, how):
    return {"workload": f"TPC-H flattened SF{sf_total:g} ({rows_total} rows), Q1-style star-join aggregate: group by l_returnflag, l_linestatus, "
                        f"l_shipdate <= 1998-09-02, count + sum(l_quantity, l_extendedprice, l_discount) + 3 averages (BASELINE configs[0]'s query at the metric's SF100 scale)",
            "sf_total": sf_total, "how": how}


def run_reference(args):
    """--impl reference: the reference's CPU path for the headline query on the SAME rows as the GPU arm (same generator,
    same SF, every shard), all host threads.  Spark / a JVM do not exist on this box, so the timed code is
    the threaded C restatement of Spark's plan (oracle/c; kind "port").  One step = the query over every shard + merge."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import psutil
    from oracle import c_oracle
    from spark_druid_olap_b200 import tpch
    c_oracle.build()
    q = shapes()["q1_style"][0]
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    sf_total = args.sf * (world if args.scaling == "weak" else 1)
    pieces = max(1, math.ceil(sf_total / PIECE_SF))
    # 40 B/row raw + 8 B/row of re-ranked codes; keep within half of what the host has free
    rows_per_piece = 6.0e6 * sf_total / pieces
    fit = int(psutil.virtual_memory().available * 0.5 / (rows_per_piece * 52))
    use = max(1, min(pieces, fit))
    threads = c_oracle.default_threads()
    t_gen = time.perf_counter()
    shards = [CpuShard(tpch.FlatTable(sf_total, Q1_COLUMNS, p, pieces)) for p in range(use)]
    t_gen = time.perf_counter() - t_gen
    n_rows = sum(s.n_rows for s in shards)

    def step():
        return c_oracle.merge_partials(q, [c_oracle.execute_partial(s.ct, q, threads) for s in shards])

    # bounded: when K + W passes over the whole table would take more than a few minutes, a step covers the first shards only
    t0 = time.perf_counter()
    step()
    t1 = time.perf_counter() - t0
    budget = 150.0
    if t1 * (args.steps + args.warmup) > budget and use > 1:
        use = max(1, int(use * budget / (t1 * (args.steps + args.warmup))))
        shards = shards[:use]
        n_rows = sum(s.n_rows for s in shards)
    for _ in range(max(0, args.warmup - 1)):
        step()
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        rows = step()
        times.append(time.perf_counter() - t0)
    ms = 1000 * sum(times) / len(times)
    value = n_rows / (ms / 1000)
    sample = (f"all {pieces} shards of SF{sf_total:g} ({n_rows} rows): the same rows as the GPU arm" if use == pieces else
              f"{use} of {pieces} shards of SF{sf_total:g} ({n_rows} rows; host memory holds no more), same generator and query")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong" if world == 1 else args.scaling, "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": workload_config(sf_total, n_rows, "CPU: threaded C restatement of Spark's plan, shard by shard + merge"),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "result_rows": len(rows), "generate_s": round(t_gen, 1), "same_rows_as_gpu_arm": use == pieces}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--sf", type=float, default=100.0, help="scale factor of the table (strong scaling: in total; weak: per GPU)")
    ap.add_argument("--scaling", choices=["strong", "weak"], default="strong")
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--headline-only", action="store_true", help="skip the other config shapes (profiling runs)")
    ap.add_argument("--no-gate", action="store_true", help="skip the parity gate and the CPU baseline (profiling runs under ncu); the line then carries parity: skipped")
    ap.add_argument("--shape", default="q1_style", help="profiling: which shape the timed step runs (the default is the bench)")
    ap.add_argument("--suite-callers", type=int, nargs="*", default=[1, 4, 16], help="concurrent callers of the suite leg (N = 1)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import ctypes as C

    import numpy as np
    import torch
    import torch.distributed as dist

    import helpers as H
    import spark_druid_olap_b200 as sdo
    from spark_druid_olap_b200 import sharded, tpch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU execution path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    gate = not args.no_gate
    if gate:
        from oracle import c_oracle
        c_oracle.build()
    eng = sdo.Engine(local)
    L = sdo.lib()
    SH = shapes()
    names = [args.shape] if args.headline_only else list(SH)
    if args.shape not in names:
        names.insert(0, args.shape)
    cols = list(Q1_COLUMNS) + ([] if names == ["q1_style"] else OTHER_COLUMNS)

    # ---- load (and, shard by shard while the raw rows are in host memory, the oracle's partial answers) -----------------
    sf_total = args.sf * (world if args.scaling == "weak" else 1)
    pieces = max(1, math.ceil(sf_total / world / PIECE_SF))  # shards held by this rank
    n_shards = world * pieces
    t_load = time.perf_counter()
    eng.create_datasource("tpch")
    n_local = 0
    partials = {n: [] for n in names if SH[n][1] == "merge"}
    ok_cnt = ok_sum = None
    shard0 = None
    t_oracle = 0.0
    cpu_q1 = []
    threads = (os.cpu_count() or 1)
    for piece in range(pieces):
        tab = tpch.FlatTable(sf_total, cols, rank * pieces + piece, n_shards)
        eng.load("tpch", tab.engine_columns(), partition=rank * pieces + piece)
        n_local += tab.n_rows
        if gate:
            t0 = time.perf_counter()
            cs = CpuShard(tab)
            for n in partials:
                t1 = time.perf_counter()
                partials[n].append(c_oracle.execute_partial(cs.ct, as_groupby(SH[n][0]), threads))
                if n == "q1_style":
                    # the CPU baseline: the same call once more, now that the table-side preparation (codes re-ranked by the
                    # dictionary's byte order) is cached -- query work only, as in the reference arm
                    t1 = time.perf_counter()
                    c_oracle.execute_partial(cs.ct, as_groupby(SH[n][0]), threads)
                    cpu_q1.append((time.perf_counter() - t1, tab.n_rows))
            if "by_orderkey_top10" in names:
                k = tab.data["o_orderkey"]
                size = int(1.5e6 * sf_total) * 4 + 64  # sparse order keys: 4 x orders (TPC-H leaves 3 of 4 key blocks unused)
                if ok_cnt is None:
                    ok_cnt, ok_sum = np.zeros(size, dtype=np.int64), np.zeros(size, dtype=np.float64)
                assert int(k.max()) < size
                ok_cnt += np.bincount(k, minlength=size)
                ok_sum += np.bincount(k, weights=tab.data["l_extendedprice"], minlength=size)
            if piece == 0 and rank == 0:
                shard0 = cs  # kept for the partition-0 gate
            t_oracle += time.perf_counter() - t0
        del tab
    t_load = time.perf_counter() - t_load - t_oracle

    if world > 1:
        sharded.init_comm(eng)  # the engine's own NCCL communicator
    prepared = {}

    def qbytes(name):
        """Query JSON bytes as every rank runs them (N > 1: dictionaries aligned, time keys on the global grid)."""
        if name not in prepared:
            q = SH[name][0]
            prepared[name] = sharded.prepare_query(eng, q) if world > 1 else json.dumps(q).encode()
        return prepared[name]

    def run_raw(name):
        """One query through the C ABI.  N = 1: sdo_query_execute.  N > 1: sdo_query_execute_sharded on every rank (scan,
        NCCL merge of the partial aggregates and the finish on rank 0 queued by one C call).  Returns a result handle, or
        None on ranks without a result."""
        qb = qbytes(name)
        h = C.c_void_p()
        if world == 1:
            sdo._check(L.sdo_query_execute(eng._h, qb, len(qb), C.byref(h)))
            return h
        sdo._check(L.sdo_query_execute_sharded(eng._h, qb, len(qb), 0, C.byref(h)))
        return h if h else None

    def run_result(name):
        r = run_raw(name)
        if r is None:
            return None
        return r if isinstance(r, sdo.Result) else sdo.Result(r)

    def free(r):
        if r is not None and not isinstance(r, sdo.Result):
            L.sdo_result_free(r)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- parity gate ---------------------------------------------------------------------------------------------------
    if gate:
        parity = {"status": None,
                  "oracle": "threaded C restatement of Spark's plan (oracle/c), cross-checked against the numpy oracle in tests/; aggregate values "
                            "are not pinned to a live Spark + Druid (none exists here): parity unpinned",
                  "hll_estimator": "no bias tables (registers bit-exact; estimates in 400 < E < 2560 are the raw estimate in engine and oracle alike)",
                  "shapes": {}}
    else:
        parity = {"status": "skipped (--no-gate)"}
    failed = []
    if gate:
        for n in names:
            q, kind = SH[n]
            res = run_result(n)
            diff = None
            if kind == "orderkey":
                # every rank may hold a share of the groups (sparse merge): collect the candidates on rank 0
                mine = H.engine_rows(res) if res is not None else []
                if world > 1:
                    box = [None] * world
                    dist.all_gather_object(box, mine)
                    mine = [r for part in box for r in part]
                    cnt_t, sum_t = torch.from_numpy(ok_cnt).to(dev), torch.from_numpy(ok_sum).to(dev)
                    dist.reduce(cnt_t, 0)
                    dist.reduce(sum_t, 0)
                    ok_cnt, ok_sum = cnt_t.cpu().numpy(), sum_t.cpu().numpy()
                    del cnt_t, sum_t
                if rank == 0:
                    got = sorted(mine, key=lambda r: (-r["price"], int(r["o_orderkey"])))[:10]
                    top = np.argsort(-ok_sum, kind="stable")[:10]  # stable: equal sums keep key order
                    want = [{"o_orderkey": str(int(k)), "n": int(ok_cnt[k]), "price": float(ok_sum[k])} for k in top]
                    diff = rows_equal(got, want, ordered=True)
                    parity["shapes"][n] = {"against": f"numpy bincount over all {n_shards} shards ({int((ok_cnt > 0).sum())} groups)", "ok": diff is None}
            elif kind == "merge":
                box = [partials[n]]
                if world > 1:
                    box = [None] * world
                    dist.all_gather_object(box, partials[n])
                if rank == 0:
                    want = finish_topn(q, c_oracle.merge_partials(as_groupby(q), [p for part in box for p in part]))
                    diff = rows_equal(H.engine_rows(res), want, ordered=q["queryType"] == "topN")
                    parity["shapes"][n] = {"against": f"C oracle over all {n_shards} shards, partials merged ({len(want)} rows)", "ok": diff is None}
            elif rank == 0:  # partition0: rank 0's first shard only, a local query
                q0 = dict(q)
                q0["intervals"] = {"type": "segments", "segments": [{"itvl": iv, "ver": "", "part": 0} for iv in q["intervals"]]}
                r0 = eng.execute(q0)
                want = c_oracle.execute(shard0.ct, q, threads)
                diff = rows_equal(H.engine_rows(r0), want, ordered=True)
                parity["shapes"][n] = {"against": f"C oracle on shard 0 of {n_shards} ({shard0.n_rows} rows); the engine restricted to that shard's segments", "ok": diff is None}
            if rank == 0 and diff is not None:
                parity["shapes"][n]["difference"] = diff
                failed.append(n)
            del res
        partials.clear()
        ok_cnt = ok_sum = shard0 = None
        parity["status"] = "failed" if failed else "exact"
    if world > 1:
        flag = torch.tensor([len(failed)], device=dev)
        dist.broadcast(flag, 0)
        if int(flag[0]):
            failed = failed or ["(rank 0)"]
    if failed:
        if rank == 0:
            print(json.dumps({"metric": METRIC, "value": None, "unit": UNIT, "n_gpus": world, "parity": parity,
                              "error": "parity gate failed for " + ", ".join(failed) + ": no value is reported"}))
        if world > 1:
            dist.destroy_process_group()
        sys.exit(1)

    # ---- the timed headline ----------------------------------------------------------------------------------------------
    head = args.shape
    plan = eng.explain(SH[head][0] if world == 1 else json.loads(qbytes(head)))
    rows_local, alg_local = plan["rowsScanned"], plan["algorithmicBytes"]

    def step():
        h = run_raw(head)
        if h is None:
            return None
        if isinstance(h, sdo.Result):
            return (json.dumps(h.stats).encode(), h.n_rows)
        out = (L.sdo_result_stats_json(h), int(L.sdo_result_num_rows(h)))
        L.sdo_result_free(h)
        return out

    # A shape's first queries run on the ahead-of-time kernel while its specialised kernel compiles in the background
    # (csrc/jit.cpp): touch every shape once, wait for the compiler, then measure the steady state.
    for n in names:
        free(run_raw(n))
    t_settle = time.perf_counter()
    while L.sdo_jit_pending() > 0 and time.perf_counter() - t_settle < 60:
        time.sleep(0.05)
    sync()
    warm = max(args.warmup, 3)
    for _ in range(warm):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    raw_stats, result_rows = [], None
    sync()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = step()
        if r is not None:
            raw_stats.append(r[0])
            result_rows = r[1]
    sync()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    stats = [json.loads(x) for x in raw_stats]
    if world > 1:
        t = torch.tensor([wall], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        wall = float(t[0])
        tot = torch.tensor([rows_local, alg_local], dtype=torch.int64, device=dev)
        dist.all_reduce(tot)
        rows_total, alg_total = int(tot[0]), int(tot[1])
    else:
        rows_total, alg_total = rows_local, alg_local

    # ---- the other shapes, reference protocol: 3 warm-up + 6 timed runs each -----------------------------------------------
    queries = {}
    peak, peak_src = peaks()
    if not args.headline_only:
        for n in names:
            for _ in range(3):
                free(run_raw(n))
            sc, dv, e2 = [], [], []
            st = None
            sync()
            for _ in range(6):
                if world == 1:
                    sync()
                t1 = time.perf_counter()
                r = run_raw(n)
                e2.append((time.perf_counter() - t1) * 1e3)  # the C-ABI call: it returns when the result is in host memory
                if r is not None:
                    res = r if isinstance(r, sdo.Result) else sdo.Result(r)
                    st = res.stats
                    sc.append(st["scanKernelMs"])
                    dv.append(st.get("deviceMs", 0.0))
                    del res
            pl = eng.explain(SH[n][0] if world == 1 else json.loads(qbytes(n)))
            tot = torch.tensor([pl["rowsScanned"], pl["algorithmicBytes"]], dtype=torch.int64, device=dev)
            e2t = torch.tensor([sum(e2) / 6, -min(e2)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tot)
                dist.all_reduce(e2t, op=dist.ReduceOp.MAX)
            if rank == 0 and st is not None:
                scan = sum(sc) / len(sc)
                e = {"strategy": st["strategy"], "rows_scanned": int(tot[0]), "bytes_per_row": pl["bytesPerRow"], "algorithmic_bytes": int(tot[1]),
                     "scan_ms": scan, "scan_ms_min": min(sc), "device_ms": sum(dv) / len(dv), "e2e_ms": float(e2t[0]), "e2e_ms_min": -float(e2t[1]),
                     "rows_per_s_e2e": int(tot[0]) / (float(e2t[0]) / 1e3), "launches": st["launches"], "result_rows": st.get("resultRows")}
                if st.get("shardPhasesMs"):
                    e["merge"], e["phases_ms_rank0"], e["exchange_bytes_rank0"] = st.get("shardMerge"), st["shardPhasesMs"], st.get("exchangeBytes")
                elif st.get("shardMerge"):
                    e["merge"] = st["shardMerge"]
                if world == 1:
                    e["achieved_gbs_algorithmic"] = int(tot[1]) / (scan / 1e3) / 1e9
                    traffic, src = measured_traffic(n, sf_total)
                    if traffic:
                        e["traffic"], e["traffic_source"] = traffic, src
                        e["frac"] = traffic / (scan / 1e3) / 1e9 / peak
                        e["frac_basis"] = "ncu DRAM bytes of the scan kernel / its CUDA-event time / measured HBM peak"
                    elif not SH[n][0].get("filter"):
                        e["frac"] = e["achieved_gbs_algorithmic"] / peak
                        e["frac_basis"] = "algorithmic bytes (no filter: every referenced column byte is read) / scan time / measured HBM peak"
                    else:
                        e["frac"] = None
                        e["frac_basis"] = ("selective filter: 32-byte sectors of key / metric columns without a surviving row are never fetched, so the "
                                           "algorithmic bytes overstate the traffic; no ncu capture at this scale")
                else:
                    # aggregate over the ranks: bytes all ranks scan / the wall time of the whole sharded query
                    e["aggregate_gbs_e2e"] = int(tot[1]) / (float(e2t[0]) / 1e3) / 1e9
                    e["frac_of_aggregate_roofline_e2e"] = e["aggregate_gbs_e2e"] / (peak * world)
                    e["scan_ms_is"] = "rank 0's scan kernel"
                queries[n] = e

    # config 1 at its own scale: SF1, Q1-style and the reference's own q1.json shape, kernel time next to end to end (N = 1)
    sf1 = None
    if not args.headline_only and world == 1:
        import queries as Q
        c1 = ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "ps_supplycost", "ps_availqty", "count", "o_orderkey"]
        tab1 = tpch.FlatTable(1.0, c1, 0, 1)
        eng.create_datasource("tpch_sf1")
        eng.load("tpch_sf1", tab1.engine_columns())
        sf1 = {"rows": tab1.n_rows}
        cs1 = CpuShard(tab1) if gate else None
        for n, spec in (("q1_style", Q.q1_style()), ("q1_json_shape", Q.basic_agg())):
            q = dict(json.loads(spec.to_json()), dataSource="tpch_sf1")
            for _ in range(3):
                eng.execute(q)
            rs = []
            for _ in range(6):
                t1 = time.perf_counter()
                r = eng.execute(q)
                rs.append(((time.perf_counter() - t1) * 1e3, r.stats["scanKernelMs"], r.stats["deviceMs"]))
            e = {"scan_ms": sum(x[1] for x in rs) / 6, "device_ms": sum(x[2] for x in rs) / 6, "e2e_ms": sum(x[0] for x in rs) / 6, "e2e_ms_min": min(x[0] for x in rs),
                 "strategy": r.stats["strategy"], "algorithmic_bytes": r.stats["algorithmicBytes"]}
            e["achieved_gbs_algorithmic"] = e["algorithmic_bytes"] / (e["scan_ms"] / 1e3) / 1e9
            if gate:
                t1 = time.perf_counter()
                want = c_oracle.execute(cs1.ct, q, threads)
                e["cpu_ms"] = (time.perf_counter() - t1) * 1e3
                diff = rows_equal(H.engine_rows(r), want, ordered=False)
                parity["shapes"]["sf1_" + n] = {"against": f"C oracle on the SF1 table ({tab1.n_rows} rows)", "ok": diff is None}
                if diff is not None:
                    parity["status"] = "failed"
                    parity["shapes"]["sf1_" + n]["difference"] = diff
            sf1[n] = e
        eng.drop_datasource("tpch_sf1")
        del tab1, cs1

    # BASELINE config 5's shape on one GPU: the suite from concurrent callers on one engine handle
    # (docs/bi-benchmark/snap-sales-demo.jmx runs 4 thread groups; DruidPlanner.scala:83-90 allows 100 connections)
    suite = None
    if not args.headline_only and world == 1 and args.suite_callers:
        suite = {}
        sq = [(n, qbytes(n)) for n in names if n != "by_orderkey_top10"]
        srows = {n: queries[n]["rows_scanned"] for n, _ in sq}
        for n_threads in args.suite_callers:
            lat = {n: [] for n, _ in sq}
            rows = [0]
            lock = threading.Lock()
            gate_b = threading.Barrier(n_threads + 1)

            def caller(t):
                gate_b.wait()
                for _ in range(4):
                    for k in range(len(sq)):
                        n, qb = sq[(k + t) % len(sq)]
                        # the C-ABI call itself (ctypes releases the GIL for its duration)
                        t1 = time.perf_counter()
                        h = eng.execute_raw(qb)
                        dt = (time.perf_counter() - t1) * 1e3
                        L.sdo_result_free(h)
                        with lock:
                            lat[n].append(dt)
                            rows[0] += srows[n]

            ths = [threading.Thread(target=caller, args=(t,)) for t in range(n_threads)]
            for th in ths:
                th.start()
            gate_b.wait()
            t1 = time.perf_counter()
            for th in ths:
                th.join()
            w = time.perf_counter() - t1
            suite[f"{n_threads}_callers"] = {"queries": sum(len(v) for v in lat.values()), "wall_s": w, "rows_scanned_per_s": rows[0] / w,
                                             "latency_ms_mean": {n: sum(v) / len(v) for n, v in lat.items()}}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    scan_ms = [x["scanKernelMs"] for x in stats]
    dev_ms = [x["deviceMs"] for x in stats]
    launches = sum(x["launches"] for x in stats)
    h2d = sum(x["h2dBytes"] for x in stats)
    d2h = sum(x["d2hBytes"] for x in stats)
    jit = bool(stats and stats[-1].get("jit"))
    e2e_ms = 1000 * wall / args.steps
    # device time of a step, same definition at every N: CUDA events on the engine's stream of the rank that finishes the
    # query (rank 0) from the plan upload to the end of the result download -- at N > 1 that span contains the wait for the
    # other ranks' partials (the NCCL merge on the same stream) and the merge kernel
    ms_per_step = sum(dev_ms) / len(dev_ms)
    value = rows_total / (ms_per_step / 1000)
    scan = sum(scan_ms) / len(scan_ms)
    achieved = alg_local / (scan / 1000) / 1e9  # rank 0's scan kernel over rank 0's bytes
    traffic, traffic_src = measured_traffic(head, sf_total) if world == 1 else (None, None)
    cfg = workload_config(sf_total, rows_total, "GPU: sdo_query_execute" + ("_sharded, partial aggregates merged over NCCL inside the library" if world > 1 else ""))
    cfg.update({"rows_per_gpu": n_local, "rows_scanned_per_step": rows_total, "bytes_per_row": plan["bytesPerRow"], "strategy": plan["strategy"],
                "l2": "inputs larger than L2 (%.0f MB scanned per step per GPU)" % (alg_local / 1e6),
                "parallelism": f"{n_shards} shards of the table over {world} GPU(s), {pieces} per GPU" + ("" if world == 1 else f" ({args.scaling} scaling)"),
                "value_timing": "CUDA events on the engine stream per query: plan upload + scan kernels (+ NCCL merge of partials on the same stream at N > 1) + result download",
                "kernel_variant": "run-time specialised (NVRTC) scan kernel" if jit else "ahead-of-time scan kernel",
                "load_s": round(t_load, 1), "gate_s": round(t_oracle, 1)})
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": args.scaling if world > 1 else "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": cfg,
        "e2e": {"value": rows_total / (e2e_ms / 1000), "unit": UNIT, "ms_per_query": e2e_ms, "h2d_bytes_per_step": h2d // max(1, args.steps),
                "d2h_bytes_per_step": d2h // max(1, args.steps),
                "what": "the C-ABI call, wall clock, max over ranks: query JSON in host memory in, columnar result in host memory out; segments HBM-resident"},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": ("sdo_scan_jit" if jit else "scan_kernel") + f"<{plan['strategy']}>", "kernel_ms": scan, "algorithmic_bytes": alg_local,
                     "bytes_per_row": plan["bytesPerRow"], "rows": rows_local, "peak_source": peak_src,
                     "note": "no selective filter: the algorithmic bytes (rows in the interval x bytes per row) are what the kernel reads"},
        "parity": parity,
        "clocks": clocks,
        "result_rows": result_rows,
    }
    if world > 1:
        line["aggregate"] = {"bytes_per_step": alg_total, "gbs_device": alg_total / (ms_per_step / 1e3) / 1e9, "gbs_e2e": alg_total / (e2e_ms / 1e3) / 1e9,
                             "frac_of_aggregate_roofline_device": alg_total / (ms_per_step / 1e3) / 1e9 / (peak * world),
                             "frac_of_aggregate_roofline_e2e": alg_total / (e2e_ms / 1e3) / 1e9 / (peak * world)}
    if world == 1 and gate and cpu_q1:
        best = min(t / r for t, r in cpu_q1)
        line["cpu_baseline"] = {"value": 1.0 / best, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"the headline query on {len(cpu_q1)} shard(s) of ~{cpu_q1[0][1]} rows each (same generator, same query JSON), best shard; "
                                          f"timed during the parity gate; `bench.py --impl reference` times the whole table"}
    if queries:
        line["queries"] = queries
    if sf1:
        line["config1_sf1"] = sf1
    if suite:
        line["concurrent_suite"] = suite
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
