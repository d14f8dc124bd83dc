"""BASELINE config 5: the TPC-H-flattened suite Q1, Q3, Q5, Q7, Q8, Q10 (tools/TpchBenchMark.scala:208-288) run concurrently
over a table resident across N GPUs.

    torchrun --nproc-per-node N tools/suite_sharded.py [sf_total] [callers ...]  ->  one JSON line on rank 0

Every rank holds shards r*L .. r*L+L-1 of the SF table.  Each of T caller threads per rank walks the suite (rotated by its
index, so different shapes are in flight together -- docs/bi-benchmark/snap-sales-demo.jmx runs 4 thread groups against a
pool of 100 connections, DruidPlanner.scala:83-90) through sdo_query_execute_sharded_at: scans of different queries overlap,
their NCCL merges are issued in turn order on every rank.  Before anything is timed every shape's merged answer is compared
with the CPU oracles run shard by shard on the same rows and merged (C oracle; numpy oracle for Q7's timeFormat dimension;
numpy bincount for the large results).

Q10 groups by o_custkey in place of c_name, c_address, c_phone, c_comment: the reference declares c_name -> the others as
functional dependencies (BaseTest.scala:40-49) and c_name is "Customer#<custkey>" -- one key column carries the group.
"""
import ctypes as C
import json
import math
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402
import helpers as H  # noqa: E402
import queries as Q  # noqa: E402
import spark_druid_olap_b200 as sdo  # noqa: E402
from oracle import c_oracle  # noqa: E402
from oracle import druid_oracle as orc  # noqa: E402
from spark_druid_olap_b200 import sharded, tpch  # noqa: E402

COLS = ["l_returnflag", "l_linestatus", "l_extendedprice", "ps_supplycost", "ps_availqty", "count", "o_orderkey", "c_mktsegment", "o_orderdate", "o_shippriority",
        "s_region", "s_nation", "c_nation", "c_region", "p_type", "order_year", "o_custkey"]


def suite():
    j = lambda s: json.loads(s.to_json())  # noqa: E731
    q10 = Q.GroupByQuerySpec(Q.DS, [Q.DefaultDimensionSpec("o_custkey"), Q.DefaultDimensionSpec("c_nation")], None, None, "all",
                             Q.And(Q.BoundFilterSpec("o_orderdate", lower="1993-10-01", lowerStrict=False, upper="1994-01-01", upperStrict=True),
                                   Q.SelectorFilterSpec("l_returnflag", "R")),
                             [Q.agg("doubleSum", "price", "l_extendedprice")], None, Q.ALL_TIME, None)
    return {"q1": j(Q.basic_agg()), "q3": j(Q.q3()), "q5": j(Q.q5()), "q7": j(Q.q7()), "q8": j(Q.q8()), "q10": j(q10)}


def log(*a):
    print(f"[rank {os.environ.get('RANK')}] {time.strftime('%H:%M:%S')}", *a, file=sys.stderr, flush=True)


def main():
    import faulthandler
    faulthandler.dump_traceback_later(int(os.environ.get("SUITE_WATCHDOG_S", "600")), exit=True)  # a stuck collective must not hold the GPUs
    sf = float(sys.argv[1]) if len(sys.argv) > 1 else 300.0
    callers = [int(x) for x in sys.argv[2:]] or [1, 4, 16]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    c_oracle.build()
    eng = sdo.Engine(local)
    L = sdo.lib()
    SU = suite()
    threads = os.cpu_count() or 1
    pieces = max(1, math.ceil(sf / world / bench.PIECE_SF))
    n_shards = world * pieces
    eng.create_datasource("tpch")
    t_load = time.perf_counter()
    n_local = 0
    small = {n: [] for n in ("q1", "q5", "q8")}      # C oracle partials, merged on rank 0
    q7_parts = []                                    # numpy oracle rows (timeFormat dimension)
    big = {"q3": [0, 0.0], "q10": [0, 0.0]}          # groups and sum(price) of the large results
    q10_keys = set()
    for piece in range(pieces):
        tab = tpch.FlatTable(sf, COLS, rank * pieces + piece, n_shards)
        eng.load("tpch", tab.engine_columns(), partition=rank * pieces + piece)
        n_local += tab.n_rows
        cs = bench.CpuShard(tab)
        for n in small:
            small[n].append(c_oracle.execute_partial(cs.ct, SU[n], threads))
        q7_parts.append(orc.execute(cs.table, SU["q7"]))
        for n in ("q3", "q10"):
            # order keys / customer keys do not repeat across generator shards for q3 (orders are dealt out in blocks); for
            # q10 a customer's orders do, so its groups are counted on the merged key set below
            rows = c_oracle.execute(cs.ct, dict(SU[n], limitSpec=None), threads)
            big[n][1] += math.fsum(r["price"] if n == "q10" else r["revenue"] for r in rows)
            big[n][0] = (big[n][0] + len(rows)) if n == "q3" else big[n][0]
            if n == "q10":
                q10_keys.update((r["o_custkey"], r["c_nation"]) for r in rows)
        del tab, cs
    t_load = time.perf_counter() - t_load
    log("loaded", n_local, "rows in", round(t_load, 1), "s")
    sharded.init_comm(eng)
    prepared = {n: sharded.prepare_query(eng, q) for n, q in SU.items()}

    def run(name, turn=None):
        h = C.c_void_p()
        qb = prepared[name]
        if turn is None:
            sdo._check(L.sdo_query_execute_sharded(eng._h, qb, len(qb), 0, C.byref(h)))
        else:
            sdo._check(L.sdo_query_execute_sharded_at(eng._h, qb, len(qb), 0, turn, C.byref(h)))
        return h if h else None

    # ---- parity gate ---------------------------------------------------------------------------------------------------
    parity, rows_scanned = {}, {}
    for n in SU:
        log("gate", n)
        h = run(n)
        res = sdo.Result(h) if h is not None else None
        tot = torch.tensor([eng.explain(json.loads(prepared[n]))["rowsScanned"]], dtype=torch.int64, device=dev)
        dist.all_reduce(tot)
        rows_scanned[n] = int(tot[0])
        if n in small or n == "q7":
            box = [None] * world
            dist.all_gather_object(box, small[n] if n in small else q7_parts)
            if rank == 0:
                parts = [p for b in box for p in b]
                if n == "q7":
                    want = orc.merge_partials(parts, ["s_nation", "c_nation", "l_year"], {"alias-1": "sum"})
                else:
                    want = c_oracle.merge_partials(SU[n], parts)
                parity[n] = bench.rows_equal(H.engine_rows(res), want, ordered=False) or "exact"
        else:
            # every rank returns its share of the groups: count them and sum the aggregate
            col = "price" if n == "q10" else "revenue"
            mine = torch.tensor([float(res.n_rows), float(np.asarray(res[col]).sum()) if res.n_rows else 0.0], dtype=torch.float64, device=dev)
            want = torch.tensor([float(big[n][0]), big[n][1]], dtype=torch.float64, device=dev)
            dist.all_reduce(mine)
            dist.all_reduce(want)
            if n == "q10":
                box = [None] * world
                dist.all_gather_object(box, q10_keys)
                want[0] = float(len(set().union(*box)))
            if rank == 0:
                ok = int(mine[0]) == int(want[0]) and abs(float(mine[1]) - float(want[1])) <= 1e-9 * abs(float(want[1]))
                parity[n] = "exact (groups %d, sum within 1e-9)" % int(mine[0]) if ok else f"FAILED: groups {int(mine[0])} vs {int(want[0])}, sum {float(mine[1])!r} vs {float(want[1])!r}"
        del res
    failed = torch.tensor([0 if rank != 0 or all(str(v).startswith("exact") for v in parity.values()) else 1], device=dev)
    dist.broadcast(failed, 0)
    if int(failed[0]):
        if rank == 0:
            print(json.dumps({"suite": "failed", "parity": parity}))
        dist.destroy_process_group()
        sys.exit(1)

    # ---- compile every shape, then the concurrent runs ------------------------------------------------------------------------
    while L.sdo_jit_pending() > 0:
        time.sleep(0.05)
    for n in SU:
        h = run(n)
        if h is not None:
            L.sdo_result_free(h)
    names = list(SU)
    out = {}
    for T in callers:
        log("callers", T)
        rounds = 6
        base = eng.sharded_next_turn()
        lat = {n: [] for n in names}
        lock = threading.Lock()
        gate = threading.Barrier(T + 1)
        errors = []

        def caller(t):
            gate.wait()
            try:
                for r in range(rounds):
                    for k in range(len(names)):
                        n = names[(k + t) % len(names)]
                        t0 = time.perf_counter()
                        h = run(n, base + (r * len(names) + k) * T + t)
                        dt = (time.perf_counter() - t0) * 1e3
                        if h is not None:
                            L.sdo_result_free(h)
                        with lock:
                            lat[n].append(dt)
            except Exception as ex:  # noqa: BLE001
                log("caller", t, "failed:", repr(ex))
                errors.append(repr(ex))

        # one untimed pass with this many callers first: contexts, pinned buffers and the memory pool grow to what T
        # queries in flight need
        warm_base = base
        base += len(names) * T

        def warm(t):
            for k in range(len(names)):
                h = run(names[(k + t) % len(names)], warm_base + k * T + t)
                if h is not None:
                    L.sdo_result_free(h)

        ws = [threading.Thread(target=warm, args=(t,)) for t in range(T)]
        for th in ws:
            th.start()
        for th in ws:
            th.join()
        ths = [threading.Thread(target=caller, args=(t,)) for t in range(T)]
        for th in ths:
            th.start()
        torch.cuda.synchronize()
        dist.barrier()
        gate.wait()
        t0 = time.perf_counter()
        for th in ths:
            th.join()
        wall = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        dist.all_reduce(wall, op=dist.ReduceOp.MAX)
        if errors:
            raise SystemExit(f"rank {rank}: {errors[:2]}")
        total_rows = sum(rows_scanned[n] * len(v) for n, v in lat.items())
        out[f"{T}_callers_per_rank"] = {"queries": sum(len(v) for v in lat.values()), "wall_s": float(wall[0]), "rows_scanned_per_s": total_rows / float(wall[0]),
                                        "latency_ms_mean_rank0": {n: sum(v) / len(v) for n, v in lat.items()},
                                        "latency_ms_max_rank0": {n: max(v) for n, v in lat.items()}}
    if rank == 0:
        print(json.dumps({"suite": "TPC-H flattened Q1, Q3, Q5, Q7, Q8, Q10 (TpchBenchMark.scala:208-288), concurrent callers per rank, sharded execution with turn-ordered merges",
                          "sf_total": sf, "n_gpus": world, "rows_per_gpu": n_local, "shards": n_shards, "load_s": round(t_load, 1), "parity": parity,
                          "rows_scanned_per_query": rows_scanned, "runs": out}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
