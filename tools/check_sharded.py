"""torchrun --nproc-per-node N tools/check_sharded.py [sf]: every rank loads shard `rank` of N, runs the corpus through the
sharded path (NCCL merge) and rank 0 (dense) / all ranks (sparse) compare with the oracle on the whole table."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import faulthandler  # noqa: E402
import torch  # noqa: E402

faulthandler.dump_traceback_later(100, exit=True)
import torch.distributed as dist  # noqa: E402
import helpers as H  # noqa: E402
import queries as Q  # noqa: E402
import spark_druid_olap_b200 as sdo  # noqa: E402
from oracle import druid_oracle as orc  # noqa: E402
from spark_druid_olap_b200 import sharded, tpch  # noqa: E402

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 0.05
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cols = ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "o_orderkey", "o_custkey", "o_orderdate", "o_shippriority",
        "c_mktsegment", "s_region", "s_nation", "count", "ps_availqty", "ps_supplycost", "l_shipmode", "c_region"]
eng = sdo.Engine(local)
eng.create_datasource("tpch")
if os.environ.get("CHECK_SKEW"):
    # rank 0 holds 1/16 of the rows, the others share the rest: the merged table on rank 0 has far more groups than
    # rank 0 scanned rows (what a root sees at 8+ ranks)
    pieces = 16
    owners = [0] + [1 + (i % max(world - 1, 1)) if world > 1 else 0 for i in range(pieces - 1)]
    mine = None
    for i in range(pieces):
        if owners[i] == rank:
            t = tpch.FlatTable(sf, cols, i, pieces)
            eng.load("tpch", t.engine_columns(), partition=i)
            mine = mine or t
    whole = H.oracle_table([tpch.FlatTable(sf, cols, i, pieces) for i in range(pieces)])
else:
    mine = tpch.FlatTable(sf, cols, rank, world)
    eng.load("tpch", mine.engine_columns(), partition=rank)
    whole = H.oracle_table([tpch.FlatTable(sf, cols, r, world) for r in range(world)])
ok = True
import host_merge  # noqa: E402
# two passes: the host-side merge over the partial API (tests/host_merge.py), then the merge inside the library
for pass_name in ("host-merge", "library"):
  if pass_name == "library":
    sharded.init_comm(eng)
  for name, spec, ordered in [("q1", Q.q1_style(), False), ("q1_block", Q.q1_style("block"), False), ("ts_month_hll", Q.timeseries("P1M"), False),
                              ("ts_day_hll", Q.timeseries("P1D"), False), ("q5t", Q.q5_topn(), True), ("by_orderkey_hash", Q.by_orderkey("hash"), False),
                              ("by_orderkey_global", Q.by_orderkey("global"), False), ("by_orderkey", Q.by_orderkey(), False), ("q3", Q.q3(), False),
                              ("q3_limit10", Q.q3(10), True), ("q1_reduce", Q.q1_style(), False), ("ts_day_reduce", Q.timeseries("P1D"), False),
                              ("q5_sparse", Q.q5(), False), ("by_orderkey_emit", Q.by_orderkey("emit"), False), ("q3_emit", Q.q3(None, "emit"), False),
                              ("q3_limit10_emit", Q.q3(10, "emit"), True), ("q5t_emit", Q.q5_topn(10, "emit"), True)] * int(os.environ.get("CHECK_REPEAT", "1")):
      if os.environ.get("CHECK_ONLY") and name not in os.environ["CHECK_ONLY"].split(","):
          continue
      q = json.loads(spec.to_json())
      if name.endswith("_emit") and pass_name != "library":
          continue
      if name.endswith("_reduce") or name.endswith("_sparse"):
          if pass_name != "library":
              continue
          q["context"] = dict(q.get("context") or {"queryId": name}, gpuShardMerge=name.rsplit("_", 1)[1])
      if pass_name == "host-merge" and name == "q3_limit10":
          continue  # the host-side merge returns per-rank shares; a limit applies per share there
      res, info = (sharded.execute_sharded(eng, q) if pass_name == "library" else host_merge.execute_host_merge(eng, q))
      want = orc.execute(whole, q)
      # dense merges and top-K queries end on rank 0, sparse merges leave every rank a disjoint share: compare the union
      parts = [None] * world
      dist.all_gather_object(parts, H.engine_rows(res) if res is not None else [])
      try:
          if rank == 0:
              holders = sum(1 for p in parts if p)
              allrows = [r for p in parts for r in p]
              H.assert_rows_equal(allrows, want, ordered=ordered and holders <= 1, what=name)
              print(f"OK   {name:22s} [{pass_name}] merge={info.get('merge', 'dense' if info.get('dense') else 'sparse')} rows={len(allrows)} from {holders} rank(s) "
                    f"bytes={info.get('merge_bytes')}", flush=True)
      except AssertionError as e:
          ok = False
          print(f"FAIL {name} [{pass_name}] rank {rank}: {e}", flush=True)
# callers on threads of their own (a JVM executor's task threads): a thread that has never touched CUDA has device 0
# current, whatever the engine's device is -- every rank but 0 exercises that here; turn numbers order the merges
import threading  # noqa: E402
thread_out, thread_err = {}, []
base = eng.sharded_next_turn()
shapes_t = [("q1", Q.q1_style(), False), ("q5t", Q.q5_topn(), True), ("q3_limit10_emit", Q.q3(10, "emit"), True), ("ts_month_hll", Q.timeseries("P1M"), False)]


def _call(i):
    try:
        name, spec, ordered = shapes_t[i % len(shapes_t)]
        res, _ = sharded.execute_sharded(eng, json.loads(spec.to_json()), turn=base + i)
        thread_out[i] = H.engine_rows(res) if res is not None else []
        if i == 0:
            thread_out["local"] = H.engine_rows(eng.execute(json.loads(Q.q1_style().to_json())))  # a plain query from a fresh thread
    except Exception as ex:  # noqa: BLE001
        thread_err.append(f"thread {i}: {ex!r}")


ths = [threading.Thread(target=_call, args=(i,)) for i in reversed(range(8))]
for th in ths:
    th.start()
for th in ths:
    th.join(90)
if thread_err or any(th.is_alive() for th in ths):
    ok = False
    print(f"FAIL threads rank {rank}: {thread_err[:2]} alive={sum(th.is_alive() for th in ths)}", flush=True)
elif rank == 0:
    try:
        for i in range(8):
            name, spec, ordered = shapes_t[i % len(shapes_t)]
            H.assert_rows_equal(thread_out[i], orc.execute(whole, json.loads(spec.to_json())), ordered=ordered, what=f"thread {i} {name}")
        print("OK   8 caller threads, turn-ordered merges", flush=True)
    except AssertionError as e:
        ok = False
        print(f"FAIL threads: {e}", flush=True)
flag = torch.tensor([0 if ok else 1], device="cuda")
dist.all_reduce(flag)
if rank == 0:
    print("sharded check:", "PASS" if int(flag) == 0 else "FAIL", f"world={world} sf={sf}")
dist.destroy_process_group()
sys.exit(0 if int(flag) == 0 else 1)
