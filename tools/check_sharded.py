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
        "c_mktsegment", "s_region", "s_nation", "count"]
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
# second pass: the same dense shapes with the merge inside the library (sdo_query_execute_sharded)
passes = [("host-merge", False), ("native-merge", True)]
for pass_name, native in passes:
  if native:
    sharded.init_comm(eng)
  for name, spec, ordered in [("q1", Q.q1_style(), False), ("q1_block", Q.q1_style("block"), False), ("ts_month_hll", Q.timeseries("P1M"), False),
                              ("ts_day_hll", Q.timeseries("P1D"), False), ("q5t", Q.q5_topn(), True), ("by_orderkey_hash", Q.by_orderkey("hash"), False),
                              ("by_orderkey_global", Q.by_orderkey("global"), False), ("q3", Q.q3(), False)] * int(os.environ.get("CHECK_REPEAT", "1")):
      if os.environ.get("CHECK_ONLY") and name not in os.environ["CHECK_ONLY"].split(","):
          continue
      q = json.loads(spec.to_json())
      print(f'[rank {rank}] {name} ...', flush=True)
      res, info = sharded.execute_sharded(eng, q)
      print(f'[rank {rank}] {name} engine done', flush=True)
      want = orc.execute(whole, q)
      try:
          if info["dense"]:
              if rank == 0:
                  H.assert_rows_equal(H.engine_rows(res), want, ordered=ordered, what=name)
          else:
              # sparse: every rank holds a disjoint share; gather and compare the union
              parts = [None] * world
              rows_here = H.engine_rows(res)
              cnt_col = [k for k in rows_here[0] if k in ('n',)] if rows_here else []
              print(f'[rank {rank}] {name}: {len(rows_here)} result rows, sum n = {sum(r[cnt_col[0]] for r in rows_here) if cnt_col else None}, local rows {mine.n_rows}, info {info}', flush=True)
              dist.all_gather_object(parts, rows_here)
              if rank == 0:
                  allrows = [r for p in parts for r in p]
                  keys = [r.get('o_orderkey') for r in allrows]
                  print(f'union {len(allrows)} rows, distinct keys {len(set(keys))}, want {len(want)}', flush=True)
                  H.assert_rows_equal(allrows, want, what=name)
          if rank == 0:
              print(f"OK   {name:22s} [{pass_name}{' in library' if info.get('native') else ''}] dense={info['dense']} keys={info['n_keys']} merge_bytes={info['merge_bytes']}")
      except AssertionError as e:
          ok = False
          print(f"FAIL {name} [{pass_name}] rank {rank}: {e}")
          if rank == 0 and info["dense"] and res is not None:
              # what differs: groups only on one side, and groups whose exact columns differ
              got_rows = H.engine_rows(res)
              dim = [k for k in want[0] if isinstance(want[0][k], str)]
              gk = {tuple(r[d] for d in dim): r for r in got_rows}
              wk = {tuple(r[d] for d in dim): r for r in want}
              only_g, only_w = sorted(set(gk) - set(wk))[:5], sorted(set(wk) - set(gk))[:5]
              diff = [(k, gk[k], wk[k]) for k in sorted(set(gk) & set(wk)) if any(gk[k][c] != wk[k][c] for c in wk[k] if not isinstance(wk[k][c], float))][:5]
              print(f"  rows got {len(got_rows)} want {len(want)}; only in result {only_g}; only in oracle {only_w}; exact-column mismatches {diff}", flush=True)
flag = torch.tensor([0 if ok else 1], device="cuda")
dist.all_reduce(flag)
if rank == 0:
    print("sharded check:", "PASS" if int(flag) == 0 else "FAIL", f"world={world} sf={sf}")
dist.destroy_process_group()
sys.exit(0 if int(flag) == 0 else 1)
