"""torchrun --nproc-per-node N tools/prof_sparse.py [sf_per_gpu]: phases of the sparse (hash-partitioned all-to-all) merge
for group-by o_orderkey + top 10, each timed with a device synchronize on both sides (max over ranks)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import faulthandler  # noqa: E402
import torch  # noqa: E402

faulthandler.dump_traceback_later(200, exit=True)
import torch.distributed as dist  # noqa: E402
import queries as Q  # noqa: E402
import spark_druid_olap_b200 as sdo  # noqa: E402
from spark_druid_olap_b200 import sharded, tpch  # noqa: E402

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = sdo.Engine(local)
eng.create_datasource("tpch")
for piece in range(2):
    tab = tpch.FlatTable(sf * world, ["o_orderkey", "l_extendedprice", "count"], rank * 2 + piece, world * 2)
    eng.load("tpch", tab.engine_columns(), partition=rank * 2 + piece)
    del tab
spec = Q.GroupByQuerySpec(Q.DS, [Q.DefaultDimensionSpec("o_orderkey")], Q.LimitSpec(10, [Q.OrderByColumnSpec("price", "descending")]), None, "all", None,
                          [Q.agg("count", "n", "count"), Q.agg("doubleSum", "price", "l_extendedprice")], None, Q.ALL_TIME, Q.ctx("hash"))
query = sharded.prepare_query(eng, spec)
dev = torch.device("cuda", local)


def run(timed):
    marks = []

    def mark(name):
        if timed:
            torch.cuda.synchronize()
            marks.append((name, time.perf_counter()))

    mark("start")
    part = eng.partial(query)
    try:
        stream = torch.cuda.ExternalStream(part.stream(), device=dev)
        with torch.cuda.stream(stream):
            mark("scan + compact")
            counts = part.partition(world)
            mark("partition")
            v = part.view()
            n = int(v.n_keys)
            keys = sharded.device_bytes(v.keys, max(n, 1) * 8, dev).view(torch.int64)[:n]
            slots = sharded.device_bytes(v.slots, max(n, 1) * v.n_slots * 8, dev).view(torch.int64)[:n * v.n_slots].view(n, v.n_slots)
            rkeys, rslots = sharded.exchange_sparse(keys, slots, counts)
            mark("all-to-all")
            stream.synchronize()
            part.merge_rows(rkeys.data_ptr(), rslots.data_ptr(), int(rkeys.shape[0]))
            mark("merge rows")
            r = part.finish()
            mark("finish (top 10)")
    finally:
        part.free()
    mark("free")
    return marks, n, int(rkeys.shape[0]), r


for _ in range(3):
    run(False)
dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    run(False)
torch.cuda.synchronize()
dist.barrier()
whole = (time.perf_counter() - t0) / 5
marks, n_local, n_recv, r = run(True)
phases = [(b[0], 1000 * (b[1] - a[1])) for a, b in zip(marks, marks[1:])]
t = torch.tensor([whole * 1000] + [p[1] for p in phases], dtype=torch.float64, device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"world={world} sf/gpu={sf:g} local groups={n_local} received={n_recv} result rows={r.n_rows}")
    print(f"untimed loop: {float(t[0]):.3f} ms per query")
    for (name, _), v in zip(phases, t[1:].tolist()):
        print(f"  {name:18s} {v:8.3f} ms")
dist.destroy_process_group()
