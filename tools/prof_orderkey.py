"""torchrun --nproc-per-node N tools/prof_orderkey.py [sf]: the high-cardinality sharded group-by (BASELINE config 4: group by
o_orderkey, top 10 by revenue) over an SF table split N ways; prints rank 0's per-phase device times."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402
import spark_druid_olap_b200 as sdo  # noqa: E402
from spark_druid_olap_b200 import sharded, tpch  # noqa: E402

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 100.0
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = sdo.Engine(local)
eng.create_datasource("tpch")
pieces = max(1, int(sf / world // 10))
for p in range(pieces):
    t = tpch.FlatTable(sf, ["o_orderkey", "l_extendedprice"], rank * pieces + p, world * pieces)
    eng.load("tpch", t.engine_columns(), partition=rank * pieces + p)
    del t
sharded.init_comm(eng)
q = bench.shapes()["by_orderkey_top10"][0]
for variant in sys.argv[2:] or ["default"]:
    qq = dict(q)
    if variant != "default":
        qq["context"] = {"queryId": "p", "gpuStrategy": variant}
    for it in range(6):
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        res, info = sharded.execute_sharded(eng, qq)
        dt = (time.perf_counter() - t0) * 1e3
        if rank == 0 and it >= 2:
            st = res.stats
            print(f"{variant:8s} wall {dt:7.3f} ms  device {st['deviceMs']:7.3f}  scan {st['scanKernelMs']:6.3f}  merge={st.get('shardMerge')} phases={st.get('shardPhasesMs')} "
                  f"exchange={st.get('exchangeBytes', 0) / 1e6:.0f} MB launches={st['launches']} rows={res.rows()[:2]}", flush=True)
dist.destroy_process_group()
