"""Where does the sharded path spend host time?  One GPU, world size 1, Q5T at SF10."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import queries as Q  # noqa: E402
import spark_druid_olap_b200 as sdo  # noqa: E402
from spark_druid_olap_b200 import sharded, tpch  # noqa: E402

torch.cuda.set_device(0)
dist.init_process_group("nccl", init_method="file:///tmp/sdo_pg_prof", rank=0, world_size=1, device_id=torch.device("cuda", 0))
eng = sdo.Engine(0)
tpch.load_into(eng, "tpch", float(sys.argv[1]) if len(sys.argv) > 1 else 10.0, ["s_region", "o_orderdate", "s_nation", "l_extendedprice"])
q = Q.q5_topn().to_json().encode()


def timeit(f, n=200):
    for _ in range(20):
        f()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


print("execute            %.4f ms" % timeit(lambda: eng.execute(q)))
print("execute_raw+free   %.4f ms" % timeit(lambda: sdo.lib().sdo_result_free(eng.execute_raw(q))))
print("sharded world1     %.4f ms" % timeit(lambda: sharded.execute_sharded(eng, q)))
print("sharded + exchange %.4f ms" % timeit(lambda: sharded.execute_sharded(eng, q, force_exchange=True)))
pr = cProfile.Profile()
pr.enable()
for _ in range(300):
    sharded.execute_sharded(eng, q, force_exchange=True)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
dist.destroy_process_group()
