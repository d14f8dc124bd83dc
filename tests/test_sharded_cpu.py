"""Host-side merge logic (tests/host_merge.py) on 2 ranks over gloo (CPU): the merge algebra the library applies over NCCL,
stated with tensors.  Mirrors the reference's broker-vs-historical equivalence test (HistoricalServerTest.scala:177-230):
N per-shard partial queries + merge must equal the one merged query."""
import json
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as H
import queries as Q
from oracle import druid_oracle as orc

SF = 0.01
WORLD = 2


def _partial_table(tbl, q, keys, ops_cols):
    """Dense partial table of one shard, laid out like the engine's: [n_keys, n_slots] int64 words."""
    rows = orc.execute(tbl, q)
    out = {}
    for r in rows:
        out[tuple(r[k] for k in keys)] = r
    return out


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import host_merge as sharded
    from spark_druid_olap_b200 import tpch
    cols = ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "ps_availqty", "o_custkey", "o_orderkey", "count"]
    mine = H.oracle_table(tpch.FlatTable(SF, cols, rank, world))
    whole = H.oracle_table([tpch.FlatTable(SF, cols, r, world) for r in range(world)])

    # ---- dense: fixed key space (3 flags x 2 statuses), slots = count, long sum, double sum, long min, double max(ord) ----
    q = {"queryType": "groupBy", "dataSource": "t", "dimensions": ["l_returnflag", "l_linestatus"], "granularity": "all",
         "aggregations": [{"type": "count", "name": "n"}, {"type": "longSum", "name": "ls", "fieldName": "ps_availqty"},
                          {"type": "doubleSum", "name": "ds", "fieldName": "l_extendedprice"}, {"type": "longMin", "name": "lmin", "fieldName": "ps_availqty"},
                          {"type": "longMax", "name": "lmax", "fieldName": "ps_availqty"}, {"type": "cardinality", "name": "dc", "fieldNames": ["o_custkey"], "byRow": True}]}
    flags, stats = ["A", "N", "R"], ["F", "O"]
    slot_ops = [sharded.SLOT_SUM_I64, sharded.SLOT_SUM_I64, sharded.SLOT_SUM_F64, sharded.SLOT_MIN_I64, sharded.SLOT_MAX_I64]
    table = torch.zeros((6, 5), dtype=torch.int64)
    table[:, 3] = np.iinfo(np.int64).max
    table[:, 4] = np.iinfo(np.int64).min
    regs = torch.zeros((6, 512), dtype=torch.uint8)
    part, pregs = orc.execute(mine, q, return_registers=True)
    for i, r in enumerate(part):
        k = flags.index(r["l_returnflag"]) * 2 + stats.index(r["l_linestatus"])
        table[k, 0], table[k, 1], table[k, 3], table[k, 4] = r["n"], r["ls"], r["lmin"], r["lmax"]
        table[k, 2] = int(np.float64(r["ds"]).view(np.int64))
        regs[k] = torch.from_numpy(pregs["dc"][i])
    sharded.merge_dense_table(table, slot_ops)
    sharded.merge_hll(regs)
    want, wregs = orc.execute(whole, q, return_registers=True)
    for i, r in enumerate(want):
        k = flags.index(r["l_returnflag"]) * 2 + stats.index(r["l_linestatus"])
        assert (int(table[k, 0]), int(table[k, 1]), int(table[k, 3]), int(table[k, 4])) == (r["n"], r["ls"], r["lmin"], r["lmax"])
        ds = float(np.int64(int(table[k, 2])).view(np.float64))
        assert abs(ds - r["ds"]) <= H.F64_RTOL * abs(r["ds"])
        assert np.array_equal(regs[k].numpy(), wregs["dc"][i])  # HLL merge = register-wise max, bit-exact

    # ---- sparse: group by o_orderkey, rows routed to an owner rank, merged there ----
    q2 = {"queryType": "groupBy", "dataSource": "t", "dimensions": ["o_orderkey"], "granularity": "all",
          "aggregations": [{"type": "count", "name": "n"}, {"type": "doubleSum", "name": "p", "fieldName": "l_extendedprice"}]}
    part = orc.execute(mine, q2)
    keys = np.array([int(r["o_orderkey"]) for r in part], dtype=np.int64)
    slots = np.stack([np.array([r["n"] for r in part], dtype=np.int64), np.array([r["p"] for r in part]).view(np.int64)], axis=1)
    owner = (keys * 2654435761 >> 7) % world
    order = np.argsort(owner, kind="stable")
    counts = [int((owner == r).sum()) for r in range(world)]
    rk, rs = sharded.exchange_sparse(torch.from_numpy(keys[order]), torch.from_numpy(slots[order]), counts)
    acc = {}
    for k, (n, pbits) in zip(rk.tolist(), rs.tolist()):
        a = acc.setdefault(k, [0, 0.0])
        a[0] += n
        a[1] += float(np.int64(pbits).view(np.float64))
    want2 = {int(r["o_orderkey"]): r for r in orc.execute(whole, q2)}
    mine_keys = {k for k in want2 if (k * 2654435761 >> 7) % world == rank}
    assert set(acc) == mine_keys
    for k, (n, p) in acc.items():
        assert n == want2[k]["n"] and abs(p - want2[k]["p"]) <= H.F64_RTOL * abs(want2[k]["p"])
    sizes = [None] * world
    dist.all_gather_object(sizes, len(acc))
    assert sum(sizes) == len(want2)
    open(os.path.join(tmp, f"ok{rank}"), "w").write("ok")
    dist.destroy_process_group()


def test_dense_and_sparse_merge_two_ranks(native_libs, tmp_path):
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(WORLD, port, str(tmp_path)), nprocs=WORLD, join=True)
    assert all(os.path.exists(tmp_path / f"ok{r}") for r in range(WORLD))
