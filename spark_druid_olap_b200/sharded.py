"""Sharded execution: one engine (process) per GPU, partial aggregates merged over NCCL.

Replaces what the reference does when it queries historicals directly: one partial query per
(server, segment group) partition (DruidRDD.scala:244-269) and a Spark-side re-aggregation of the
partials (sources/druid/DruidStrategy.scala:353-359) with the merge algebra of
sources/druid/PostAggregate.scala:39-60 (count/sum -> sum, min -> min, max -> max; HLL -> register max).

Dense key spaces: every rank holds the same fixed-layout table, so the merge is an all-reduce per
slot class (int64 sum, float64 sum, int64 min, int64 max) plus a uint8 max for HLL registers.
Sparse key spaces: rows are split by owner = mix(key) % world, exchanged with one all-to-all, and
merged into the owner's hash table; every rank then finishes its disjoint share.

The tensor-level functions work on any device/backend (the CPU tests drive them over gloo);
execute_sharded binds them to an engine's device buffers without copies.
"""
import torch
import torch.distributed as dist

SLOT_SUM_I64, SLOT_SUM_F64, SLOT_MIN_I64, SLOT_MAX_I64 = 0, 1, 2, 3


class _DevArray:
    """Zero-copy view of engine-owned device memory for torch (CUDA array interface)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def device_bytes(ptr, nbytes, device):
    return torch.as_tensor(_DevArray(ptr, nbytes), device=device)


def merge_dense_table(table, slot_ops, group=None):
    """All-reduce a [n_keys, n_slots] int64-typed table in place, one collective per slot class."""
    n_slots = table.shape[1]
    assert len(slot_ops) == n_slots
    for op, red, as_float in ((SLOT_SUM_I64, dist.ReduceOp.SUM, False), (SLOT_SUM_F64, dist.ReduceOp.SUM, True),
                              (SLOT_MIN_I64, dist.ReduceOp.MIN, False), (SLOT_MAX_I64, dist.ReduceOp.MAX, False)):
        cols = [i for i, o in enumerate(slot_ops) if o == op]
        if not cols:
            continue
        buf = table[:, cols].contiguous()
        if as_float:
            fb = buf.view(torch.float64)
            dist.all_reduce(fb, op=red, group=group)
        else:
            dist.all_reduce(buf, op=red, group=group)
        table[:, cols] = buf
    return table


def merge_hll(registers, group=None):
    """Register-wise max of HLL++ sketches (uint8), in place."""
    dist.all_reduce(registers, op=dist.ReduceOp.MAX, group=group)
    return registers


def exchange_sparse(keys, slots, counts, group=None):
    """All-to-all of (key, slots) rows already grouped by destination rank.
    keys [n] int64, slots [n, n_slots] int64, counts: rows per destination.  Returns (keys, slots) received."""
    world = dist.get_world_size(group)
    dev = keys.device
    send = torch.tensor(counts, dtype=torch.int64, device=dev)
    recv = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_to_all_single(recv, send, group=group)
    rc = [int(x) for x in recv.tolist()]
    n_slots = slots.shape[1]
    rkeys = torch.empty(sum(rc), dtype=torch.int64, device=dev)
    dist.all_to_all_single(rkeys, keys, output_split_sizes=rc, input_split_sizes=list(counts), group=group)
    rslots = torch.empty((sum(rc), n_slots), dtype=torch.int64, device=dev)
    dist.all_to_all_single(rslots, slots.contiguous(), output_split_sizes=rc, input_split_sizes=list(counts), group=group)
    return rkeys, rslots


def execute_sharded(engine, query, group=None, root=0):
    """Run `query` over every rank's local segments and return the merged Result on `root`
    (dense key spaces) or this rank's share of the groups (sparse key spaces; see `sparse` flag).
    Returns (result_or_None, info)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    part = engine.partial(query)
    try:
        v = part.view()
        dev = torch.device("cuda", engine.device)
        stream = torch.cuda.ExternalStream(part.stream(), device=dev)
        info = {"dense": bool(v.dense), "n_keys": int(v.n_keys), "n_slots": int(v.n_slots), "merge_bytes": 0}
        with torch.cuda.stream(stream):
            if v.dense:
                ops = [v.slot_ops[i] for i in range(v.n_slots)]
                if world > 1 and v.n_keys > 0:
                    nb = v.n_keys * v.n_slots * 8
                    table = device_bytes(v.slots, nb, dev).view(torch.int64).view(v.n_keys, v.n_slots)
                    merge_dense_table(table, ops, group)
                    info["merge_bytes"] += nb
                    if v.n_hll and v.hll:
                        hb = v.n_keys * v.n_hll * 512
                        merge_hll(device_bytes(v.hll, hb, dev), group)
                        info["merge_bytes"] += hb
                stream.synchronize()
                if rank == root:
                    return part.finish(), info
                return None, info
            # sparse: hash-partitioned all-to-all, every rank keeps the groups it owns
            if world > 1:
                counts = part.partition(world)
                v = part.view()
                n = int(v.n_keys)
                keys = device_bytes(v.keys, max(n, 1) * 8, dev).view(torch.int64)[:n]
                slots = device_bytes(v.slots, max(n, 1) * v.n_slots * 8, dev).view(torch.int64)[:n * v.n_slots].view(n, v.n_slots)
                rkeys, rslots = exchange_sparse(keys, slots, counts, group)
                info["merge_bytes"] += (n - counts[rank]) * (1 + v.n_slots) * 8
                stream.synchronize()
                part.merge_rows(rkeys.data_ptr(), rslots.data_ptr(), int(rkeys.shape[0]))
            return part.finish(), info
    finally:
        part.free()
