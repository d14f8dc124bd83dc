"""Host-side merge of partial aggregates (TEST INFRASTRUCTURE): what a host layer without the in-library merge would do with
the sdo_query_partial* entry points -- export / view the partial state on the device, merge it with the host's own
collectives (here torch.distributed: NCCL on GPUs, gloo in the CPU tests), hand the merged state back, finish.
The product path is spark_druid_olap_b200.sharded.execute_sharded (sdo_query_execute_sharded: everything inside the
library); this module keeps the partial API and the merge algebra (PostAggregate.scala:39-60) under test."""
import torch
import torch.distributed as dist

import spark_druid_olap_b200 as sdo
from spark_druid_olap_b200.sharded import prepare_query

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


_buffers = {}
GATHER_MAX_BYTES = 4 << 20  # dense partials up to this size are merged by all-gather + one fold kernel


def _gather_buffers(dev, nbytes, world):
    """Reusable (local, gathered) byte tensors for the small-table merge."""
    key = (dev.index, nbytes, world)
    if key not in _buffers:
        _buffers[key] = (torch.empty(nbytes, dtype=torch.uint8, device=dev), torch.empty(nbytes * world, dtype=torch.uint8, device=dev))
    return _buffers[key]


def execute_host_merge(engine, query, group=None, root=0, force_exchange=False):
    """Run `query` over every rank's local segments and return the merged Result on `root`
    (dense key spaces) or this rank's share of the groups (sparse key spaces; see `sparse` flag).
    Returns (result_or_None, info)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world > 1:
        query = prepare_query(engine, query, group)
    part = engine.partial(query)
    try:
        v = part.view()
        dev = torch.device("cuda", engine.device)
        stream = torch.cuda.ExternalStream(part.stream(), device=dev)
        info = {"dense": bool(v.dense), "n_keys": int(v.n_keys), "n_slots": int(v.n_slots), "merge_bytes": 0}
        with torch.cuda.stream(stream):
            if v.dense and (world > 1 or force_exchange) and 0 < v.blob_bytes <= GATHER_MAX_BYTES:
                # small table: one all-gather of the raw blobs (any mix of sum / min / max slots and HLL registers),
                # folded in rank order by one kernel on the root
                nb = int(v.blob_bytes)
                local, gathered = _gather_buffers(dev, nb, world)
                part.export(local.data_ptr(), nb)
                dist.all_gather_into_tensor(gathered, local, group=group)
                info["merge_bytes"] += nb * (world - 1)
                if rank == root:
                    part.merge_gathered(gathered.data_ptr(), world)
                    return part.finish(), info
                stream.synchronize()
                return None, info
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
            # (force_exchange runs the partition / exchange / merge steps even on one rank: tests)
            if world > 1 or force_exchange:
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
