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
import json
import os

import torch
import torch.distributed as dist

import spark_druid_olap_b200 as sdo

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


_boundary_cache = {}
_aligned = set()
_buffers = {}
GATHER_MAX_BYTES = 4 << 20  # dense partials up to this size are merged by all-gather + one fold kernel


def _gather_buffers(dev, nbytes, world):
    """Reusable (local, gathered) byte tensors for the small-table merge."""
    key = (dev.index, nbytes, world)
    if key not in _buffers:
        _buffers[key] = (torch.empty(nbytes, dtype=torch.uint8, device=dev), torch.empty(nbytes * world, dtype=torch.uint8, device=dev))
    return _buffers[key]


def align_datasource(engine, datasource, group=None, force=False):
    """Give every rank's shard of `datasource` the same id space: string dimensions get the union of all
    ranks' dictionaries (device re-encode where ids move), integer dimensions the largest cardinality.
    Packed group keys, dense table shapes and HLL lookup tables then mean the same thing on every rank.
    Done once per engine and datasource; call again with force=True after loading more segments."""
    key = (id(engine), datasource)
    if key in _aligned and not force:
        return
    world = dist.get_world_size(group)
    md = engine.metadata(datasource)
    dims = sorted(n for n, c in md["columns"].items() if c["type"] == "STRING")
    names = [None] * world
    dist.all_gather_object(names, dims, group=group)
    if any(n != dims for n in names):
        raise ValueError(f"ranks disagree on the dimensions of {datasource}: {names}")
    for d in dims:
        values, card = engine.dimension_dictionary(datasource, d)
        gathered = [None] * world
        dist.all_gather_object(gathered, (values, card), group=group)
        union = set()
        for v, _ in gathered:
            union.update(v)
        extra = sorted(union.difference(values))
        engine.extend_dimension(datasource, d, extra, max(c for _, c in gathered) if not union else 0)
    _aligned.add(key)
    _boundary_cache.pop(key, None)


def global_time_boundary(engine, datasource, group=None):
    """[min_ms, max_ms] of the datasource over all ranks (cached per engine and datasource).
    Time-derived group keys are numbered on this grid so that every rank uses the same key space."""
    key = (id(engine), datasource)
    if key not in _boundary_cache:
        tb = engine.metadata(datasource).get("timeBoundary")
        lo, hi = (tb["minMs"], tb["maxMs"]) if tb else (2 ** 62, -2 ** 62)
        dev = torch.device("cuda", engine.device) if engine.device >= 0 else torch.device("cpu")
        t = torch.tensor([-lo, hi], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        _boundary_cache[key] = (-int(t[0]), int(t[1]))
    return _boundary_cache[key]


def with_time_boundary(query, boundary):
    """The query as a dict with context.gpuTimeBoundary set."""
    if hasattr(query, "to_json"):
        query = json.loads(query.to_json())
    elif isinstance(query, (bytes, str)):
        query = json.loads(query)
    else:
        query = dict(query)
    ctx = dict(query.get("context") or {})
    ctx["gpuTimeBoundary"] = [int(boundary[0]), int(boundary[1])]
    query["context"] = ctx
    return query


_comms = {}     # id(engine) -> (world, group id) of the communicator the engine owns
_prepared = {}  # (id(engine), query text) -> query JSON bytes with the global time boundary set


def _nccl_library():
    """Path of the libnccl.so.2 PyTorch ships (the engine reuses the copy already loaded into the process if any)."""
    import glob
    import site
    for sp in site.getsitepackages():
        hits = glob.glob(os.path.join(sp, "nvidia", "nccl", "lib", "libnccl.so.2"))
        if hits:
            return hits[0].encode()
    return None


def init_comm(engine, group=None):
    """Give `engine` its own NCCL communicator over the ranks of `group` (collective; once per engine), so that
    sdo_query_execute_sharded can scan, gather and fold inside the library.  The unique id travels over `group`."""
    import ctypes as C
    if id(engine) in _comms:
        return
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    ident = C.create_string_buffer(sdo.COMM_ID_BYTES)
    path = _nccl_library()
    if rank == 0:
        sdo._check(sdo.lib().sdo_comm_unique_id(ident, path))
    box = [bytes(ident.raw)]
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    sdo._check(sdo.lib().sdo_engine_comm_init(engine._h, box[0], rank, world, path))
    _comms[id(engine)] = (world, id(group))


def prepare_query(engine, query, group=None):
    """Query JSON bytes every rank can run: dictionaries aligned, time keys on the global grid (cached per query)."""
    text = query if isinstance(query, (bytes, str)) else (query.to_json() if hasattr(query, "to_json") else json.dumps(query, sort_keys=True))
    key = (id(engine), text)
    if key not in _prepared:
        q = with_time_boundary(query, (0, 0))
        align_datasource(engine, q["dataSource"], group)
        _prepared[key] = json.dumps(with_time_boundary(q, global_time_boundary(engine, q["dataSource"], group))).encode()
    return _prepared[key]


def execute_sharded(engine, query, group=None, root=0, force_exchange=False):
    """Run `query` over every rank's local segments and return the merged Result on `root`
    (dense key spaces) or this rank's share of the groups (sparse key spaces; see `sparse` flag).
    Returns (result_or_None, info)."""
    import ctypes as C
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world > 1:
        query = prepare_query(engine, query, group)
        if _comms.get(id(engine)) == (world, id(group)) and not force_exchange:
            # merge inside the library: scan + NCCL all-gather + fold queued by one C call
            out = C.c_void_p()
            status = sdo.lib().sdo_query_execute_sharded(engine._h, query, len(query), root, C.byref(out))
            if status == sdo.SDO_OK:
                return (sdo.Result(out) if rank == root else None), {"dense": True, "native": True, "n_keys": -1, "n_slots": -1, "merge_bytes": -1}
            if status != sdo.SDO_UNSUPPORTED or b"sdo_query_partial" not in sdo.lib().sdo_last_error():
                sdo._check(status)
            # a shape for the host-side merge below (every rank takes the same branch: the decision is made from the plan)
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
