"""Sharded execution: one engine (process) per GPU, partial aggregates merged over NCCL INSIDE the library.

Replaces what the reference does when it queries historicals directly: one partial query per
(server, segment group) partition (DruidRDD.scala:244-269) and a Spark-side re-aggregation of the
partials (sources/druid/DruidStrategy.scala:353-359) with the merge algebra of
sources/druid/PostAggregate.scala:39-60 (count/sum -> sum, min -> min, max -> max; HLL -> register max).

This module is set-up glue around the C ABI, which a JVM host would do the same way over its own channel:
  * align_datasource: give every rank the same id space (sdo_dimension_dictionary / sdo_dimension_extend);
  * global_time_boundary / prepare_query: number time-derived keys on one grid (context.gpuTimeBoundary);
  * init_comm: hand every engine the NCCL unique id (sdo_comm_unique_id / sdo_engine_comm_init);
  * execute_sharded: ONE call, sdo_query_execute_sharded[_at] -- scan, merge (all-gather + fold, per-slot
    ncclReduce, or hash-partitioned ncclSend / ncclRecv exchange) and finish are queued by the library.
torch.distributed is used for the set-up steps only (an object all-gather of dictionaries, a max over time
boundaries, the broadcast of the unique id); no collective of a query goes through it.
"""
import json
import os

import torch
import torch.distributed as dist

import spark_druid_olap_b200 as sdo

_boundary_cache = {}
_aligned = set()


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


def execute_sharded(engine, query, group=None, root=0, turn=None):
    """Run `query` over every rank's local segments with the merge inside the library (sdo_query_execute_sharded).
    Returns (result_or_None, info).  Dense key spaces and top-K queries: the merged result on `root`, None elsewhere.
    Sparse key spaces without a top-K: every rank gets its disjoint share of the groups (the caller unions them, as Spark
    unions partitions).  `turn` (>= 0, the same on every rank, consecutive per engine) orders the collectives of queries
    that run concurrently on one engine; None = the caller issues sharded queries one at a time, in the same order."""
    import ctypes as C
    if id(engine) not in _comms:
        init_comm(engine, group)
    qb = prepare_query(engine, query, group)
    out = C.c_void_p()
    L = sdo.lib()
    if turn is None:
        sdo._check(L.sdo_query_execute_sharded(engine._h, qb, len(qb), root, C.byref(out)))
    else:
        sdo._check(L.sdo_query_execute_sharded_at(engine._h, qb, len(qb), root, int(turn), C.byref(out)))
    if not out:
        return None, {"native": True, "merge": None}
    res = sdo.Result(out)
    return res, {"native": True, "merge": res.stats.get("shardMerge"), "merge_bytes": res.stats.get("exchangeBytes", 0)}
