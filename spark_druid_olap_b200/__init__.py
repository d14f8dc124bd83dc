"""ctypes binding of libsdo_engine.so (include/sdo_engine.h) -- what tests and bench.py use in
place of the JNI shim a Spark executor would load.

The binding is deliberately thin: every method is one C-ABI call.  There is no Python or CPU
execution path; if the CUDA library is missing or no device is present, calls fail loudly.

Error mapping follows the reference: any failure surfaces as DruidDataSourceException
(/root/reference/src/main/scala/org/sparklinedata/druid/DruidDataSourceException.scala), with
UnsupportedQuery (status SDO_UNSUPPORTED) as the signal to leave the Spark plan un-rewritten.
"""
import ctypes as C
import json
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ENGINE_LIB = os.path.join(_HERE, "libsdo_engine.so")

SDO_OK, SDO_ERROR, SDO_UNSUPPORTED, SDO_INVALID, SDO_NOT_FOUND, SDO_CANCELLED, SDO_NO_DEVICE, SDO_OUT_OF_MEMORY = range(8)
COL_TIME, COL_DIM, COL_METRIC = 0, 1, 2
T_INT64, T_FLOAT64, T_STRING, T_CODES, T_INT32 = 0, 1, 2, 3, 4
SRC_STRING, SRC_INT32, SRC_INT64 = 0, 1, 2
RES_INT64, RES_FLOAT64, RES_STRING, RES_DIM_INT64, RES_TIMESTAMP = 0, 1, 2, 3, 4
HLL_M = 512


class DruidDataSourceException(Exception):
    def __init__(self, msg, status=SDO_ERROR):
        super().__init__(msg)
        self.status = status


class UnsupportedQuery(DruidDataSourceException):
    pass


class InvalidQuery(DruidDataSourceException):
    pass


class NoDevice(DruidDataSourceException):
    pass


class QueryCancelled(DruidDataSourceException):
    pass


class _Column(C.Structure):
    _fields_ = [("name", C.c_char_p), ("kind", C.c_int), ("type", C.c_int), ("data", C.c_void_p), ("bytes", C.c_void_p),
                ("dict_offsets", C.c_void_p), ("n_dict", C.c_int64), ("src_type", C.c_int)]


class _ResultColumn(C.Structure):
    _fields_ = [("name", C.c_char_p), ("type", C.c_int), ("values", C.c_void_p), ("dict_offsets", C.c_void_p),
                ("dict_bytes", C.c_void_p), ("n_dict", C.c_int64)]


class PartialView(C.Structure):
    _fields_ = [("dense", C.c_int), ("n_keys", C.c_int64), ("n_slots", C.c_int), ("keys", C.c_void_p), ("slots", C.c_void_p),
                ("slot_ops", C.POINTER(C.c_int)), ("hll", C.c_void_p), ("n_hll", C.c_int), ("blob_bytes", C.c_int64)]


_lib = None


def lib():
    """Load the CUDA library; raise if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(ENGINE_LIB):
        raise ImportError(f"{ENGINE_LIB} is missing: run `python -m spark_druid_olap_b200.build` (nvcc, sm_100a)")
    L = C.CDLL(ENGINE_LIB)
    vp, cp, i64, sz = C.c_void_p, C.c_char_p, C.c_int64, C.c_size_t
    sig = {
        "sdo_engine_create": (C.c_int, [C.c_int, C.POINTER(vp)]),
        "sdo_engine_destroy": (None, [vp]),
        "sdo_last_error": (cp, []),
        "sdo_version": (cp, []),
        "sdo_datasource_create": (C.c_int, [vp, cp, cp]),
        "sdo_datasource_drop": (C.c_int, [vp, cp]),
        "sdo_segment_load": (C.c_int, [vp, cp, i64, C.c_int, C.POINTER(_Column), C.c_int]),
        "sdo_datasource_metadata": (C.c_int, [vp, cp, C.POINTER(cp)]),
        "sdo_datasource_segments": (C.c_int, [vp, cp, C.POINTER(cp)]),
        "sdo_query_execute": (C.c_int, [vp, cp, sz, C.POINTER(vp)]),
        "sdo_query_explain": (C.c_int, [vp, cp, sz, C.POINTER(cp)]),
        "sdo_query_cancel": (C.c_int, [vp, cp]),
        "sdo_result_num_rows": (i64, [vp]),
        "sdo_result_num_columns": (C.c_int, [vp]),
        "sdo_result_column_get": (C.c_int, [vp, C.c_int, C.POINTER(_ResultColumn)]),
        "sdo_result_hll_registers": (vp, [vp, C.c_int]),
        "sdo_result_stats_json": (cp, [vp]),
        "sdo_result_free": (None, [vp]),
        "sdo_query_partial": (C.c_int, [vp, cp, sz, C.POINTER(vp)]),
        "sdo_partial_view_get": (C.c_int, [vp, C.POINTER(PartialView)]),
        "sdo_partial_export": (C.c_int, [vp, vp, i64]),
        "sdo_partial_merge_gathered": (C.c_int, [vp, vp, C.c_int]),
        "sdo_partial_partition": (C.c_int, [vp, C.c_int, C.POINTER(i64)]),
        "sdo_partial_merge_rows": (C.c_int, [vp, vp, vp, i64]),
        "sdo_partial_stream": (vp, [vp]),
        "sdo_partial_finish": (C.c_int, [vp, C.POINTER(vp)]),
        "sdo_partial_free": (None, [vp]),
        "sdo_groupby_sum_noalloc": (C.c_int, [C.c_int, vp, vp, i64, C.c_int, vp, vp, vp]),
        "sdo_dimension_dictionary": (C.c_int, [vp, cp, cp, C.POINTER(i64), C.POINTER(vp), C.POINTER(vp), C.POINTER(i64)]),
        "sdo_dimension_extend": (C.c_int, [vp, cp, cp, i64, vp, vp, i64]),
        "sdo_datasource_save": (C.c_int, [vp, cp, cp]),
        "sdo_datasource_restore": (C.c_int, [vp, cp, cp]),
        "sdo_comm_unique_id": (C.c_int, [vp, cp]),
        "sdo_engine_comm_init": (C.c_int, [vp, vp, C.c_int, C.c_int, cp]),
        "sdo_query_execute_sharded": (C.c_int, [vp, cp, sz, C.c_int, C.POINTER(vp)]),
        "sdo_query_execute_sharded_at": (C.c_int, [vp, cp, sz, C.c_int, i64, C.POINTER(vp)]),
        "sdo_engine_comm_next_turn": (C.c_int, [vp, C.POINTER(i64)]),
        "sdo_jit_pending": (C.c_int, []),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype = res
        f.argtypes = args
    _lib = L
    return L


EXPORTED_SYMBOLS = [
    "sdo_engine_create", "sdo_engine_destroy", "sdo_last_error", "sdo_version", "sdo_datasource_create", "sdo_datasource_drop",
    "sdo_segment_load", "sdo_datasource_metadata", "sdo_datasource_segments", "sdo_query_execute", "sdo_query_explain",
    "sdo_query_cancel", "sdo_result_num_rows", "sdo_result_num_columns", "sdo_result_column_get", "sdo_result_hll_registers",
    "sdo_result_stats_json", "sdo_result_free", "sdo_query_partial", "sdo_partial_view_get", "sdo_partial_partition",
    "sdo_partial_merge_rows", "sdo_partial_stream", "sdo_partial_finish", "sdo_partial_free", "sdo_groupby_sum_noalloc",
    "sdo_dimension_dictionary", "sdo_dimension_extend", "sdo_partial_export", "sdo_partial_merge_gathered",
    "sdo_comm_unique_id", "sdo_engine_comm_init", "sdo_query_execute_sharded", "sdo_query_execute_sharded_at", "sdo_engine_comm_next_turn", "sdo_jit_pending", "sdo_datasource_save", "sdo_datasource_restore",
]

NOALLOC_MAX_KEYS = 24
COMM_ID_BYTES = 128
NOALLOC_MAX_ROWS = 1 << 32


def groupby_sum_noalloc(device, keys_ptr, values_ptr, n_rows, n_keys, counts_out_ptr, sums_out_ptr, stream=None):
    """sdo_groupby_sum_noalloc: count and double sum per one-byte key without any device allocation.
    For hosts that may compute on the GPU but not allocate there; all pointers are caller-owned device
    memory.  See include/sdo_engine.h for the limits, the first-call set-up and the sharing rule."""
    _check(lib().sdo_groupby_sum_noalloc(device, keys_ptr, values_ptr, n_rows, n_keys, counts_out_ptr, sums_out_ptr, stream))


def _check(status):
    if status == SDO_OK:
        return
    msg = lib().sdo_last_error().decode("utf-8", "replace")
    cls = {SDO_UNSUPPORTED: UnsupportedQuery, SDO_INVALID: InvalidQuery, SDO_NO_DEVICE: NoDevice,
           SDO_CANCELLED: QueryCancelled}.get(status, DruidDataSourceException)
    raise cls(msg, status)


def _as_json_bytes(q):
    if isinstance(q, bytes):
        return q
    if isinstance(q, str):
        return q.encode("utf-8")
    if hasattr(q, "to_json"):
        return q.to_json().encode("utf-8")
    return json.dumps(q).encode("utf-8")


# ---- load-path column helpers ------------------------------------------------------------------
class Column:
    """One column of a load call; keeps the numpy buffers alive for the duration of the call."""

    def __init__(self, name, kind, typ, data, bytes_=None, dict_offsets=None, n_dict=0, src_type=SRC_STRING):
        self.name, self.kind, self.type = name, kind, typ
        self.data, self.bytes, self.dict_offsets, self.n_dict, self.src_type = data, bytes_, dict_offsets, n_dict, src_type

    def c_struct(self):
        s = _Column()
        s.name = self.name.encode()
        s.kind, s.type, s.src_type, s.n_dict = self.kind, self.type, self.src_type, self.n_dict
        s.data = self.data.ctypes.data if self.data is not None else None
        s.bytes = self.bytes.ctypes.data if self.bytes is not None else None
        s.dict_offsets = self.dict_offsets.ctypes.data if self.dict_offsets is not None else None
        return s


def _pack_strings(values):
    enc = [v.encode("utf-8") if isinstance(v, str) else bytes(v) for v in values]
    offsets = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        offsets[1:] = np.cumsum([len(b) for b in enc])
    blob = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8).copy()
    return offsets, blob


def time_column(name, millis):
    return Column(name, COL_TIME, T_INT64, np.ascontiguousarray(millis, dtype=np.int64))


def dim_codes(name, codes, dictionary, src_type=SRC_STRING):
    offsets, blob = _pack_strings(dictionary)
    return Column(name, COL_DIM, T_CODES, np.ascontiguousarray(codes, dtype=np.int32), blob, offsets, len(dictionary), src_type)


def dim_strings(name, values, src_type=SRC_STRING):
    offsets, blob = _pack_strings(values)
    return Column(name, COL_DIM, T_STRING, offsets, blob, None, 0, src_type)


def dim_int(name, values, src_type=SRC_INT32):
    return Column(name, COL_DIM, T_INT64, np.ascontiguousarray(values, dtype=np.int64), src_type=src_type)


def metric(name, values):
    a = np.asarray(values)
    if a.dtype.kind == "f":
        return Column(name, COL_METRIC, T_FLOAT64, np.ascontiguousarray(a, dtype=np.float64))
    return Column(name, COL_METRIC, T_INT64, np.ascontiguousarray(a, dtype=np.int64))


# ---- result ------------------------------------------------------------------------------------
class Result:
    """Columnar query result copied out of an sdo_result (which is freed immediately)."""

    def __init__(self, handle):
        L = lib()
        try:
            self.n_rows = int(L.sdo_result_num_rows(handle))
            self.names, self.types, self.columns, self.hll = [], {}, {}, {}
            n = self.n_rows
            for i in range(L.sdo_result_num_columns(handle)):
                rc = _ResultColumn()
                _check(L.sdo_result_column_get(handle, i, C.byref(rc)))
                name = rc.name.decode()
                self.names.append(name)
                self.types[name] = rc.type
                if rc.type in (RES_INT64, RES_DIM_INT64, RES_TIMESTAMP):
                    v = np.ctypeslib.as_array(C.cast(rc.values, C.POINTER(C.c_int64)), (n,)).copy() if n else np.zeros(0, np.int64)
                elif rc.type == RES_FLOAT64:
                    v = np.ctypeslib.as_array(C.cast(rc.values, C.POINTER(C.c_double)), (n,)).copy() if n else np.zeros(0, np.float64)
                else:
                    ids = np.ctypeslib.as_array(C.cast(rc.values, C.POINTER(C.c_uint32)), (n,)).copy() if n else np.zeros(0, np.uint32)
                    nd = int(rc.n_dict)
                    offs = np.ctypeslib.as_array(C.cast(rc.dict_offsets, C.POINTER(C.c_int64)), (nd + 1,))
                    raw = C.string_at(rc.dict_bytes, int(offs[nd])) if nd else b""
                    used = np.unique(ids)
                    strs = {int(u): raw[offs[u]:offs[u + 1]].decode("utf-8") for u in used}
                    v = np.array([strs[int(x)] for x in ids], dtype=object)
                self.columns[name] = v
                regs = L.sdo_result_hll_registers(handle, i)
                if regs and n:
                    self.hll[name] = np.ctypeslib.as_array(C.cast(regs, C.POINTER(C.c_uint8)), (n, HLL_M)).copy()
            self.stats = json.loads(L.sdo_result_stats_json(handle).decode())
        finally:
            L.sdo_result_free(handle)

    def __getitem__(self, name):
        return self.columns[name]

    def rows(self):
        """Rows as tuples in column order (dimension values as str, as the reference's ResultRow events)."""
        cols = [self.columns[n] for n in self.names]
        return [tuple(c[i].item() if hasattr(c[i], "item") else c[i] for c in cols) for i in range(self.n_rows)]


class Partial:
    """Device-resident partial aggregate of one rank (sdo_query_partial)."""

    def __init__(self, handle):
        self.handle = handle

    def view(self):
        v = PartialView()
        _check(lib().sdo_partial_view_get(self.handle, C.byref(v)))
        return v

    def stream(self):
        return lib().sdo_partial_stream(self.handle)

    def export(self, dst_ptr, dst_bytes):
        _check(lib().sdo_partial_export(self.handle, dst_ptr, dst_bytes))

    def merge_gathered(self, gathered_ptr, n_ranks):
        _check(lib().sdo_partial_merge_gathered(self.handle, gathered_ptr, n_ranks))

    def partition(self, n_ranks):
        counts = (C.c_int64 * n_ranks)()
        _check(lib().sdo_partial_partition(self.handle, n_ranks, counts))
        return [int(c) for c in counts]

    def merge_rows(self, keys_ptr, slots_ptr, n):
        _check(lib().sdo_partial_merge_rows(self.handle, keys_ptr, slots_ptr, n))

    def finish(self):
        out = C.c_void_p()
        _check(lib().sdo_partial_finish(self.handle, C.byref(out)))
        return Result(out)

    def free(self):
        if self.handle:
            lib().sdo_partial_free(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Engine:
    """One engine per GPU (device >= 0) or a planning-only engine (device < 0)."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        _check(lib().sdo_engine_create(device, C.byref(self._h)))
        self.device = device

    def close(self):
        if self._h:
            lib().sdo_engine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def create_datasource(self, name, **options):
        _check(lib().sdo_datasource_create(self._h, name.encode(), json.dumps(options).encode() if options else None))

    def drop_datasource(self, name):
        _check(lib().sdo_datasource_drop(self._h, name.encode()))

    def save_datasource(self, name, path):
        """Encoded segments, dictionaries and time index of `name` into one file."""
        _check(lib().sdo_datasource_save(self._h, name.encode(), os.fsencode(path)))

    def restore_datasource(self, name, path):
        """A file written by save_datasource back into HBM under `name` (no re-encoding, no time sort)."""
        _check(lib().sdo_datasource_restore(self._h, name.encode(), os.fsencode(path)))

    def load(self, datasource, columns, partition=0):
        n_rows = None
        for c in columns:
            n = len(c.data) - 1 if c.type == T_STRING else len(c.data)
            if n_rows is None:
                n_rows = n
            elif n != n_rows:
                raise ValueError(f"column {c.name} has {n} rows, expected {n_rows}")
        arr = (_Column * len(columns))(*[c.c_struct() for c in columns])
        _check(lib().sdo_segment_load(self._h, datasource.encode(), n_rows or 0, len(columns), arr, partition))

    def metadata(self, name):
        out = C.c_char_p()
        _check(lib().sdo_datasource_metadata(self._h, name.encode(), C.byref(out)))
        return json.loads(out.value.decode())

    def segments(self, name):
        out = C.c_char_p()
        _check(lib().sdo_datasource_segments(self._h, name.encode(), C.byref(out)))
        return json.loads(out.value.decode())

    def dimension_dictionary(self, datasource, dimension):
        """(list of dictionary strings, cardinality); the list is empty for integer dimensions."""
        n, card = C.c_int64(), C.c_int64()
        offs, raw = C.c_void_p(), C.c_void_p()
        _check(lib().sdo_dimension_dictionary(self._h, datasource.encode(), dimension.encode(), C.byref(n), C.byref(offs), C.byref(raw), C.byref(card)))
        if n.value == 0:
            return [], int(card.value)
        o = np.ctypeslib.as_array(C.cast(offs, C.POINTER(C.c_int64)), (n.value + 1,))
        b = C.string_at(raw, int(o[n.value]))
        return [b[o[i]:o[i + 1]].decode("utf-8") for i in range(n.value)], int(card.value)

    def extend_dimension(self, datasource, dimension, values=(), min_cardinality=0):
        """Merge `values` into a string dimension's dictionary / raise an integer dimension's cardinality."""
        offsets, blob = _pack_strings(list(values))
        _check(lib().sdo_dimension_extend(self._h, datasource.encode(), dimension.encode(), len(values), offsets.ctypes.data if len(values) else None,
                                          blob.ctypes.data if len(values) else None, int(min_cardinality)))

    def execute(self, query):
        """QuerySpec.executeQuery: query JSON in, columnar result out (DruidQuerySpec.scala:598-599)."""
        q = _as_json_bytes(query)
        out = C.c_void_p()
        _check(lib().sdo_query_execute(self._h, q, len(q), C.byref(out)))
        return Result(out)

    def execute_raw(self, qbytes):
        """Execute and return the opaque sdo_result handle (bench: avoids the numpy copy-out)."""
        out = C.c_void_p()
        _check(lib().sdo_query_execute(self._h, qbytes, len(qbytes), C.byref(out)))
        return out

    def explain(self, query):
        q = _as_json_bytes(query)
        out = C.c_char_p()
        _check(lib().sdo_query_explain(self._h, q, len(q), C.byref(out)))
        return json.loads(out.value.decode())

    def sharded_next_turn(self):
        """The next free turn of this engine's communicator (sdo_engine_comm_next_turn)."""
        t = C.c_int64()
        _check(lib().sdo_engine_comm_next_turn(self._h, C.byref(t)))
        return int(t.value)

    def cancel(self, query_id):
        _check(lib().sdo_query_cancel(self._h, query_id.encode()))

    def partial(self, query):
        q = _as_json_bytes(query)
        out = C.c_void_p()
        _check(lib().sdo_query_partial(self._h, q, len(q), C.byref(out)))
        return Partial(out)
