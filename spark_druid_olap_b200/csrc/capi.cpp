// extern "C" surface of libsdo_engine.so -- see include/sdo_engine.h for the contract and the
// reference call sites each entry point replaces.  No exception crosses this file.
#include <cstring>
#include <new>
#include <string>

#include "../../include/sdo_engine.h"
#include "engine.hpp"
#include "jit.hpp"

using namespace sdo;

struct sdo_engine { Engine* e; };
struct sdo_result { Result* r; };
struct sdo_partial { Partial* p; };

namespace {
thread_local std::string g_err;
thread_local std::string g_json;

template <class F>
int guard(F f) {
  try {
    g_err.clear();
    f();
    return SDO_OK;
  } catch (const Unsupported& e) { g_err = e.what(); return SDO_UNSUPPORTED; }
  catch (const Invalid& e) { g_err = e.what(); return SDO_INVALID; }
  catch (const JsonError& e) { g_err = e.what(); return SDO_INVALID; }
  catch (const NotFound& e) { g_err = e.what(); return SDO_NOT_FOUND; }
  catch (const Cancelled& e) { g_err = e.what(); return SDO_CANCELLED; }
  catch (const NoDevice& e) { g_err = e.what(); return SDO_NO_DEVICE; }
  catch (const DeviceError& e) {
    g_err = e.what();
    return e.code == cudaErrorMemoryAllocation ? SDO_OUT_OF_MEMORY : SDO_ERROR;
  } catch (const std::bad_alloc&) { g_err = "out of host memory"; return SDO_OUT_OF_MEMORY; }
  catch (const std::exception& e) { g_err = e.what(); return SDO_ERROR; }
  catch (...) { g_err = "unknown error"; return SDO_ERROR; }
}
int bad(const char* m) { g_err = m; return SDO_INVALID; }
}  // namespace

namespace sdo {
// lets other translation units (noalloc.cu) report through sdo_last_error()
void set_last_error(const std::string& m) { g_err = m; }
}  // namespace sdo

extern "C" {

int sdo_engine_create(int device, sdo_engine** out) {
  if (!out) return bad("sdo_engine_create: out is NULL");
  *out = nullptr;
  return guard([&] {
    Engine* e = new Engine(device);
    *out = new sdo_engine{e};
  });
}
void sdo_engine_destroy(sdo_engine* e) {
  if (!e) return;
  try { delete e->e; } catch (...) {}
  delete e;
}
const char* sdo_last_error(void) { return g_err.c_str(); }
const char* sdo_version(void) { return "0.9.1-b200-r1"; }

int sdo_datasource_create(sdo_engine* e, const char* name, const char* options_json) {
  if (!e || !name) return bad("sdo_datasource_create: NULL argument");
  return guard([&] { e->e->store().create(name, options_json ? options_json : ""); });
}
int sdo_datasource_drop(sdo_engine* e, const char* name) {
  if (!e || !name) return bad("sdo_datasource_drop: NULL argument");
  return guard([&] { e->e->store().drop(name); });
}
int sdo_segment_load(sdo_engine* e, const char* datasource, int64_t n_rows, int n_cols, const sdo_column* cols, int partition_num) {
  if (!e || !datasource || (n_cols > 0 && !cols)) return bad("sdo_segment_load: NULL argument");
  return guard([&] {
    std::vector<LoadColumn> lc;
    for (int i = 0; i < n_cols; i++) {
      if (!cols[i].name) throw Invalid("sdo_segment_load: column without a name");
      LoadColumn c;
      c.name = cols[i].name;
      c.kind = cols[i].kind;
      c.type = cols[i].type;
      c.src_type = cols[i].src_type;
      c.data = cols[i].data;
      c.bytes = cols[i].bytes;
      c.dict_offsets = cols[i].dict_offsets;
      c.n_dict = cols[i].n_dict;
      if (c.kind < 0 || c.kind > 2) throw Invalid("sdo_segment_load: bad column kind for '" + c.name + "'");
      lc.push_back(c);
    }
    e->e->store().load(datasource, n_rows, lc, partition_num);
  });
}
int sdo_datasource_metadata(sdo_engine* e, const char* name, const char** json_out) {
  if (!e || !name || !json_out) return bad("sdo_datasource_metadata: NULL argument");
  return guard([&] {
    g_json = e->e->store().metadata_json(name);
    *json_out = g_json.c_str();
  });
}
int sdo_datasource_segments(sdo_engine* e, const char* name, const char** json_out) {
  if (!e || !name || !json_out) return bad("sdo_datasource_segments: NULL argument");
  return guard([&] {
    g_json = e->e->store().segments_json(name);
    *json_out = g_json.c_str();
  });
}

int sdo_query_execute(sdo_engine* e, const char* query_json, size_t len, sdo_result** out) {
  if (!e || !query_json || !out) return bad("sdo_query_execute: NULL argument");
  *out = nullptr;
  return guard([&] {
    Result* r = e->e->execute(query_json, len);
    *out = new sdo_result{r};
  });
}
int sdo_query_explain(sdo_engine* e, const char* query_json, size_t len, const char** json_out) {
  if (!e || !query_json || !json_out) return bad("sdo_query_explain: NULL argument");
  return guard([&] {
    g_json = e->e->explain(query_json, len);
    *json_out = g_json.c_str();
  });
}
int sdo_query_cancel(sdo_engine* e, const char* query_id) {
  if (!e || !query_id) return bad("sdo_query_cancel: NULL argument");
  return guard([&] { e->e->cancel(query_id); });
}

int64_t sdo_result_num_rows(const sdo_result* r) { return r ? r->r->n_rows : 0; }
int sdo_result_num_columns(const sdo_result* r) { return r ? (int)r->r->columns.size() : 0; }
int sdo_result_column_get(const sdo_result* r, int i, sdo_result_column* out) {
  if (!r || !out) return bad("sdo_result_column_get: NULL argument");
  if (i < 0 || i >= (int)r->r->columns.size()) return bad("sdo_result_column_get: column index out of range");
  const ResultColumn& c = r->r->columns[i];
  memset(out, 0, sizeof *out);
  out->name = c.name.c_str();
  out->type = c.type;
  switch (c.type) {
    case SDO_RES_INT64: case SDO_RES_DIM_INT64: case SDO_RES_TIMESTAMP: out->values = c.raw ? c.raw : (const void*)c.i64.data(); break;
    case SDO_RES_FLOAT64: out->values = c.raw ? c.raw : (const void*)c.f64.data(); break;
    case SDO_RES_STRING:
      out->values = c.raw ? c.raw : (const void*)c.ids.data();
      out->dict_offsets = c.dict->offsets.data();
      out->dict_bytes = c.dict->bytes.data();
      out->n_dict = c.dict->size();
      break;
  }
  return SDO_OK;
}
const uint8_t* sdo_result_hll_registers(const sdo_result* r, int i) {
  if (!r || i < 0 || i >= (int)r->r->columns.size()) return nullptr;
  ResultColumn& c = r->r->columns[i];
  if (c.hll.empty() && c.hll_rows && !c.hll_pick.empty()) {
    // first request: put the sketches into result order (a result is read by the one task that owns it)
    c.hll.resize(c.hll_pick.size() * 512);
    for (size_t k = 0; k < c.hll_pick.size(); k++) memcpy(&c.hll[k * 512], c.hll_rows + c.hll_pick[k] * 512, 512);
    c.hll_rows = nullptr;
    c.hll_pick.clear();
  }
  return c.hll.empty() ? nullptr : c.hll.data();
}
const char* sdo_result_stats_json(const sdo_result* r) { return r ? r->r->stats.c_str() : ""; }
void sdo_result_free(sdo_result* r) {
  if (!r) return;
  delete r->r;
  delete r;
}

int sdo_dimension_dictionary(sdo_engine* e, const char* datasource, const char* dimension, int64_t* n_out, const int64_t** offsets_out,
                             const char** bytes_out, int64_t* cardinality_out) {
  if (!e || !datasource || !dimension || !n_out || !offsets_out || !bytes_out) return bad("sdo_dimension_dictionary: NULL argument");
  return guard([&] {
    auto ds = e->e->store().get(datasource);
    int di = ds->dim_index(dimension);
    if (di < 0) throw NotFound(std::string("unknown dimension '") + dimension + "'");
    // keep the snapshot alive in a per-thread slot: dictionaries are replaced, never mutated
    thread_local std::shared_ptr<Dictionary> hold;
    hold = e->e->store().dictionary(datasource, dimension);
    const bool is_dict = ds->dims[di].enc == DimEncoding::Dict;
    *n_out = is_dict ? hold->size() : 0;
    *offsets_out = hold->offsets.data();
    *bytes_out = hold->bytes.data();
    if (cardinality_out) *cardinality_out = ds->dims[di].cardinality;
  });
}
int sdo_dimension_extend(sdo_engine* e, const char* datasource, const char* dimension, int64_t n_values, const int64_t* value_offsets,
                         const char* value_bytes, int64_t min_cardinality) {
  if (!e || !datasource || !dimension || (n_values > 0 && (!value_offsets || !value_bytes))) return bad("sdo_dimension_extend: NULL argument");
  return guard([&] {
    Dictionary d;
    for (int64_t i = 0; i < n_values; i++) d.push(value_bytes + value_offsets[i], (size_t)(value_offsets[i + 1] - value_offsets[i]));
    e->e->store().extend_dimension(datasource, dimension, n_values > 0 ? &d : nullptr, min_cardinality);
  });
}

int sdo_query_partial(sdo_engine* e, const char* query_json, size_t len, sdo_partial** out) {
  if (!e || !query_json || !out) return bad("sdo_query_partial: NULL argument");
  *out = nullptr;
  return guard([&] {
    Partial* p = e->e->partial(query_json, len);
    *out = new sdo_partial{p};
  });
}
int sdo_partial_view_get(sdo_partial* p, sdo_partial_view* out) {
  if (!p || !out) return bad("sdo_partial_view_get: NULL argument");
  return guard([&] { p->p->engine->partial_view(p->p, out); });
}
int sdo_partial_export(sdo_partial* p, void* dst_dev, int64_t dst_bytes) {
  if (!p || !dst_dev) return bad("sdo_partial_export: NULL argument");
  return guard([&] { p->p->engine->partial_export(p->p, dst_dev, dst_bytes); });
}
int sdo_partial_merge_gathered(sdo_partial* p, const void* gathered_dev, int n_ranks) {
  if (!p || !gathered_dev) return bad("sdo_partial_merge_gathered: NULL argument");
  return guard([&] { p->p->engine->partial_merge_gathered(p->p, gathered_dev, n_ranks); });
}
int sdo_partial_partition(sdo_partial* p, int n_ranks, int64_t* counts_out) {
  if (!p || !counts_out) return bad("sdo_partial_partition: NULL argument");
  return guard([&] { p->p->engine->partial_partition(p->p, n_ranks, counts_out); });
}
int sdo_partial_merge_rows(sdo_partial* p, const void* keys_dev, const void* slots_dev, int64_t n) {
  if (!p || (n > 0 && (!keys_dev || !slots_dev))) return bad("sdo_partial_merge_rows: NULL argument");
  return guard([&] { p->p->engine->partial_merge_rows(p->p, keys_dev, slots_dev, n); });
}
void* sdo_partial_stream(sdo_partial* p) { return p && p->p->ctx ? (void*)p->p->ctx->stream : nullptr; }
int sdo_partial_finish(sdo_partial* p, sdo_result** out) {
  if (!p || !out) return bad("sdo_partial_finish: NULL argument");
  *out = nullptr;
  return guard([&] {
    Result* r = p->p->engine->partial_finish(p->p);
    *out = new sdo_result{r};
  });
}
void sdo_partial_free(sdo_partial* p) {
  if (!p) return;
  try { p->p->engine->partial_free(p->p); } catch (...) {}
  delete p;
}

int sdo_datasource_save(sdo_engine* e, const char* name, const char* path) {
  if (!e || !name || !path) return bad("sdo_datasource_save: NULL argument");
  return guard([&] { e->e->store().save(name, path); });
}
int sdo_datasource_restore(sdo_engine* e, const char* name, const char* path) {
  if (!e || !name || !path) return bad("sdo_datasource_restore: NULL argument");
  return guard([&] { e->e->store().restore(name, path); });
}

int sdo_comm_unique_id(void* id_out, const char* nccl_library) {
  if (!id_out) return bad("sdo_comm_unique_id: NULL argument");
  return guard([&] {
    nccl_load(nccl_library);
    nccl_unique_id(id_out);
  });
}
int sdo_engine_comm_init(sdo_engine* e, const void* id, int rank, int world, const char* nccl_library) {
  if (!e || !id) return bad("sdo_engine_comm_init: NULL argument");
  return guard([&] {
    nccl_load(nccl_library);
    e->e->comm_init(id, rank, world);
  });
}
int sdo_query_execute_sharded(sdo_engine* e, const char* query_json, size_t len, int root, sdo_result** out) {
  if (!e || !query_json || !out) return bad("sdo_query_execute_sharded: NULL argument");
  *out = nullptr;
  return guard([&] {
    Result* r = e->e->execute_sharded(query_json, len, root);
    if (r) *out = new sdo_result{r};
  });
}

int sdo_query_execute_sharded_at(sdo_engine* e, const char* query_json, size_t len, int root, int64_t turn, sdo_result** out) {
  if (!e || !query_json || !out) return bad("sdo_query_execute_sharded_at: NULL argument");
  if (turn < 0) return bad("sdo_query_execute_sharded_at: turn must be >= 0");
  *out = nullptr;
  return guard([&] {
    Result* r = e->e->execute_sharded(query_json, len, root, turn);
    if (r) *out = new sdo_result{r};
  });
}

int sdo_jit_pending(void) { return jit_pending(); }

int sdo_engine_comm_next_turn(sdo_engine* e, int64_t* turn_out) {
  if (!e || !turn_out) return bad("sdo_engine_comm_next_turn: NULL argument");
  return guard([&] { *turn_out = e->e->next_turn(); });
}

}  // extern "C"
