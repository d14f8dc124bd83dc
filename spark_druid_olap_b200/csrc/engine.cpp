#include "engine.hpp"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <numeric>

#include "../../include/sdo_engine.h"
#include "jit.hpp"
#include "xxh64.hpp"

namespace sdo {

static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// Spark HyperLogLogPlusPlusHelper.query for p = 9 (relativeSD = 0.05), restated: raw estimate
// alpha*m^2 / sum(2^-M[j]), linear counting below the p = 9 threshold (400), rounded to a long.
// The empirical bias-correction tables of HLL++ (used by Spark when E < 5m) are not available
// offline; in that band (roughly 400 < E < 2560) the estimate is the uncorrected E and is
// reported as such -- the registers themselves (sdo_result_hll_registers) are exact.
double hll_estimate(const uint8_t* regs) {
  const double m = HLL_M;
  static const struct Pow2 {
    double inv[64];
    Pow2() { for (int i = 0; i < 64; i++) inv[i] = 1.0 / (double)(1ull << i); }  // exact powers of two: same sum as dividing
  } pow2;
  // Four independent chains (the terms are powers of two >= 2^-56 and their partial sums stay far inside 53 bits
  // for any real sketch, so the sum is exact in every order -- same value as the sequential loop, a quarter of the latency).
  double z0 = 0, z1 = 0, z2 = 0, z3 = 0;
  int zeros = 0;
  for (int j = 0; j < HLL_M; j += 4) {
    z0 += pow2.inv[regs[j] & 63];
    z1 += pow2.inv[regs[j + 1] & 63];
    z2 += pow2.inv[regs[j + 2] & 63];
    z3 += pow2.inv[regs[j + 3] & 63];
    zeros += (regs[j] == 0) + (regs[j + 1] == 0) + (regs[j + 2] == 0) + (regs[j + 3] == 0);
  }
  return hll_estimate_from((z0 + z1) + (z2 + z3), zeros);
}
// true when Spark would have subtracted an empirical bias from this sketch's estimate (raw E <= 5m and linear counting
// not chosen): the only inputs for which the value column differs from Spark's
bool hll_needs_bias_tables(double zinv, double V) {
  const double m = HLL_M;
  const double E = ((0.7213 / (1.0 + 1.079 / m)) * m * m) / zinv;
  if (E > 5.0 * m) return false;
  if (V > 0 && m * std::log(m / V) <= 400.0) return false;
  return true;
}
double hll_estimate_from(double zinv, double V) {
  const double m = HLL_M;
  const double alpha_m2 = (0.7213 / (1.0 + 1.079 / m)) * m * m;
  const double E = alpha_m2 / zinv;
  double est = E;
  if (V > 0) {
    double H = m * std::log(m / V);
    if (H <= 400.0) est = H;
  }
  return std::floor(est + 0.5);  // Math.round
}

Engine::Engine(int device) : device_(device), store_(device) {
  if (device_ >= 0) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) throw NoDevice(std::string("no CUDA device available: ") + cudaGetErrorString(e));
    if (device_ >= n) throw NoDevice("CUDA device " + std::to_string(device_) + " does not exist (" + std::to_string(n) + " visible)");
    SDO_CUDA(cudaSetDevice(device_));
    cudaDeviceProp prop;
    SDO_CUDA(cudaGetDeviceProperties(&prop, device_));
    sm_count_ = prop.multiProcessorCount;
    if (prop.major < 10) throw NoDevice(std::string("device '") + prop.name + "' is not sm_100 class; this library is built for sm_100a only");
    cudaMemPool_t pool;
    SDO_CUDA(cudaDeviceGetDefaultMemPool(&pool, device_));
    uint64_t thr = UINT64_MAX;
    SDO_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
  }
}

Engine::~Engine() {
  if (device_ >= 0) {
    cudaSetDevice(device_);
    cudaDeviceSynchronize();
    if (comm_) nccl_comm_destroy(comm_);
    for (QueryCtx* c : free_ctx_) {
      if (c->stream) cudaStreamDestroy(c->stream);
      for (auto& e : c->ev)
        if (e) cudaEventDestroy(e);
      if (c->pinned_in) cudaFreeHost(c->pinned_in);
      if (c->pinned_out) cudaFreeHost(c->pinned_out);
      if (c->counters) cudaFree(c->counters);
      if (c->scratch) cudaFree(c->scratch);
      delete c;
    }
  }
}

QueryCtx* Engine::acquire_ctx() {
  // every query starts here: a caller thread that has never touched CUDA has device 0 current, whatever the engine's is
  if (device_ >= 0) SDO_CUDA(cudaSetDevice(device_));
  {
    std::lock_guard<std::mutex> lk(ctx_mu_);
    if (!free_ctx_.empty()) {
      QueryCtx* c = free_ctx_.back();
      free_ctx_.pop_back();
      return c;
    }
  }
  QueryCtx* c = new QueryCtx();
  SDO_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  for (auto& e : c->ev) SDO_CUDA(cudaEventCreate(&e));
  SDO_CUDA(cudaMalloc((void**)&c->counters, COUNTER_BYTES));
  SDO_CUDA(cudaMemset(c->counters, 0, COUNTER_BYTES));
  c->scratch_bytes = 2u << 20;
  SDO_CUDA(cudaMalloc((void**)&c->scratch, c->scratch_bytes));
  ensure_pinned(&c->pinned_in, &c->pinned_in_bytes, 1 << 20);
  ensure_pinned(&c->pinned_out, &c->pinned_out_bytes, 4 << 20);
  return c;
}
void* Engine::take_scratch(QueryCtx* c, size_t bytes) {
  const size_t at = round_up(c->scratch_used, 256);
  if (at + bytes > c->scratch_bytes) return nullptr;
  c->scratch_used = at + bytes;
  return c->scratch + at;
}
void Engine::release_ctx(QueryCtx* c) {
  std::lock_guard<std::mutex> lk(ctx_mu_);
  free_ctx_.push_back(c);
}
void Engine::ensure_pinned(void** p, size_t* cur, size_t need) {
  if (*cur >= need) return;
  if (*p) cudaFreeHost(*p);
  *p = nullptr;
  size_t n = std::max(need, *cur * 2);
  SDO_CUDA(cudaMallocHost(p, n));
  *cur = n;
}

// Cancellation (replaces CancellableHolder / TaskCancelHandler, DruidClient.scala:82-103, DruidRDD.scala:453-481).
// A query registers its id while it runs; cancel() marks a running id, and the mark dies with the query.  A cancel
// that arrives for an id that is not running (the canceller can race ahead of the execute call) is kept for
// CANCEL_KEEP_MS only, so cancelling a finished or unknown query cannot abort a later retry that reuses the id
// (Spark task retries and speculative attempts reuse <queryId>-<partition>, DruidRDD.scala:80).  The flag is polled
// before the scan is queued and after it has been queued (before results are read back).
static constexpr double CANCEL_KEEP_MS = 5000.0;
void Engine::cancel(const std::string& qid) {
  std::lock_guard<std::mutex> lk(cancel_mu_);
  const double now = now_ms();
  for (auto it = cancelled_.begin(); it != cancelled_.end();)
    it = (now - it->second > CANCEL_KEEP_MS && !running_.count(it->first)) ? cancelled_.erase(it) : std::next(it);
  cancelled_[qid] = now;
}
void Engine::check_cancel(const std::string& qid) {
  if (qid.empty()) return;
  std::lock_guard<std::mutex> lk(cancel_mu_);
  auto it = cancelled_.find(qid);
  if (it == cancelled_.end()) return;
  const bool fresh = now_ms() - it->second <= CANCEL_KEEP_MS || running_since_.count(qid) && it->second >= running_since_[qid];
  cancelled_.erase(it);
  if (fresh) throw Cancelled("query '" + qid + "' was cancelled");
}
Engine::RunningQuery::RunningQuery(Engine* e, const std::string& qid) : e_(e), qid_(qid) {
  if (qid_.empty()) return;
  std::lock_guard<std::mutex> lk(e_->cancel_mu_);
  if (e_->running_[qid_]++ == 0) e_->running_since_[qid_] = now_ms();
}
Engine::RunningQuery::~RunningQuery() {
  if (qid_.empty()) return;
  std::lock_guard<std::mutex> lk(e_->cancel_mu_);
  auto it = e_->running_.find(qid_);
  if (it != e_->running_.end() && --it->second == 0) {
    e_->running_.erase(it);
    e_->running_since_.erase(qid_);
    e_->cancelled_.erase(qid_);  // a cancel that came too late dies with the query
  }
}

const uint16_t* Engine::ensure_hll_lut(DataSource& ds, int di, cudaStream_t st) {
  std::lock_guard<std::mutex> lk(lut_mu_);
  DimMeta& dm = ds.dims[di];
  if (dm.hll_lut && dm.hll_lut_card == dm.cardinality) return (const uint16_t*)dm.hll_lut;
  if (dm.hll_lut) { cudaFree(dm.hll_lut); dm.hll_lut = nullptr; }
  const Dictionary& d = *dm.dict;
  void *dbytes = nullptr, *doffs = nullptr, *lut = nullptr;
  SDO_CUDA(cudaMalloc(&dbytes, std::max<size_t>(d.bytes.size(), 16)));
  SDO_CUDA(cudaMalloc(&doffs, d.offsets.size() * 8));
  SDO_CUDA(cudaMalloc(&lut, std::max<size_t>((size_t)d.size() * 2, 16)));
  SDO_CUDA(cudaMemcpyAsync(dbytes, d.bytes.data(), d.bytes.size(), cudaMemcpyHostToDevice, st));
  SDO_CUDA(cudaMemcpyAsync(doffs, d.offsets.data(), d.offsets.size() * 8, cudaMemcpyHostToDevice, st));
  launch_hll_lut_strings((const char*)dbytes, (const int64_t*)doffs, d.size(), (uint16_t*)lut, st);
  SDO_CUDA(cudaStreamSynchronize(st));
  cudaFree(dbytes);
  cudaFree(doffs);
  dm.hll_lut = lut;
  dm.hll_lut_card = dm.cardinality;
  return (const uint16_t*)lut;
}

// The dimension's sketch column of one segment (uint16 per row), built on first use: the per-row analogue of
// Druid's ingest-time hyperUnique metric.  Depends on row values only, so dictionary re-encodes leave it valid.
const void* Engine::ensure_hll_column(DataSource& ds, int di, Segment& sg, cudaStream_t st) {
  {
    std::lock_guard<std::mutex> lk(lut_mu_);
    if ((int)sg.dim_hll.size() > di && sg.dim_hll[di]) return sg.dim_hll[di];
  }
  const DimMeta& dm = ds.dims[di];
  const uint16_t* lut = nullptr;
  int mode;
  if (dm.enc == DimEncoding::Direct) mode = dm.src_type == 1 ? HLL_DIRECT_INT : HLL_DIRECT_LONG;
  else { mode = HLL_LUT; lut = ensure_hll_lut(ds, di, st); }
  std::lock_guard<std::mutex> lk(lut_mu_);
  if ((int)sg.dim_hll.size() <= di) sg.dim_hll.resize(ds.dims.size(), nullptr);
  if (sg.dim_hll[di]) return sg.dim_hll[di];
  const size_t nb = round_up((size_t)sg.n_rows * 2 + 64, 256);
  void* col = nullptr;
  SDO_CUDA(cudaMalloc(&col, nb));
  SDO_CUDA(cudaMemsetAsync(col, 0, nb, st));
  launch_hll_column(sg.dim_data[di], sg.dim_width[di], mode, lut, sg.n_rows, (uint16_t*)col, st);
  SDO_CUDA(cudaStreamSynchronize(st));
  sg.dim_hll[di] = col;
  sg.device_bytes += nb;
  return col;
}

static uint64_t pow2_ceil(uint64_t v) {
  uint64_t p = 1;
  while (p < v) p <<= 1;
  return p;
}

// Upload the per-query blob (spans, coarse index, segment pointer table, bitsets, id remaps) and fill the parts of
// ScanParams every kernel over the plan's rows needs: spans, columns, compiled filter, dimensions, aggregators.
void Engine::prepare_params(QueryCtx* ctx, const Plan& plan, ScanState* st) {
  if (device_ < 0) throw NoDevice("this engine was created without a CUDA device (planning only); there is no CPU execution path");
  SDO_CUDA(cudaSetDevice(device_));
  cudaStream_t s = ctx->stream;
  DataSource& ds = *plan.ds;
  const size_t n_cols = plan.columns.size();
  const size_t n_segs = plan.seg_index.size();
  const size_t n_spans = plan.spans.size();

  // ---- per-query device blob: spans + sentinel | coarse index | segment pointer table | bitsets ----
  size_t off_spans = 0;
  size_t off_coarse = off_spans + (n_spans + 1) * sizeof(DevSpan);
  size_t off_segptr = round_up(off_coarse + plan.coarse.size() * 4, 16);
  size_t off = round_up(off_segptr + std::max<size_t>(1, n_segs * n_cols) * 8, 16);
  std::vector<size_t> off_bits(plan.leaves.size(), 0);
  for (size_t i = 0; i < plan.leaves.size(); i++)
    if (plan.leaves[i].leaf.kind == LEAF_BITSET) {
      off_bits[i] = off;
      off = round_up(off + plan.leaves[i].bitset.size() * 4, 16);
    }
  // id remaps of dimensions under an extraction function, in dev_dims order
  std::vector<const PlanDim*> scan_dims;
  for (auto& pd : plan.dims)
    if (!pd.is_time) scan_dims.push_back(&pd);
  std::vector<size_t> off_remap(scan_dims.size(), 0);
  for (size_t i = 0; i < scan_dims.size(); i++)
    if (!scan_dims[i]->remap.empty()) {
      off_remap[i] = off;
      off = round_up(off + scan_dims[i]->remap.size() * 4, 16);
    }
  // small dense tables ride along: their initial image (slot identities, zeroed registers) is part of the blob, so the
  // scan needs neither an allocation nor a fill kernel before it
  const uint64_t cells_pre = plan.n_keys * (uint64_t)plan.n_slots;
  const size_t tb_pre = round_up(std::max<uint64_t>(cells_pre, 1) * 8, 256);
  const size_t hb_pre = plan.n_hll ? (size_t)plan.n_keys * plan.n_hll * HLL_M : 0;
  size_t off_table = 0;
  const bool preinit = plan.strategy != STRAT_HASH && plan.n_keys <= 4096 && tb_pre + hb_pre <= 16 * 1024;
  if (preinit) {
    off_table = round_up(off, 256);
    off = off_table + tb_pre + hb_pre;
  }
  const size_t blob_bytes = off;
  ensure_pinned(&ctx->pinned_in, &ctx->pinned_in_bytes, blob_bytes);
  char* h = (char*)ctx->pinned_in;
  if (preinit) {
    unsigned long long* img = (unsigned long long*)(h + off_table);
    for (uint64_t x = 0; x < cells_pre; x++) img[x] = plan.slot_init[x % plan.n_slots];
    memset(h + off_table + cells_pre * 8, 0, tb_pre + hb_pre - cells_pre * 8);
  }
  memcpy(h + off_spans, plan.spans.data(), n_spans * sizeof(DevSpan));
  DevSpan sentinel{};
  sentinel.unit_begin = plan.n_units;
  memcpy(h + off_spans + n_spans * sizeof(DevSpan), &sentinel, sizeof(DevSpan));
  memcpy(h + off_coarse, plan.coarse.data(), plan.coarse.size() * 4);
  const void** sp = (const void**)(h + off_segptr);
  for (size_t si = 0; si < n_segs; si++) {
    const Segment& sg = *ds.segments[plan.seg_index[si]];
    for (size_t c = 0; c < n_cols; c++) {
      const PlanColumn& pc = plan.columns[c];
      const void* p = pc.is_metric ? sg.metric_data[pc.index] : sg.dim_data[pc.index];
      if (!p) throw Unsupported("segment " + format_iso_instant(sg.start_ms) + " has no data for column '" + (pc.is_metric ? ds.metrics[pc.index].name : ds.dims[pc.index].name) + "'");
      if (pc.is_hll) p = ensure_hll_column(ds, pc.index, *ds.segments[plan.seg_index[si]], s);
      else if (!pc.is_metric && sg.dim_width[pc.index] != pc.width) throw std::runtime_error("internal: segment id width differs from datasource width");
      sp[si * n_cols + c] = p;
    }
  }
  for (size_t i = 0; i < plan.leaves.size(); i++)
    if (plan.leaves[i].leaf.kind == LEAF_BITSET) memcpy(h + off_bits[i], plan.leaves[i].bitset.data(), plan.leaves[i].bitset.size() * 4);

  for (size_t i = 0; i < scan_dims.size(); i++)
    if (!scan_dims[i]->remap.empty()) memcpy(h + off_remap[i], scan_dims[i]->remap.data(), scan_dims[i]->remap.size() * 4);

  SDO_CUDA(cudaEventRecord(ctx->ev[0], s));
  st->blob = take_scratch(ctx, blob_bytes);
  if (!st->blob) SDO_CUDA(cudaMallocAsync(&st->blob, blob_bytes, s));
  SDO_CUDA(cudaMemcpyAsync(st->blob, h, blob_bytes, cudaMemcpyHostToDevice, s));
  if (preinit) {
    st->table_preinit = true;
    st->table = (unsigned long long*)((char*)st->blob + off_table);
  }
  st->h2d_bytes += blob_bytes;

  ScanParams& p = st->params;
  memset(&p, 0, sizeof p);
  p.ticket_shift = 10;  // refined per shape below; the hash count pass runs with this
  char* d = (char*)st->blob;
  p.spans = (const DevSpan*)(d + off_spans);
  p.n_spans = (uint32_t)n_spans;
  p.n_units = plan.n_units;
  p.units_per_chunk = 2;
  p.n_cols = (uint32_t)n_cols;
  p.segptr = (const void* const*)(d + off_segptr);
  p.n_leaves = (int)plan.leaves.size();
  for (size_t i = 0; i < plan.leaves.size(); i++) {
    p.leaves[i] = plan.leaves[i].leaf;
    if (p.leaves[i].kind == LEAF_BITSET) p.leaves[i].bits = (const uint32_t*)(d + off_bits[i]);
  }
  p.n_ops = (int)plan.op_code.size();
  for (size_t i = 0; i < plan.op_code.size(); i++) { p.op_code[i] = plan.op_code[i]; p.op_arg[i] = plan.op_arg[i]; }
  p.n_dims = (int)plan.dev_dims.size();
  for (size_t i = 0; i < plan.dev_dims.size(); i++) {
    p.dims[i] = plan.dev_dims[i];
    if (!scan_dims[i]->remap.empty()) p.dims[i].remap = (const uint32_t*)(d + off_remap[i]);
  }
  p.n_aggs = (int)plan.dev_aggs.size();
  for (size_t i = 0; i < plan.dev_aggs.size(); i++) {
    p.aggs[i] = plan.dev_aggs[i];
    if (p.aggs[i].op == AGG_HLL && p.aggs[i].hll_mode == HLL_LUT) p.aggs[i].hll_lut = ensure_hll_lut(ds, plan.hll_ds_dim[i], s);
  }
  p.n_exprs = (int)plan.exprs.size();
  for (size_t i = 0; i < plan.exprs.size(); i++) p.exprs[i] = plan.exprs[i];
  p.n_slots = plan.n_slots;
  p.n_hll = plan.n_hll;
  p.n_keys = plan.n_keys;
  for (int i = 0; i < plan.n_slots; i++) { p.slot_op[i] = plan.slot_op[i]; p.slot_init[i] = plan.slot_init[i]; }
  p.counters = ctx->counters;  // zeroed when the context was created and by free_state after every query
  if (plan.rows_scanned >= (1ll << 32)) throw Unsupported("a single launch scans at most 2^32 rows per GPU");
}

void Engine::run_scan(QueryCtx* ctx, const Plan& plan, ScanState* st) {
  prepare_params(ctx, plan, st);
  cudaStream_t s = ctx->stream;
  ScanParams& p = st->params;
  int strat = plan.strategy;
  st->strategy = strat;
  const uint64_t cells = plan.n_keys * (uint64_t)plan.n_slots;

  // rows that pass the filter, counted exactly by a scan without keys or aggregates (cached per plan by the callers)
  auto count_passing = [&]() -> uint64_t {
      ScanParams c = p;
    c.n_dims = 0;
    c.n_aggs = 0;
    c.n_keys = 1;
    c.n_slots = 1;
    c.n_hll = 0;
    c.ignore_key_base = 1;
    size_t smem = scan_smem_bytes(STRAT_LANE, 1, 1);
    int grid = scan_grid_size(STRAT_LANE, smem, sm_count_);
    unsigned long long* part = nullptr;
    SDO_CUDA(cudaMallocAsync((void**)&part, (size_t)(grid + 1) * 8, s));
    c.table = part;
    launch_scan(STRAT_LANE, c, grid, smem, s);
    launch_reduce_partials(part, grid, 1, 1, c, part + grid, s);
    unsigned long long cnt = 0;
    SDO_CUDA(cudaMemcpyAsync(&cnt, part + grid, 8, cudaMemcpyDeviceToHost, s));
    SDO_CUDA(cudaStreamSynchronize(s));
    SDO_CUDA(cudaFreeAsync(part, s));
    SDO_CUDA(cudaMemsetAsync(ctx->counters, 0, COUNTER_BYTES, s));
    st->launches += 2;
    st->d2h_bytes += 8;
    return cnt;
  };
  if (strat == STRAT_EMIT) {
    // No table: the surviving rows come out as groups of one row each, n_slots - 1 words per row, split by owner when the
    // caller (a sharded query) names the owners.  A region holds its fair share of the rows and a quarter more; keys
    // skewed beyond that raise the overflow flag and the scan is repeated with regions that hold everything.
    uint64_t cap_all = (uint64_t)std::max<int64_t>(plan.rows_scanned, 1);
    if (p.n_ops > 0) {
      // a filter can shrink the output a hundredfold: size it from the exact count of passing rows (first run of the plan)
      int64_t known = plan.passing_rows.load();
      if (known < 0) { known = (int64_t)count_passing(); plan.passing_rows.store(known); }
      cap_all = (uint64_t)std::max<int64_t>(known, 1);
    }
    const int owners = std::max(1, st->emit_owners);
    const int nsr = plan.n_slots - 1;
    st->emit_cap = owners > 1 && !st->emit_full_regions ? std::min<uint64_t>(cap_all, cap_all / owners + cap_all / (4 * owners) + 65536) : cap_all;
    SDO_CUDA(cudaMallocAsync((void**)&st->rows_keys, (uint64_t)owners * st->emit_cap * 8, s));
    SDO_CUDA(cudaMallocAsync((void**)&st->rows_slots, std::max<uint64_t>((uint64_t)owners * st->emit_cap * nsr, 1) * 8, s));
    st->emit_cursors_dev = (unsigned long long*)take_scratch(ctx, 64 * 8);
    if (!st->emit_cursors_dev) throw std::runtime_error("internal: context scratch exhausted");
    SDO_CUDA(cudaMemsetAsync(st->emit_cursors_dev, 0, 64 * 8, s));
    p.emit_keys = st->rows_keys;
    p.emit_slots = st->rows_slots;
    p.emit_cursors = st->emit_cursors_dev;
    p.emit_cap = st->emit_cap;
    p.emit_owners = owners;
    p.emit_interleave = st->emit_interleave;
    st->table_rows = 0;
  } else if (strat == STRAT_HASH) {
    // size the table from an exact count of passing rows when a filter can shrink it
    uint64_t rows_upper = (uint64_t)plan.rows_scanned;
    const int64_t known = plan.passing_rows.load();
    if (p.n_ops > 0 && known >= 0) {
      rows_upper = (uint64_t)known;  // counted by an earlier run of this cached plan
    } else if (p.n_ops > 0) {
      const unsigned long long cnt = count_passing();
      rows_upper = cnt;
      plan.passing_rows.store((int64_t)cnt);
    }
    uint64_t groups_upper = std::min<uint64_t>(plan.n_keys, rows_upper);
    uint64_t cap = pow2_ceil(std::max<uint64_t>(1024, groups_upper + groups_upper / 2));
    st->table_rows = cap;
    const uint32_t stride = (uint32_t)pow2_ceil((uint64_t)plan.n_slots + 1);  // key + slots, sector-aligned buckets
    SDO_CUDA(cudaMallocAsync((void**)&st->table, cap * (uint64_t)stride * 8, s));
    launch_fill_hash(st->table, cap, stride, plan.n_slots, p, s);
    st->launches += 1;
    p.hkeys = st->table;
    p.hmask = cap - 1;
    p.hstride = stride;
    p.table = st->table;
  } else if (strat != STRAT_EMIT) {
    st->table_rows = plan.n_keys;
    // dense table and its HLL registers share one block: a partial is then a single contiguous blob
    const size_t tb = round_up(std::max<uint64_t>(cells, 1) * 8, 256);
    const size_t hb = plan.n_hll ? (size_t)plan.n_keys * plan.n_hll * HLL_M : 0;
    if (hb > opt_.hll_max_bytes) throw Unsupported("cardinality aggregation needs " + std::to_string(hb) + " bytes of registers");
    if (!st->table_preinit) SDO_CUDA(cudaMallocAsync((void**)&st->table, tb + hb, s));
    st->dense_blob_bytes = tb + hb;
    st->dense_table_bytes = tb;
    if (hb) {
      st->hll = (uint8_t*)st->table + tb;
      st->hll_in_table = true;
      if (!st->table_preinit) SDO_CUDA(cudaMemsetAsync(st->hll, 0, hb, s));
      p.hll = st->hll;
    }
    if (strat != STRAT_LANE) {
      if (!st->table_preinit) {
        launch_fill_slots(st->table, plan.n_keys, plan.n_slots, nullptr, p, s);
        st->launches += 1;
      }
      p.table = st->table;
    }
  }
  if (plan.n_hll && !st->hll_in_table) {
    size_t hb = (size_t)st->table_rows * plan.n_hll * HLL_M;
    if (hb > opt_.hll_max_bytes) throw Unsupported("cardinality aggregation needs " + std::to_string(hb) + " bytes of registers");
    SDO_CUDA(cudaMallocAsync((void**)&st->hll, hb, s));
    SDO_CUDA(cudaMemsetAsync(st->hll, 0, hb, s));
    p.hll = st->hll;
  }

  if (strat == STRAT_BLOCK) {
    // replicas of the shared table while they stay within 24 KB; an odd stride spreads the replicas over the banks
    const uint32_t stride = (uint32_t)cells | 1u;
    uint32_t rep = 32;
    while (rep > 1 && (uint64_t)rep * stride * 8 > 24 * 1024) rep >>= 1;
    if (rep > 1) { p.rep_mask = rep - 1; p.rep_stride = stride; }
  }
  size_t smem = scan_smem_bytes(strat, plan.n_keys, plan.n_slots, p.rep_mask + 1, p.rep_stride);
  p.hll_mode = HLL_IN_GLOBAL;
  if (plan.n_hll && strat != STRAT_HASH) {
    // keep the sketches in shared memory when they fit next to the aggregate table
    const std::string& force = plan.spec.gpu_hll_place;
    const uint64_t all_bytes = plan.n_keys * (uint64_t)plan.n_hll * HLL_M;
    // Sketches per block cost a warm-up of every register in every block, and shared memory costs occupancy: all groups
    // go to shared memory only while 4 blocks per SM still fit (measured on SF10: 84 month sketches 0.28 ms shared vs
    // 0.47 ms global; 175 sketches 1.00 ms shared vs 0.65 ms global).  One word per register when that fits too.
    const uint64_t budget = (force == "shared" ? 200 : 56) * 1024;
    // No grouping dimension (timeseries): the key is constant over a span, so one cached sketch per block is enough.
    if (force == "global") {
    } else if (plan.dev_dims.empty() && force.empty() && (size_t)plan.n_hll * HLL_M * 4 * (SCAN_THREADS / 32) + smem <= 200 * 1024) { p.hll_mode = HLL_IN_SMEM_SPAN; p.hll_wide = 1; }
    else if (all_bytes * 4 + smem <= budget) { p.hll_mode = HLL_IN_SMEM_ALL; p.hll_wide = 1; }
    else if (all_bytes + smem <= budget) p.hll_mode = HLL_IN_SMEM_ALL;
    else if (plan.dev_dims.empty() && (size_t)plan.n_hll * HLL_M * 4 * (SCAN_THREADS / 32) + smem <= 200 * 1024) { p.hll_mode = HLL_IN_SMEM_SPAN; p.hll_wide = 1; }
    if (p.hll_mode != HLL_IN_GLOBAL) smem = round_up(smem, 16) + scan_hll_smem_bytes(p.hll_mode, p.hll_wide != 0, plan.n_keys, plan.n_hll);
  }
  // Rows per ticket (measured on SF10, ms at 512 / 1024 / 2048 / 4096 rows): month timeseries + HLL 0.44 / 0.25 / 0.18 /
  // 0.17 (a warp merges its cached sketch into HBM per ticket), TopN with a selective filter 0.127 / 0.094 / 0.088 /
  // 0.134, hash group-by + limit 0.064 / 0.079 / 0.133 / 0.237, 15 M-group direct table 1.77 / 1.89 / 2.26 / 2.53.
  // On SF100 (600 M rows) the filtered TopN runs 0.552 ms at 2048 rows per ticket and 0.539 at 4096: long scans amortise the
  // ticket draw, short ones need the finer tail.
  p.ticket_shift = p.hll_mode == HLL_IN_SMEM_SPAN ? 12 : (strat == STRAT_HASH || strat == STRAT_GLOBAL) ? 9
                   : (p.n_ops > 0 || strat == STRAT_SPAN ? (plan.rows_scanned >= (1ll << 28) ? 12 : 11) : 10);
  if (plan.spec.gpu_ticket_rows) p.ticket_shift = plan.spec.gpu_ticket_rows == 512 ? 9 : plan.spec.gpu_ticket_rows == 1024 ? 10 : plan.spec.gpu_ticket_rows == 2048 ? 11 : 12;
  // run-time specialised kernel for this query shape (large scans, or when the query asks for it)
  const JitKernel* jk = nullptr;
  {
    static const char* env = getenv("SDO_JIT");
    const bool env_off = env && env[0] == '0', env_on = env && env[0] == '1';
    bool want = plan.spec.gpu_jit > 0 || (plan.spec.gpu_jit < 0 && !env_off && (env_on || plan.rows_scanned >= opt_.jit_min_rows));
    if (want) {
      double ms = 0;
      // only a query that asks for the specialised kernel waits for the compiler; the others run ahead-of-time code
      // until the background compilation of their shape has finished
      static const char* sync_env = getenv("SDO_JIT_SYNC");
      jk = jit_get(strat, p, &ms, plan.spec.gpu_jit > 0 || (sync_env && sync_env[0] == '1'));
      st->jit_compile_ms = ms;
    }
  }
  st->jit = jk != nullptr;
  int grid = 0;
  if (jk) grid = jit_blocks_per_sm(jk, smem) * sm_count_;
  if (grid <= 0) { jk = nullptr; st->jit = false; grid = scan_grid_size(strat, smem, sm_count_); }
  // Other scans in flight on this engine: a persistent grid at full occupancy would make every later kernel wait for this
  // one's last block, so a short query queued behind a long one takes the long one's time.  Each scan takes its share of the
  // resident blocks instead (never less than one block per SM) and the kernels run side by side.
  if (!st->counted) { st->counted = true; st->others_at_launch = scans_in_flight_.fetch_add(1); }
  if (st->others_at_launch > 0 && grid > sm_count_) {
    const int per_sm = grid / sm_count_;
    grid = std::max(1, per_sm / std::min(per_sm, st->others_at_launch + 1)) * sm_count_;
  }
  // a block's 8 warps each draw quarter-unit tickets: no more blocks than there is work for
  const uint32_t n_chunks = (uint32_t)((((uint64_t)plan.n_units << (12 - p.ticket_shift)) + 7) / 8);
  grid = (int)std::max<uint32_t>(1, std::min<uint32_t>((uint32_t)grid, n_chunks));
  if (strat == STRAT_LANE) {
    st->partials = (unsigned long long*)take_scratch(ctx, (size_t)grid * std::max<uint64_t>(cells, 1) * 8);
    if (!st->partials) SDO_CUDA(cudaMallocAsync((void**)&st->partials, (size_t)grid * std::max<uint64_t>(cells, 1) * 8, s));
    p.table = st->partials;
  }
  SDO_CUDA(cudaEventRecord(ctx->ev[1], s));
  if (!jk || !jit_launch(jk, p, grid, smem, s)) {
    st->jit = false;
    launch_scan(strat, p, grid, smem, s);
  }
  SDO_CUDA(cudaEventRecord(ctx->ev[2], s));
  st->launches += 1;
  if (strat == STRAT_EMIT) {
    const int owners = std::max(1, st->emit_owners);
    unsigned long long cur[64];
    unsigned int flags[4];
    SDO_CUDA(cudaMemcpyAsync(cur, st->emit_cursors_dev, (size_t)owners * 8, cudaMemcpyDeviceToHost, s));
    SDO_CUDA(cudaMemcpyAsync(flags, ctx->counters, 16, cudaMemcpyDeviceToHost, s));
    SDO_CUDA(cudaStreamSynchronize(s));
    st->d2h_bytes += (size_t)owners * 8 + 16;
    if (flags[1]) {
      if (owners == 1 || st->emit_full_regions) throw std::runtime_error("internal: more rows emitted than the spans hold");
      // skewed owners: once more with regions that hold every row
      SDO_CUDA(cudaFreeAsync(st->rows_keys, s));
      SDO_CUDA(cudaFreeAsync(st->rows_slots, s));
      st->rows_keys = st->rows_slots = nullptr;
      if (st->blob && !((char*)st->blob >= ctx->scratch && (char*)st->blob < ctx->scratch + ctx->scratch_bytes)) SDO_CUDA(cudaFreeAsync(st->blob, s));
      st->blob = nullptr;
      ctx->scratch_used = 0;
      SDO_CUDA(cudaMemsetAsync(ctx->counters, 0, COUNTER_BYTES, s));
      st->emit_full_regions = true;
      return run_scan(ctx, plan, st);
    }
    st->emit_counts.assign(cur, cur + owners);
    st->n_rows = 0;
    for (int o = 0; o < owners; o++) st->n_rows += cur[o];
    st->emit_regions = owners > 1;
    st->rows_valid = owners == 1;
  }
  if (strat == STRAT_LANE) {
    launch_reduce_partials(st->partials, grid, (uint32_t)cells, plan.n_slots, p, st->table, s);
    st->launches += 1;
    p.table = st->table;
  }
}

// The group table of a finished scan as the top-k select reads it: compacted rows, hash buckets or a direct table.
static SelView select_view(const ScanState& st, int ns, int slot, int is_f64, int desc) {
  SelView v{};
  v.is_f64 = is_f64;
  v.desc = desc;
  if (st.rows_valid) {
    v.kind = SEL_ROWS; v.base = st.rows_slots; v.keys = st.rows_keys; v.n = st.n_rows; v.stride = (uint32_t)ns; v.slot_off = (uint32_t)slot; v.slot0_off = 0;
  } else if (st.strategy == STRAT_HASH) {
    v.kind = SEL_HASH; v.base = st.table; v.n = st.table_rows; v.stride = st.params.hstride; v.slot_off = 1u + (uint32_t)slot; v.slot0_off = 1;
  } else {
    v.kind = SEL_DENSE; v.base = st.table; v.n = st.table_rows; v.stride = (uint32_t)ns; v.slot_off = (uint32_t)slot; v.slot0_off = 0;
    v.key_off = st.dense_key_off;
    v.key_mul = st.dense_key_mul;
  }
  return v;
}

// Gather the occupied groups of a large dense table or a hash table into row arrays.
void Engine::compact_rows(QueryCtx* ctx, const Plan& plan, ScanState* st) {
  cudaStream_t s = ctx->stream;
  const int ns = plan.n_slots;
  unsigned long long* counter = (unsigned long long*)(ctx->counters + 4);
  unsigned long long* src = nullptr;
  // upper bound on rows: table rows (dense) or capacity (hash); allocate exactly after a count
  // pass would cost another launch, so size by min(table rows, rows scanned)
  // (a table merged with other ranks' partials can hold more groups than this rank scanned rows)
  uint64_t cap_rows = st->merged ? st->table_rows : std::min<uint64_t>(st->table_rows, (uint64_t)std::max<int64_t>(plan.rows_scanned, 1));
  SDO_CUDA(cudaMallocAsync((void**)&st->rows_keys, std::max<uint64_t>(cap_rows, 1) * 8, s));
  SDO_CUDA(cudaMallocAsync((void**)&st->rows_slots, std::max<uint64_t>(cap_rows, 1) * ns * 8, s));
  if (plan.n_hll) SDO_CUDA(cudaMallocAsync((void**)&src, std::max<uint64_t>(cap_rows, 1) * 8, s));
  SDO_CUDA(cudaMemsetAsync(counter, 0, 8, s));
  if (st->strategy == STRAT_HASH) launch_compact_hash(st->table, st->table_rows, ns, st->params.hstride, st->rows_keys, st->rows_slots, src, counter, s);
  else launch_compact_dense(st->table, st->table_rows, ns, st->rows_keys, st->rows_slots, src, counter, s, st->dense_key_off, st->dense_key_mul);
  st->launches += 1;
  struct { unsigned int c[4]; unsigned long long n; } host;
  SDO_CUDA(cudaMemcpyAsync(&host, ctx->counters, 24, cudaMemcpyDeviceToHost, s));
  SDO_CUDA(cudaStreamSynchronize(s));
  st->d2h_bytes += 24;
  if (host.c[1]) throw std::runtime_error("internal: hash table overflow");
  st->n_rows = host.n;
  if (plan.n_hll) {
    SDO_CUDA(cudaMallocAsync((void**)&st->rows_hll, std::max<uint64_t>(st->n_rows, 1) * plan.n_hll * HLL_M, s));
    launch_gather_hll(st->hll, src, st->n_rows, plan.n_hll, st->rows_hll, s);
    st->launches += 1;
    SDO_CUDA(cudaFreeAsync(src, s));
  }
  st->rows_valid = true;
}

void Engine::free_state(QueryCtx* ctx, ScanState* st) {
  if (st->counted) { st->counted = false; scans_in_flight_.fetch_sub(1); }
  if (device_ < 0) return;
  cudaStream_t s = ctx->stream;
  void* ptrs[] = {st->blob, st->table, st->hkeys, st->hll_in_table ? nullptr : (void*)st->hll, st->partials, st->rows_keys, st->rows_slots, st->rows_hll, st->hll_stats};
  for (void* p : ptrs)
    if (p && !((char*)p >= ctx->scratch && (char*)p < ctx->scratch + ctx->scratch_bytes)) cudaFreeAsync(p, s);
  ctx->scratch_used = 0;
  // the next query of this context finds its ticket counters zeroed (off its critical path)
  cudaMemsetAsync(ctx->counters, 0, COUNTER_BYTES, s);
  st->table_preinit = false;
  st->blob = nullptr; st->table = nullptr; st->hkeys = nullptr; st->hll = nullptr; st->partials = nullptr;
  st->rows_keys = nullptr; st->rows_slots = nullptr; st->rows_hll = nullptr; st->hll_stats = nullptr;
}

namespace {

struct HostRows {
  uint64_t n = 0;
  int ns = 1;
  std::vector<uint64_t> keys, slots;
  uint8_t* hll = nullptr;          // [sketch][512] in a pooled pinned buffer, handed to the result
  size_t hll_bytes = 0, hll_cap = 0;
  PinnedPool* pool = nullptr;
  ~HostRows() { if (hll && pool) pool->put(hll, hll_cap); }
  std::vector<uint64_t> hll_set;   // per row: index of its first sketch in hll / hll_stats (sketches of a row are adjacent)
  std::vector<double> hll_stats;   // per sketch: sum of 2^-register, zero registers
};

inline double ord_to_f64(int64_t raw) {
  if (raw == INT64_MAX) return INFINITY;   // empty min
  if (raw == INT64_MIN) return -INFINITY;  // empty max
  int64_t bits = raw ^ ((raw >> 63) & 0x7FFFFFFFFFFFFFFFll);
  double d;
  memcpy(&d, &bits, 8);
  return d;
}

struct Assembler {
  const Plan& plan;
  const HostRows& rows;
  std::vector<std::vector<int64_t>> agg_i;  // per agg (int outputs)
  std::vector<std::vector<double>> agg_f;   // per agg (float outputs)
  std::vector<std::vector<double>> post;    // per post-agg

  double agg_value(size_t a, uint64_t r) const { return plan.aggs[a].out_float ? agg_f[a][r] : (double)agg_i[a][r]; }
  int find_agg(const std::string& n) const {
    for (size_t i = 0; i < plan.aggs.size(); i++)
      if (plan.aggs[i].name == n) return (int)i;
    return -1;
  }
  int find_post(const std::string& n) const {
    for (size_t i = 0; i < plan.spec.post_aggregations.size(); i++)
      if (plan.spec.post_aggregations[i].name == n) return (int)i;
    return -1;
  }
  bool named_value(const std::string& n, uint64_t r, double* v, size_t post_limit) const {
    int a = find_agg(n);
    if (a >= 0) { *v = agg_value((size_t)a, r); return true; }
    int p = find_post(n);
    if (p >= 0 && (size_t)p < post_limit) { *v = post[(size_t)p][r]; return true; }
    return false;
  }
  // PostAggregationSpec semantics: DruidQuerySpec.scala:379-430 (Druid arithmetic: x / 0 = 0)
  double eval_post(const PostAggSpec& ps, uint64_t r, size_t post_limit) const {
    switch (ps.kind) {
      case PostAggSpec::Constant: return ps.value;
      case PostAggSpec::FieldAccess: case PostAggSpec::HyperUniqueCardinality: {
        double v;
        if (!named_value(ps.field, r, &v, post_limit)) throw Invalid("postAggregation refers to unknown field '" + ps.field + "'");
        return v;
      }
      case PostAggSpec::Arithmetic: {
        double acc = eval_post(ps.fields[0], r, post_limit);
        for (size_t i = 1; i < ps.fields.size(); i++) {
          double x = eval_post(ps.fields[i], r, post_limit);
          if (ps.fn == "+") acc += x;
          else if (ps.fn == "-") acc -= x;
          else if (ps.fn == "*") acc *= x;
          else if (ps.fn == "/") acc = x == 0.0 ? 0.0 : acc / x;
          else acc = acc / x;
        }
        return acc;
      }
    }
    return 0;
  }
  bool eval_having(const HavingNode& h, uint64_t r) const {
    switch (h.kind) {
      case HavingNode::And: for (auto& c : h.specs) if (!eval_having(c, r)) return false; return true;
      case HavingNode::Or: for (auto& c : h.specs) if (eval_having(c, r)) return true; return false;
      case HavingNode::Not: return !eval_having(h.specs.at(0), r);
      default: break;
    }
    double v;
    if (!named_value(h.aggregation, r, &v, post.size())) throw Invalid("having refers to unknown aggregation '" + h.aggregation + "'");
    if (h.kind == HavingNode::GreaterThan) return v > h.value;
    if (h.kind == HavingNode::LessThan) return v < h.value;
    return v == h.value;
  }
};

}  // namespace

Result* Engine::finalize(QueryCtx* ctx, const Plan& plan, ScanState* st, double t_start_ms) {
  cudaStream_t s = ctx ? ctx->stream : nullptr;
  const int ns = plan.n_slots;
  HostRows rows;
  rows.ns = ns;
  rows.pool = &pinned_pool_;
  bool device_topk = false;
  uint64_t groups_total = 0;  // groups in the table when only the top k came back
  float ms_h2d = 0, ms_scan = 0, ms_dev = 0;

  if (!plan.always_false) {
    const bool dense = st->strategy != STRAT_HASH && !st->rows_valid;
    if (plan.zero_fill && !(dense && st->table_rows <= 65536)) throw Unsupported("skipEmptyBuckets = false with more than 65536 buckets");
    if (dense && st->table_rows <= 65536 && (uint64_t)st->table_rows * std::max(1, plan.n_hll) * HLL_M <= (64u << 20)) {
      // small dense table: bring it back whole and drop empty groups on the host
      const size_t tb = (size_t)st->table_rows * ns * 8;
      const size_t n_sets = plan.n_hll ? (size_t)st->table_rows * plan.n_hll : 0;
      const size_t hb = n_sets * HLL_M, sb = n_sets * 16;
      ensure_pinned(&ctx->pinned_out, &ctx->pinned_out_bytes, tb + sb);
      SDO_CUDA(cudaEventRecord(ctx->ev[3], s));
      SDO_CUDA(cudaMemcpyAsync(ctx->pinned_out, st->table, tb, cudaMemcpyDeviceToHost, s));
      if (hb) {
        // estimator inputs are computed on the device; the registers themselves go straight into the result's backing
        SDO_CUDA(cudaMallocAsync((void**)&st->hll_stats, sb, s));
        launch_hll_stats(st->hll, n_sets, st->hll_stats, s);
        st->launches += 1;
        SDO_CUDA(cudaMemcpyAsync((char*)ctx->pinned_out + tb, st->hll_stats, sb, cudaMemcpyDeviceToHost, s));
        rows.hll = (uint8_t*)pinned_pool_.get(hb, &rows.hll_cap);
        rows.hll_bytes = hb;
        SDO_CUDA(cudaMemcpyAsync(rows.hll, st->hll, hb, cudaMemcpyDeviceToHost, s));
      }
      SDO_CUDA(cudaEventRecord(ctx->ev[4], s));
      SDO_CUDA(cudaStreamSynchronize(s));
      st->d2h_bytes += tb + hb + sb;
      const uint64_t* t = (const uint64_t*)ctx->pinned_out;
      if (hb) {
        const double* hs = (const double*)((const char*)ctx->pinned_out + tb);
        rows.hll_stats.assign(hs, hs + n_sets * 2);
      }
      for (uint64_t k = 0; k < st->table_rows; k++) {
        if (t[k * ns] == 0 && !plan.zero_fill) continue;
        rows.keys.push_back(k * st->dense_key_mul + st->dense_key_off);
        rows.slots.insert(rows.slots.end(), t + k * ns, t + (k + 1) * ns);
        if (hb) rows.hll_set.push_back(k * plan.n_hll);
      }
      rows.n = rows.keys.size();
    } else {
      // device top-k when only a few rows of a large result are wanted
      int topk_slot = -1, topk_f64 = 0, topk_desc = 1;
      int64_t k_want = 0;
      if (!topk_request(plan, &topk_slot, &topk_f64, &topk_desc, &k_want)) topk_slot = -1;
      // First choice: radix select over the table as it stands (compacted rows, direct table or hash table) -- no
      // compaction pass, a tournament over the few candidates only.  Ties beyond the candidate list fall through.
      if (topk_slot >= 0 && (st->rows_valid ? st->n_rows : st->table_rows) > 65536) {
        const int k = (int)k_want;
        SDO_CUDA(cudaEventRecord(ctx->ev[3], s));
        unsigned long long *gk = nullptr, *gs = nullptr;
        void* scratch = nullptr;
        SDO_CUDA(cudaMallocAsync((void**)&gk, (size_t)k * 8, s));
        SDO_CUDA(cudaMallocAsync((void**)&gs, (size_t)k * ns * 8, s));
        SDO_CUDA(cudaMallocAsync(&scratch, select_scratch_bytes(k), s));
        launch_select_topk(select_view(*st, ns, topk_slot, topk_f64, topk_desc), ns, k, scratch, gk, gs, s);
        st->launches += 16;
        SelState sel;
        rows.keys.assign((size_t)k, 0);
        rows.slots.assign((size_t)k * ns, 0);
        SDO_CUDA(cudaMemcpyAsync(rows.keys.data(), gk, (size_t)k * 8, cudaMemcpyDeviceToHost, s));
        SDO_CUDA(cudaMemcpyAsync(rows.slots.data(), gs, (size_t)k * ns * 8, cudaMemcpyDeviceToHost, s));
        SDO_CUDA(cudaMemcpyAsync(&sel, scratch, sizeof sel, cudaMemcpyDeviceToHost, s));
        SDO_CUDA(cudaEventRecord(ctx->ev[4], s));
        SDO_CUDA(cudaStreamSynchronize(s));
        st->d2h_bytes += (size_t)k * (ns + 1) * 8 + sizeof sel;
        for (void* p : {(void*)gk, (void*)gs, scratch}) SDO_CUDA(cudaFreeAsync(p, s));
        if (!sel.overflow) {
          size_t w = 0;  // fewer than k groups: the rest is padding
          for (size_t i = 0; i < (size_t)k; i++)
            if (rows.keys[i] != ~0ull) {
              rows.keys[w] = rows.keys[i];
              memmove(&rows.slots[w * ns], &rows.slots[i * ns], (size_t)ns * 8);
              w++;
            }
          rows.keys.resize(w);
          rows.slots.resize(w * ns);
          rows.n = w;
          groups_total = sel.live;
          device_topk = true;
        }
      }
      if (!device_topk) {
      if (!st->rows_valid) compact_rows(ctx, plan, st);
      uint64_t n = st->n_rows;
      if (n <= 65536) topk_slot = -1;
      // A large result that needs no host-side work (no post-aggregation, having, ordering or sketches)
      // is turned into its columns on the device and copied straight into pinned result buffers.
      {
        const QuerySpec& q = plan.spec;
        const bool ordered = q.type == QuerySpec::TopN || (q.type == QuerySpec::GroupBy && q.limit.present);
        if (n > 65536 && !ordered && !q.has_having && q.post_aggregations.empty() && plan.n_hll == 0) return finalize_columnar(ctx, plan, st, t_start_ms);
      }
      SDO_CUDA(cudaEventRecord(ctx->ev[3], s));
      if (topk_slot >= 0 && !plan.n_hll) {
        int k = (int)k_want;
        size_t se = topk_scratch_elems(n, k);
        unsigned long long *sa = nullptr, *sb = nullptr, *oi = nullptr, *gk = nullptr, *gs = nullptr;
        SDO_CUDA(cudaMallocAsync((void**)&sa, se * 8, s));
        SDO_CUDA(cudaMallocAsync((void**)&sb, se * 8, s));
        SDO_CUDA(cudaMallocAsync((void**)&oi, (size_t)k * 8, s));
        SDO_CUDA(cudaMallocAsync((void**)&gk, (size_t)k * 8, s));
        SDO_CUDA(cudaMallocAsync((void**)&gs, (size_t)k * ns * 8, s));
        launch_topk(st->rows_keys, st->rows_slots, n, ns, topk_slot, topk_f64, topk_desc, k, sa, sb, oi, s);
        launch_gather_rows(st->rows_keys, st->rows_slots, oi, (uint64_t)k, ns, gk, gs, s);
        st->launches += 2;
        rows.keys.assign((size_t)k, 0);
        rows.slots.assign((size_t)k * ns, 0);
        SDO_CUDA(cudaMemcpyAsync(rows.keys.data(), gk, (size_t)k * 8, cudaMemcpyDeviceToHost, s));
        SDO_CUDA(cudaMemcpyAsync(rows.slots.data(), gs, (size_t)k * ns * 8, cudaMemcpyDeviceToHost, s));
        SDO_CUDA(cudaEventRecord(ctx->ev[4], s));
        SDO_CUDA(cudaStreamSynchronize(s));
        st->d2h_bytes += (size_t)k * (ns + 1) * 8;
        for (void* p : {(void*)sa, (void*)sb, (void*)oi, (void*)gk, (void*)gs}) SDO_CUDA(cudaFreeAsync(p, s));
        // drop padding (fewer than k groups cannot happen here since n > 65536 >= k, kept for safety)
        size_t w = 0;
        for (size_t i = 0; i < (size_t)k; i++)
          if (rows.keys[i] != ~0ull) {
            rows.keys[w] = rows.keys[i];
            memmove(&rows.slots[w * ns], &rows.slots[i * ns], (size_t)ns * 8);
            w++;
          }
        rows.keys.resize(w);
        rows.slots.resize(w * ns);
        rows.n = w;
        groups_total = n;
        device_topk = true;
      } else {
        rows.keys.assign((size_t)n, 0);
        rows.slots.assign((size_t)n * ns, 0);
        if (n) {
          SDO_CUDA(cudaMemcpyAsync(rows.keys.data(), st->rows_keys, (size_t)n * 8, cudaMemcpyDeviceToHost, s));
          SDO_CUDA(cudaMemcpyAsync(rows.slots.data(), st->rows_slots, (size_t)n * ns * 8, cudaMemcpyDeviceToHost, s));
          if (plan.n_hll) {
            const size_t n_sets = (size_t)n * plan.n_hll;
            rows.hll_bytes = n_sets * HLL_M;
            rows.hll = (uint8_t*)pinned_pool_.get(rows.hll_bytes, &rows.hll_cap);
            rows.hll_stats.assign(n_sets * 2, 0.0);
            SDO_CUDA(cudaMallocAsync((void**)&st->hll_stats, n_sets * 16, s));
            launch_hll_stats(st->rows_hll, n_sets, st->hll_stats, s);
            st->launches += 1;
            SDO_CUDA(cudaMemcpyAsync(rows.hll_stats.data(), st->hll_stats, n_sets * 16, cudaMemcpyDeviceToHost, s));
            SDO_CUDA(cudaMemcpyAsync(rows.hll, st->rows_hll, rows.hll_bytes, cudaMemcpyDeviceToHost, s));
            rows.hll_set.resize(n);
            for (uint64_t r = 0; r < n; r++) rows.hll_set[r] = r * plan.n_hll;
          }
        }
        SDO_CUDA(cudaEventRecord(ctx->ev[4], s));
        SDO_CUDA(cudaStreamSynchronize(s));
        st->d2h_bytes += (size_t)n * (ns + 1) * 8 + rows.hll_bytes + rows.hll_stats.size() * 8;
        rows.n = n;
        if (!plan.n_hll) {
          // top-k candidates gathered from several ranks are padded with the empty key where a share held fewer than K groups
          size_t w = 0;
          for (size_t i = 0; i < (size_t)n; i++)
            if (rows.keys[i] != ~0ull) {
              if (w != i) { rows.keys[w] = rows.keys[i]; memmove(&rows.slots[w * ns], &rows.slots[i * ns], (size_t)ns * 8); }
              w++;
            }
          rows.keys.resize(w);
          rows.slots.resize(w * ns);
          rows.n = w;
        }
      }
    }  // !device_topk
    }
    cudaEventElapsedTime(&ms_h2d, ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&ms_scan, ctx->ev[1], ctx->ev[2]);
    cudaEventElapsedTime(&ms_dev, ctx->ev[0], ctx->ev[4]);
    st->scan_ms = ms_scan;
  }

  // ---- host assembly -------------------------------------------------------------------------------
  const double t_assemble = now_ms();
  const QuerySpec& q = plan.spec;
  const uint64_t n = rows.n;
  Assembler as{plan, rows, {}, {}, {}};
  as.agg_i.resize(plan.aggs.size());
  as.agg_f.resize(plan.aggs.size());
  uint64_t hll_uncorrected = 0;  // estimates Spark would have bias-corrected (tables not available: hll_estimate above)
  for (size_t a = 0; a < plan.aggs.size(); a++) {
    const PlanAgg& pa = plan.aggs[a];
    if (pa.out_float) as.agg_f[a].resize(n);
    else as.agg_i[a].resize(n);
    for (uint64_t r = 0; r < n; r++) {
      if (pa.hll_index >= 0) {
        const uint64_t set = rows.hll_set[r] + pa.hll_index;
        as.agg_f[a][r] = hll_estimate_from(rows.hll_stats[set * 2], rows.hll_stats[set * 2 + 1]);
        hll_uncorrected += hll_needs_bias_tables(rows.hll_stats[set * 2], rows.hll_stats[set * 2 + 1]);
        continue;
      }
      uint64_t raw = rows.slots[r * ns + pa.slot];
      if (pa.const_times_count) as.agg_i[a][r] = (int64_t)raw * pa.const_value;
      else if (pa.slot_is_f64_ord) as.agg_f[a][r] = ord_to_f64((int64_t)raw);
      else if (pa.out_float) { double d; memcpy(&d, &raw, 8); as.agg_f[a][r] = d; }
      else as.agg_i[a][r] = (int64_t)raw;
    }
  }
  as.post.resize(q.post_aggregations.size());
  for (size_t pi = 0; pi < q.post_aggregations.size(); pi++) {
    as.post[pi].resize(n);
    for (uint64_t r = 0; r < n; r++) as.post[pi][r] = as.eval_post(q.post_aggregations[pi], r, pi);
  }

  // decode keys
  std::vector<uint32_t> tk(plan.has_time_key ? n : 0);
  std::vector<std::vector<uint32_t>> dim_ids(plan.dims.size());
  for (size_t d = 0; d < plan.dims.size(); d++)
    if (!plan.dims[d].is_time) dim_ids[d].resize(n);
  for (uint64_t r = 0; r < n; r++) {
    uint64_t k = rows.keys[r];
    if (plan.has_time_key) tk[r] = (uint32_t)(k / plan.tk_stride);
    for (size_t d = 0; d < plan.dims.size(); d++) {
      const PlanDim& pd = plan.dims[d];
      if (pd.is_time) continue;
      dim_ids[d][r] = (uint32_t)((k / pd.stride) % pd.card);
    }
  }

  // having, ordering, limit
  std::vector<uint64_t> order(n);
  std::iota(order.begin(), order.end(), 0);
  if (q.has_having) {
    size_t w = 0;
    for (uint64_t r = 0; r < n; r++)
      if (as.eval_having(q.having, r)) order[w++] = r;
    order.resize(w);
  }
  int64_t limit = -1;
  auto dim_cmp = [&](size_t d, uint64_t a, uint64_t b) -> int {
    const PlanDim& pd = plan.dims[d];
    if (pd.is_time) {
      const std::string& x = plan.time_keys[tk[a]].formatted[pd.time_fmt];
      const std::string& y = plan.time_keys[tk[b]].formatted[pd.time_fmt];
      return utf8_cmp(x.data(), x.size(), y.data(), y.size());
    }
    return dim_ids[d][a] < dim_ids[d][b] ? -1 : (dim_ids[d][a] > dim_ids[d][b] ? 1 : 0);
  };
  auto key_less = [&](uint64_t a, uint64_t b) { return rows.keys[a] < rows.keys[b]; };
  if (q.type == QuerySpec::TopN) {
    // TopNMetricSpec: numeric = descending, inverted = ascending, lexicographic by dimension value
    // (QuerySpecTransforms.scala:281-298)
    limit = q.threshold;
    if (q.metric.kind == TopNMetric::Numeric) {
      std::vector<double> v(n);
      for (uint64_t r : order)
        if (!as.named_value(q.metric.metric, r, &v[r], as.post.size())) throw Invalid("topN metric '" + q.metric.metric + "' is not an aggregation");
      const bool desc = !q.metric.inverted;
      std::sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) {
        if (v[a] != v[b]) return desc ? v[a] > v[b] : v[a] < v[b];
        return key_less(a, b);
      });
    } else {
      const bool desc = q.metric.inverted;
      if (!q.metric.previous_stop.empty()) {
        const std::string& ps = q.metric.previous_stop;
        size_t w = 0;
        for (uint64_t r : order) {
          std::string val;
          const PlanDim& pd = plan.dims[0];
          if (pd.is_time) val = plan.time_keys[tk[r]].formatted[pd.time_fmt];
          else if (auto dict = dim_dictionary(plan, pd)) val = dict->get(dim_ids[0][r]);
          else val = std::to_string(dim_ids[0][r]);
          int c = utf8_cmp(val.data(), val.size(), ps.data(), ps.size());
          if (desc ? c < 0 : c > 0) order[w++] = r;
        }
        order.resize(w);
      }
      std::sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) {
        int c = dim_cmp(0, a, b);
        if (c) return desc ? c > 0 : c < 0;
        return key_less(a, b);
      });
    }
  } else if (q.type == QuerySpec::GroupBy && q.limit.present) {
    // LimitSpec / OrderByColumnSpec (DruidQuerySpec.scala:437-456); sort without limit = Int.MaxValue
    limit = q.limit.limit;
    struct Ord { int dim; int agg; int post; bool asc; bool ts; };
    std::vector<Ord> ords;
    for (auto& c : q.limit.columns) {
      Ord o{-1, -1, -1, c.ascending, false};
      for (size_t d = 0; d < plan.dims.size(); d++)
        if (plan.dims[d].output_name == c.dimension) o.dim = (int)d;
      if (o.dim < 0) o.agg = as.find_agg(c.dimension);
      if (o.dim < 0 && o.agg < 0) o.post = as.find_post(c.dimension);
      if (o.dim < 0 && o.agg < 0 && o.post < 0) {
        if (c.dimension == "timestamp" || c.dimension == "__time") o.ts = true;
        else throw Invalid("limitSpec orders by unknown column '" + c.dimension + "'");
      }
      ords.push_back(o);
    }
    if (!device_topk || ords.size() != 1) {
      std::sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) {
        for (auto& o : ords) {
          int c = 0;
          if (o.dim >= 0) c = dim_cmp((size_t)o.dim, a, b);
          else if (o.ts) {
            int64_t x = plan.time_keys[tk[a]].bucket, y = plan.time_keys[tk[b]].bucket;
            c = x < y ? -1 : (x > y ? 1 : 0);
          } else {
            double x, y;
            if (o.agg >= 0) {
              if (!plan.aggs[o.agg].out_float) {
                int64_t xi = as.agg_i[o.agg][a], yi = as.agg_i[o.agg][b];
                c = xi < yi ? -1 : (xi > yi ? 1 : 0);
                if (c) return o.asc ? c < 0 : c > 0;
                continue;
              }
              x = as.agg_f[o.agg][a]; y = as.agg_f[o.agg][b];
            } else { x = as.post[o.post][a]; y = as.post[o.post][b]; }
            c = x < y ? -1 : (x > y ? 1 : 0);
          }
          if (c) return o.asc ? c < 0 : c > 0;
        }
        return key_less(a, b);
      });
    }
  } else if (n <= (4u << 20)) {
    std::sort(order.begin(), order.end(), key_less);  // deterministic output for small results
  }
  if (limit >= 0 && (uint64_t)limit < order.size()) order.resize((size_t)limit);

  // ---- columns (DruidQuerySpec.scala:614-627: dimensions, aggregations, post-aggregations) --------
  std::unique_ptr<Result> res(new Result());
  const uint64_t m = order.size();
  res->n_rows = (int64_t)m;
  res->pool = &pinned_pool_;
  res->hll_backing = rows.hll;  // the result now owns the sketches
  res->hll_backing_bytes = rows.hll_cap;
  const uint8_t* hll_rows = rows.hll;
  rows.hll = nullptr;
  if (plan.has_bucket) {
    ResultColumn c;
    c.name = "timestamp";
    c.type = SDO_RES_TIMESTAMP;
    c.i64.resize(m);
    for (uint64_t i = 0; i < m; i++) c.i64[i] = plan.time_keys[tk[order[i]]].bucket;
    res->columns.push_back(std::move(c));
  }
  for (size_t d = 0; d < plan.dims.size(); d++) {
    const PlanDim& pd = plan.dims[d];
    ResultColumn c;
    c.name = pd.output_name;
    if (pd.is_time) {
      c.type = SDO_RES_STRING;
      c.dict = std::make_shared<Dictionary>();
      std::map<std::string, uint32_t> seen;
      c.ids.resize(m);
      for (uint64_t i = 0; i < m; i++) {
        const std::string& v = plan.time_keys[tk[order[i]]].formatted[pd.time_fmt];
        auto it = seen.find(v);
        if (it == seen.end()) {
          it = seen.emplace(v, (uint32_t)c.dict->size()).first;
          c.dict->push(v.data(), v.size());
        }
        c.ids[i] = it->second;
      }
    } else if (auto dict = dim_dictionary(plan, pd)) {
      c.type = SDO_RES_STRING;
      c.dict = dict;
      c.ids.resize(m);
      for (uint64_t i = 0; i < m; i++) c.ids[i] = dim_ids[d][order[i]];
    } else {
      c.type = SDO_RES_DIM_INT64;
      c.i64.resize(m);
      for (uint64_t i = 0; i < m; i++) c.i64[i] = (int64_t)dim_ids[d][order[i]];
    }
    res->columns.push_back(std::move(c));
  }
  for (size_t a = 0; a < plan.aggs.size(); a++) {
    const PlanAgg& pa = plan.aggs[a];
    ResultColumn c;
    c.name = pa.name;
    if (pa.out_float) {
      c.type = SDO_RES_FLOAT64;
      c.f64.resize(m);
      for (uint64_t i = 0; i < m; i++) c.f64[i] = as.agg_f[a][order[i]];
      if (pa.hll_index >= 0) {
        // registers are gathered into result order only if sdo_result_hll_registers asks for them
        c.hll_rows = hll_rows;
        c.hll_pick.resize(m);
        for (uint64_t i = 0; i < m; i++) c.hll_pick[i] = rows.hll_set[order[i]] + pa.hll_index;
      }
    } else {
      c.type = SDO_RES_INT64;
      c.i64.resize(m);
      for (uint64_t i = 0; i < m; i++) c.i64[i] = as.agg_i[a][order[i]];
    }
    res->columns.push_back(std::move(c));
  }
  for (size_t pi = 0; pi < q.post_aggregations.size(); pi++) {
    ResultColumn c;
    c.name = q.post_aggregations[pi].name;
    c.type = SDO_RES_FLOAT64;
    c.f64.resize(m);
    for (uint64_t i = 0; i < m; i++) c.f64[i] = as.post[pi][order[i]];
    res->columns.push_back(std::move(c));
  }

  // ---- stats (DruidQueryExecutionView analogue, DruidRDD.scala:194-216) ------------------------------
  // cardinality columns: Spark's HLL++ subtracts an empirical bias for raw estimates up to 5m when linear counting does
  // not apply; those tables are not available here, so such values are the raw estimate and the caller is told how many
  std::string hll_note;
  if (plan.n_hll) hll_note = ",\"hllBiasCorrected\":false,\"hllUncorrectedValues\":" + std::to_string((unsigned long long)hll_uncorrected);
  if (st->shard_merge) hll_note += std::string(",\"shardMerge\":\"") + st->shard_merge + "\",\"exchangeBytes\":" + std::to_string((unsigned long long)st->exchange_bytes) + st->extra_stats;
  char buf[1600];
  snprintf(buf, sizeof buf,
           "{\"strategy\":\"%s\",\"device\":%d,\"rowsScanned\":%lld,\"bytesPerRow\":%d,\"algorithmicBytes\":%lld,\"groups\":%llu,"
           "\"resultRows\":%llu,\"scanKernelMs\":%.6f,\"deviceMs\":%.6f,\"h2dMs\":%.6f,\"totalMs\":%.6f,\"launches\":%d,"
           "\"h2dBytes\":%zu,\"d2hBytes\":%zu,\"deviceTopK\":%s,\"spans\":%zu,\"segments\":%zu,\"keySpace\":%llu,\"jit\":%s,\"jitCompileMs\":%.3f,"
           "\"planMs\":%.4f,\"assembleMs\":%.4f%s}",
           plan.always_false ? "none" : strategy_name(st->strategy), device_, (long long)plan.rows_scanned, plan.bytes_per_row,
           (long long)plan.algorithmic_bytes, (unsigned long long)(groups_total ? groups_total : n), (unsigned long long)m, ms_scan, ms_dev, ms_h2d, now_ms() - t_start_ms,
           st->launches, st->h2d_bytes, st->d2h_bytes, device_topk ? "true" : "false", plan.spans.size(), plan.seg_index.size(),
           (unsigned long long)plan.n_keys, st->jit ? "true" : "false", st->jit_compile_ms, st->plan_ms, now_ms() - t_assemble, hll_note.c_str());
  res->stats = buf;
  return res.release();
}

// ---- pinned result buffers ------------------------------------------------------------------------------
PinnedPool::~PinnedPool() {
  for (auto& b : free_) cudaFreeHost(b.first);
}
void* PinnedPool::get(size_t bytes, size_t* got) {
  bytes = round_up(std::max<size_t>(bytes, 1), 1 << 20);
  {
    std::lock_guard<std::mutex> lk(mu_);
    size_t best = free_.size();
    for (size_t i = 0; i < free_.size(); i++)
      if (free_[i].second >= bytes && free_[i].second <= 2 * bytes && (best == free_.size() || free_[i].second < free_[best].second)) best = i;
    if (best < free_.size()) {
      void* p = free_[best].first;
      *got = free_[best].second;
      pooled_ -= *got;
      free_.erase(free_.begin() + best);
      return p;
    }
  }
  void* p = nullptr;
  SDO_CUDA(cudaMallocHost(&p, bytes));
  *got = bytes;
  return p;
}
void PinnedPool::put(void* p, size_t bytes) {
  std::lock_guard<std::mutex> lk(mu_);
  if (pooled_ + bytes > (16ull << 30)) { cudaFreeHost(p); return; }
  free_.emplace_back(p, bytes);
  pooled_ += bytes;
}

// Large, unordered result without host-side post-processing: decode keys and aggregate slots into
// result columns on the device, then one D2H copy per column into pinned buffers the result owns.
Result* Engine::finalize_columnar(QueryCtx* ctx, const Plan& plan, ScanState* st, double t_start_ms) {
  cudaStream_t s = ctx->stream;
  const uint64_t n = st->n_rows;
  const QuerySpec& q = plan.spec;
  std::unique_ptr<Result> res(new Result());
  res->pool = &pinned_pool_;
  res->n_rows = (int64_t)n;
  MatParams mp;
  memset(&mp, 0, sizeof mp);
  mp.n = n;
  mp.ns = plan.n_slots;
  mp.has_tk = plan.has_time_key ? 1 : 0;
  mp.tk_stride = plan.tk_stride;
  std::vector<void*> dev_bufs;
  struct Copy { void* dev; size_t bytes; size_t col; };
  std::vector<Copy> copies;
  auto dev_alloc = [&](size_t bytes) {
    void* p = nullptr;
    SDO_CUDA(cudaMallocAsync(&p, std::max<size_t>(bytes, 16), s));
    dev_bufs.push_back(p);
    return p;
  };
  auto add_column = [&](const std::string& name, int type, void* dev, size_t bytes) {
    ResultColumn c;
    c.name = name;
    c.type = type;
    res->columns.push_back(std::move(c));
    copies.push_back({dev, bytes, res->columns.size() - 1});
  };
  SDO_CUDA(cudaEventRecord(ctx->ev[3], s));
  if (plan.has_bucket) {
    std::vector<long long> lut(plan.time_keys.size());
    for (size_t i = 0; i < lut.size(); i++) lut[i] = plan.time_keys[i].bucket;
    void* dl = dev_alloc(lut.size() * 8);
    SDO_CUDA(cudaMemcpyAsync(dl, lut.data(), lut.size() * 8, cudaMemcpyHostToDevice, s));
    SDO_CUDA(cudaStreamSynchronize(s));  // `lut` is pageable and goes out of scope
    mp.bucket_lut = (const long long*)dl;
    mp.ts_out = (long long*)dev_alloc(n * 8);
    add_column("timestamp", SDO_RES_TIMESTAMP, mp.ts_out, n * 8);
  }
  bool any_time_dim = false;
  for (auto& pd : plan.dims) any_time_dim |= pd.is_time;
  if (any_time_dim) mp.tk_out = (uint32_t*)dev_alloc(n * 4);
  for (auto& pd : plan.dims) {
    if (pd.is_time) {
      // ids = time key; dictionary entry i = the formatted value of time key i
      add_column(pd.output_name, SDO_RES_STRING, mp.tk_out, n * 4);
      auto d = std::make_shared<Dictionary>();
      for (auto& tk : plan.time_keys) d->push(tk.formatted[pd.time_fmt].data(), tk.formatted[pd.time_fmt].size());
      res->columns.back().dict = d;
      continue;
    }
    const int k = mp.n_dims++;
    mp.stride[k] = pd.stride;
    mp.card[k] = pd.card;
    if (auto dict = dim_dictionary(plan, pd)) {
      mp.dim_is64[k] = 0;
      mp.dim_out[k] = dev_alloc(n * 4);
      add_column(pd.output_name, SDO_RES_STRING, mp.dim_out[k], n * 4);
      res->columns.back().dict = dict;
    } else {
      mp.dim_is64[k] = 1;
      mp.dim_out[k] = dev_alloc(n * 8);
      add_column(pd.output_name, SDO_RES_DIM_INT64, mp.dim_out[k], n * 8);
    }
  }
  for (auto& pa : plan.aggs) {
    const int k = mp.n_aggs++;
    mp.agg_slot[k] = pa.slot;
    mp.agg_mul[k] = pa.const_value;
    mp.agg_mode[k] = pa.const_times_count ? MAT_I64_TIMES : (pa.slot_is_f64_ord ? MAT_F64_ORD : (pa.out_float ? MAT_F64_BITS : MAT_I64));
    mp.agg_out[k] = dev_alloc(n * 8);
    add_column(pa.name, pa.out_float ? SDO_RES_FLOAT64 : SDO_RES_INT64, mp.agg_out[k], n * 8);
  }
  launch_materialize(st->rows_keys, st->rows_slots, mp, s);
  st->launches += 1;
  // the same device buffer may back several columns (time-format dimensions share tk_out)
  for (auto& c : copies) {
    ResultColumn& col = res->columns[c.col];
    col.raw = pinned_pool_.get(c.bytes, &col.raw_bytes);
    SDO_CUDA(cudaMemcpyAsync(col.raw, c.dev, c.bytes, cudaMemcpyDeviceToHost, s));
    st->d2h_bytes += c.bytes;
  }
  SDO_CUDA(cudaEventRecord(ctx->ev[4], s));
  SDO_CUDA(cudaStreamSynchronize(s));
  for (void* p : dev_bufs) SDO_CUDA(cudaFreeAsync(p, s));
  float ms_h2d = 0, ms_scan = 0, ms_dev = 0;
  cudaEventElapsedTime(&ms_h2d, ctx->ev[0], ctx->ev[1]);
  cudaEventElapsedTime(&ms_scan, ctx->ev[1], ctx->ev[2]);
  cudaEventElapsedTime(&ms_dev, ctx->ev[0], ctx->ev[4]);
  st->scan_ms = ms_scan;
  char buf[1024];
  snprintf(buf, sizeof buf,
           "{\"strategy\":\"%s\",\"device\":%d,\"rowsScanned\":%lld,\"bytesPerRow\":%d,\"algorithmicBytes\":%lld,\"groups\":%llu,"
           "\"resultRows\":%llu,\"scanKernelMs\":%.6f,\"deviceMs\":%.6f,\"h2dMs\":%.6f,\"totalMs\":%.6f,\"launches\":%d,"
           "\"h2dBytes\":%zu,\"d2hBytes\":%zu,\"deviceTopK\":false,\"spans\":%zu,\"segments\":%zu,\"keySpace\":%llu,\"jit\":%s,\"jitCompileMs\":%.3f,"
           "\"columnar\":true%s}",
           strategy_name(st->strategy), device_, (long long)plan.rows_scanned, plan.bytes_per_row, (long long)plan.algorithmic_bytes,
           (unsigned long long)n, (unsigned long long)n, ms_scan, ms_dev, ms_h2d, now_ms() - t_start_ms, st->launches, st->h2d_bytes, st->d2h_bytes,
           plan.spans.size(), plan.seg_index.size(), (unsigned long long)plan.n_keys, st->jit ? "true" : "false", st->jit_compile_ms,
           st->shard_merge ? (std::string(",\"shardMerge\":\"") + st->shard_merge + "\",\"exchangeBytes\":" + std::to_string((unsigned long long)st->exchange_bytes)).c_str() : "");
  res->stats = buf;
  (void)q;
  return res.release();
}

// Plans are cached by the query's JSON text: dashboards repeat queries, and planning a day-granularity query
// (thousands of spans, calendar maths, dictionary searches) can cost more than running it.  A cached plan is
// reused only while the datasource object and its version are the ones it was made for.  Results are never cached.
std::shared_ptr<const Plan> Engine::plan_for(const char* json, size_t len, std::shared_lock<std::shared_mutex>* lock_out, QuerySpec* search_out) {
  std::string key(json, len);
  std::shared_ptr<const Plan> cached;
  {
    std::lock_guard<std::mutex> lk(plan_mu_);
    auto it = plan_cache_.find(key);
    if (it != plan_cache_.end()) cached = it->second;
  }
  if (cached) {
    auto ds = store_.get(cached->spec.data_source);
    std::shared_lock<std::shared_mutex> lk(ds->mu);
    if (ds.get() == cached->ds.get() && ds->version == cached->ds_version) {
      *lock_out = std::move(lk);
      return cached;
    }
  }
  QuerySpec spec = parse_query_spec(json, len);
  if ((spec.type == QuerySpec::Search || spec.type == QuerySpec::Select) && search_out) {
    *search_out = std::move(spec);
    return nullptr;
  }
  auto ds = store_.get(spec.data_source);
  std::shared_lock<std::shared_mutex> lk(ds->mu);
  auto plan = std::make_shared<Plan>(make_plan(spec, ds, opt_));
  plan->ds_version = ds->version;
  {
    std::lock_guard<std::mutex> g(plan_mu_);
    if (plan_cache_.size() >= 256) plan_cache_.clear();
    plan_cache_[key] = plan;
  }
  *lock_out = std::move(lk);
  return plan;
}

// Search (SearchQuerySpec, DruidQuerySpec.scala:870-927; emitted by the planner for a group-by on one dimension without
// aggregations, QuerySpecTransforms.scala:224-278): per search dimension one group-by with a row count, the search terms
// added to the filter as predicates on that dimension's dictionary.  Result columns as Druid's: dimension, value, count;
// ordered by value (lexicographic or strlen) then dimension, cut at `limit`.
Result* Engine::execute_search(const QuerySpec& spec, double t0) {
  auto ds = store_.get(spec.data_source);
  std::shared_lock<std::shared_mutex> lk(ds->mu);
  check_cancel(spec.query_id);
  std::vector<std::string> dims = spec.search_dimensions;
  if (dims.empty())
    for (auto& d : ds->dims) dims.push_back(d.name);
  struct Hit { std::string dim, value; int64_t count; };
  std::vector<Hit> hits;
  double scan_ms = 0, dev_ms = 0;
  int64_t rows_scanned = 0, launches = 0;
  for (auto& dname : dims) {
    if (ds->dim_index(dname) < 0) continue;  // Druid: an unknown dimension has no values to report
    QuerySpec g;
    g.type = QuerySpec::GroupBy;
    g.data_source = spec.data_source;
    DimensionSpec d;
    d.dimension = d.output_name = dname;
    g.dimensions.push_back(d);
    AggregationSpec cnt;
    cnt.kind = AggregationSpec::Count;
    cnt.name = "count";
    g.aggregations.push_back(cnt);
    g.intervals = spec.intervals;
    g.by_segments = spec.by_segments;
    g.segments = spec.segments;
    g.query_id = spec.query_id;
    g.gpu_strategy = spec.gpu_strategy;
    g.gpu_jit = spec.gpu_jit;
    FilterNode all;
    all.kind = FilterNode::And;
    if (spec.has_filter) all.fields.push_back(spec.filter);
    for (FilterNode t : spec.search_terms) {
      t.dimension = dname;
      all.fields.push_back(t);
    }
    if (!all.fields.empty()) { g.has_filter = true; g.filter = all; }
    Plan plan = make_plan(g, ds, opt_);
    std::unique_ptr<Result> r(execute_plan(plan, now_ms()));
    const ResultColumn& vc = r->columns.at(0);
    const ResultColumn& cc = r->columns.at(1);
    for (int64_t i = 0; i < r->n_rows; i++) {
      std::string v;
      if (vc.type == SDO_RES_STRING) {
        const uint32_t id = vc.raw ? ((const uint32_t*)vc.raw)[i] : vc.ids[i];
        v = vc.dict->get(id);
      } else {
        v = std::to_string(vc.raw ? ((const int64_t*)vc.raw)[i] : vc.i64[i]);
      }
      if (v.empty()) continue;  // Druid's search does not report the null value
      hits.push_back({dname, std::move(v), cc.raw ? ((const int64_t*)cc.raw)[i] : cc.i64[i]});
    }
    // carry the scan statistics of the sub-queries
    const std::string& sj = r->stats;
    auto num = [&](const char* key) {
      size_t k = sj.find(std::string("\"") + key + "\":");
      return k == std::string::npos ? 0.0 : atof(sj.c_str() + k + strlen(key) + 3);
    };
    scan_ms += num("scanKernelMs");
    dev_ms += num("deviceMs");
    rows_scanned += (int64_t)num("rowsScanned");
    launches += (int64_t)num("launches");
  }
  const bool by_len = spec.search_sort_strlen;
  std::sort(hits.begin(), hits.end(), [by_len](const Hit& a, const Hit& b) {
    if (by_len && a.value.size() != b.value.size()) return a.value.size() < b.value.size();
    int c = utf8_cmp(a.value.data(), a.value.size(), b.value.data(), b.value.size());
    if (c) return c < 0;
    return a.dim < b.dim;
  });
  if ((int64_t)hits.size() > spec.search_limit) hits.resize((size_t)spec.search_limit);
  std::unique_ptr<Result> res(new Result());
  res->n_rows = (int64_t)hits.size();
  ResultColumn cd, cv, cn;
  cd.name = "dimension";
  cv.name = "value";
  cn.name = "count";
  cd.type = cv.type = SDO_RES_STRING;
  cn.type = SDO_RES_INT64;
  cd.dict = std::make_shared<Dictionary>();
  cv.dict = std::make_shared<Dictionary>();
  std::map<std::string, uint32_t> dim_ids;
  for (auto& h : hits) {
    auto it = dim_ids.find(h.dim);
    if (it == dim_ids.end()) {
      it = dim_ids.emplace(h.dim, (uint32_t)cd.dict->size()).first;
      cd.dict->push(h.dim.data(), h.dim.size());
    }
    cd.ids.push_back(it->second);
    cv.ids.push_back((uint32_t)cv.dict->size());
    cv.dict->push(h.value.data(), h.value.size());
    cn.i64.push_back(h.count);
  }
  res->columns.push_back(std::move(cd));
  res->columns.push_back(std::move(cv));
  res->columns.push_back(std::move(cn));
  char buf[512];
  snprintf(buf, sizeof buf,
           "{\"queryType\":\"search\",\"strategy\":\"per-dimension group-by\",\"rowsScanned\":%lld,\"resultRows\":%lld,\"scanKernelMs\":%.6f,"
           "\"deviceMs\":%.6f,\"totalMs\":%.6f,\"launches\":%lld,\"dimensions\":%d}",
           (long long)rows_scanned, (long long)res->n_rows, scan_ms, dev_ms, now_ms() - t0, (long long)launches, (int)dims.size());
  res->stats = buf;
  return res.release();
}

// Select (SelectSpec, DruidQuerySpec.scala:980-1098): the rows that pass interval and filter, a page at a time, in
// segment order then row order (= time order inside a segment), with the requested dimension and metric values.
// Three small kernels: per-unit counts of passing rows; the host then knows which units hold the page
// [start offset per segment, threshold rows in all]; those units emit (segment, row) in row order; one gather per column.
// Result columns: timestamp, segmentId, offset, dimensions..., metrics...; stats carry the pagingIdentifiers for the
// next page (pass them back with fromNext = true).
Result* Engine::execute_select(const QuerySpec& spec, double t0) {
  if (device_ < 0) throw NoDevice("this engine was created without a CUDA device (planning only); there is no CPU execution path");
  auto ds = store_.get(spec.data_source);
  std::shared_lock<std::shared_mutex> lk(ds->mu);
  check_cancel(spec.query_id);
  // the rows of a group-by without dimensions over the same intervals and filter
  QuerySpec g;
  g.type = QuerySpec::GroupBy;
  g.data_source = spec.data_source;
  g.intervals = spec.intervals;
  g.by_segments = spec.by_segments;
  g.segments = spec.segments;
  g.has_filter = spec.has_filter;
  g.filter = spec.filter;
  g.query_id = spec.query_id;
  Plan plan = make_plan(g, ds, opt_);
  struct OutCol { std::string name; bool metric; int index; int slot; int width; };
  std::vector<OutCol> out_cols;
  auto add_dim = [&](const std::string& n) {
    if (n == "__time" || n == "timestamp") return;  // the timestamp column is always there
    int di = ds->dim_index(n);
    if (di < 0) throw NotFound("unknown dimension '" + n + "' in datasource '" + ds->name + "'");
    out_cols.push_back({n, false, di, (int)plan.columns.size(), ds->dims[di].width});
    plan.columns.push_back({false, di, ds->dims[di].width});
  };
  auto add_metric = [&](const std::string& n) {
    int mi = ds->metric_index(n);
    if (mi < 0) throw NotFound("unknown metric '" + n + "' in datasource '" + ds->name + "'");
    out_cols.push_back({n, true, mi, (int)plan.columns.size(), 8});
    plan.columns.push_back({true, mi, 8});
  };
  if (spec.select_dims.empty() && spec.select_metrics.empty()) {
    for (auto& d : ds->dims) add_dim(d.name);
    for (auto& m : ds->metrics) add_metric(m.name);
  } else {
    for (auto& n : spec.select_dims) add_dim(n);
    for (auto& n : spec.select_metrics) add_metric(n);
  }
  if (plan.columns.size() > 250) throw Unsupported("select of more than 250 columns");

  std::map<std::string, int64_t> paging(spec.paging_ids.begin(), spec.paging_ids.end());
  std::vector<uint32_t> h_seg, h_row;
  std::vector<int64_t> h_off;
  std::vector<std::vector<unsigned long long>> h_cols(out_cols.size());
  int64_t launches = 0;
  if (!plan.always_false && plan.n_units > 0) {
    QueryCtx* ctx = acquire_ctx();
    ScanState st;
    uint32_t* d_counts = nullptr;
    SelectUnit* d_units = nullptr;
    uint32_t *d_seg = nullptr, *d_row = nullptr;
    void* d_col = nullptr;
    auto cleanup = [&] {
      cudaStreamSynchronize(ctx->stream);
      for (void* q : {(void*)d_counts, (void*)d_units, (void*)d_seg, (void*)d_row, d_col})
        if (q) cudaFreeAsync(q, ctx->stream);
      free_state(ctx, &st);
      release_ctx(ctx);
    };
    try {
      cudaStream_t s = ctx->stream;
      prepare_params(ctx, plan, &st);
      const ScanParams& p = st.params;
      SDO_CUDA(cudaMallocAsync((void**)&d_counts, (size_t)plan.n_units * 4, s));
      launch_select_count(p, d_counts, s);
      std::vector<uint32_t> counts(plan.n_units);
      SDO_CUDA(cudaMemcpyAsync(counts.data(), d_counts, counts.size() * 4, cudaMemcpyDeviceToHost, s));
      SDO_CUDA(cudaStreamSynchronize(s));
      launches += 1;
      // which units hold the page: walk the spans in order, keeping each segment's count of matching rows so far
      std::vector<SelectUnit> units;
      std::map<uint32_t, int64_t> seg_pos;  // plan segment -> matching rows before the current unit
      int64_t remaining = spec.paging_threshold;
      uint32_t out_base = 0;
      for (size_t si = 0; si < plan.spans.size() && remaining > 0; si++) {
        const DevSpan& sp = plan.spans[si];
        const uint32_t u_end = si + 1 < plan.spans.size() ? plan.spans[si + 1].unit_begin : plan.n_units;
        const Segment& sg = *ds->segments[plan.seg_index[sp.seg]];
        auto it = paging.find(segment_identifier(*ds, sg));
        const int64_t start = it == paging.end() ? 0 : it->second + (spec.paging_from_next ? 1 : 0);
        int64_t& pos = seg_pos[sp.seg];
        for (uint32_t u = sp.unit_begin; u < u_end && remaining > 0; u++) {
          const int64_t c = counts[u], lo = pos;
          pos += c;
          if (pos <= start || c == 0) continue;
          const int64_t skip = std::max<int64_t>(start, lo) - lo;
          const int64_t take = std::min<int64_t>(c - skip, remaining);
          units.push_back({u, (uint32_t)skip, (uint32_t)take, out_base});
          for (int64_t k = 0; k < take; k++) h_off.push_back(lo + skip + k);
          out_base += (uint32_t)take;
          remaining -= take;
        }
      }
      const uint32_t n_out = out_base;
      if (n_out) {
        SDO_CUDA(cudaMallocAsync((void**)&d_units, units.size() * sizeof(SelectUnit), s));
        SDO_CUDA(cudaMemcpyAsync(d_units, units.data(), units.size() * sizeof(SelectUnit), cudaMemcpyHostToDevice, s));
        SDO_CUDA(cudaMallocAsync((void**)&d_seg, (size_t)n_out * 4, s));
        SDO_CUDA(cudaMallocAsync((void**)&d_row, (size_t)n_out * 4, s));
        launch_select_emit(p, d_units, (uint32_t)units.size(), d_seg, d_row, s);
        h_seg.resize(n_out);
        h_row.resize(n_out);
        SDO_CUDA(cudaMemcpyAsync(h_seg.data(), d_seg, (size_t)n_out * 4, cudaMemcpyDeviceToHost, s));
        SDO_CUDA(cudaMemcpyAsync(h_row.data(), d_row, (size_t)n_out * 4, cudaMemcpyDeviceToHost, s));
        SDO_CUDA(cudaMallocAsync(&d_col, (size_t)n_out * 8, s));
        for (size_t c = 0; c < out_cols.size(); c++) {
          const OutCol& oc = out_cols[c];
          launch_select_gather(p, oc.slot, oc.width, d_seg, d_row, n_out, d_col, s);
          h_cols[c].resize(n_out);  // ids arrive as uint32: two per word
          SDO_CUDA(cudaMemcpyAsync(h_cols[c].data(), d_col, (size_t)n_out * (oc.width == 8 ? 8 : 4), cudaMemcpyDeviceToHost, s));
          SDO_CUDA(cudaStreamSynchronize(s));  // d_col is reused by the next column
        }
        SDO_CUDA(cudaStreamSynchronize(s));
        launches += 1 + (int64_t)out_cols.size();
      }
    } catch (...) {
      cleanup();
      throw;
    }
    cleanup();
  }
  // ---- result columns ---------------------------------------------------------------------------------------------
  const size_t n = h_seg.size();
  std::unique_ptr<Result> res(new Result());
  res->n_rows = (int64_t)n;
  ResultColumn ts, sid, off;
  ts.name = "timestamp";
  ts.type = SDO_RES_TIMESTAMP;
  sid.name = "segmentId";
  sid.type = SDO_RES_STRING;
  sid.dict = std::make_shared<Dictionary>();
  off.name = "offset";
  off.type = SDO_RES_INT64;
  std::map<uint32_t, uint32_t> sid_ids;
  for (size_t i = 0; i < n; i++) {
    const Segment& sg = *ds->segments[plan.seg_index[h_seg[i]]];
    // the row's (truncated) time: the run of the time index that holds it
    const size_t run = (size_t)(std::upper_bound(sg.time_rows.begin(), sg.time_rows.end(), (int64_t)h_row[i]) - sg.time_rows.begin()) - 1;
    ts.i64.push_back(sg.time_vals.at(run));
    auto it = sid_ids.find(h_seg[i]);
    if (it == sid_ids.end()) {
      it = sid_ids.emplace(h_seg[i], (uint32_t)sid.dict->size()).first;
      const std::string id = segment_identifier(*ds, sg);
      sid.dict->push(id.data(), id.size());
      paging[id] = h_off[i];
    }
    sid.ids.push_back(it->second);
    off.i64.push_back(h_off[i]);
    paging[sid.dict->get(it->second)] = std::max(paging[sid.dict->get(it->second)], h_off[i]);
  }
  res->columns.push_back(std::move(ts));
  res->columns.push_back(std::move(sid));
  res->columns.push_back(std::move(off));
  for (size_t c = 0; c < out_cols.size(); c++) {
    const OutCol& oc = out_cols[c];
    ResultColumn col;
    col.name = oc.name;
    if (oc.metric) {
      if (ds->metrics[oc.index].is_float) {
        col.type = SDO_RES_FLOAT64;
        col.f64.resize(n);
        if (n) memcpy(col.f64.data(), h_cols[c].data(), n * 8);
      } else {
        col.type = SDO_RES_INT64;
        col.i64.resize(n);
        if (n) memcpy(col.i64.data(), h_cols[c].data(), n * 8);
      }
    } else {
      const uint32_t* ids = (const uint32_t*)h_cols[c].data();
      if (ds->dims[oc.index].enc == DimEncoding::Dict) {
        col.type = SDO_RES_STRING;
        col.dict = ds->dims[oc.index].dict;
        col.ids.assign(ids, ids + n);
      } else {
        col.type = SDO_RES_DIM_INT64;
        col.i64.resize(n);
        for (size_t i = 0; i < n; i++) col.i64[i] = (int64_t)ids[i];
      }
    }
    res->columns.push_back(std::move(col));
  }
  std::string pj = "{";
  for (auto& kv : paging) {
    if (pj.size() > 1) pj += ",";
    json_escape(pj, kv.first);
    pj += ":" + std::to_string((long long)kv.second);
  }
  pj += "}";
  char buf[256];
  snprintf(buf, sizeof buf, "{\"queryType\":\"select\",\"rowsScanned\":%lld,\"resultRows\":%lld,\"launches\":%lld,\"totalMs\":%.6f,\"pagingIdentifiers\":",
           (long long)plan.rows_scanned, (long long)n, (long long)launches, now_ms() - t0);
  res->stats = std::string(buf) + pj + "}";
  return res.release();
}

// A plan whose filter or intervals match nothing still has to produce its (empty) tables when something reads them:
// the other ranks of a sharded query, or a timeseries that reports empty buckets.  Same plan, zero units to scan.
static Plan scan_nothing_variant(const Plan& plan) {
  Plan v = plan;
  v.spans.clear();
  v.n_units = 0;
  v.coarse.assign(2, 0);
  v.rows_scanned = 0;
  v.always_false = false;
  return v;
}

Result* Engine::execute_plan(const Plan& plan_in, double t0) {
  if (device_ < 0) throw NoDevice("this engine was created without a CUDA device (planning only); there is no CPU execution path");
  Plan empty_variant;
  const Plan* use = &plan_in;
  if (plan_in.always_false && plan_in.zero_fill) { empty_variant = scan_nothing_variant(plan_in); use = &empty_variant; }
  const Plan& plan = *use;
  QueryCtx* ctx = acquire_ctx();
  ScanState st;
  st.plan_ms = now_ms() - t0;
  Result* r = nullptr;
  try {
    if (!plan.always_false) run_scan(ctx, plan, &st);
    check_cancel(plan.spec.query_id);
    r = finalize(ctx, plan, &st, t0);
  } catch (...) {
    cudaStreamSynchronize(ctx->stream);
    free_state(ctx, &st);
    release_ctx(ctx);
    throw;
  }
  free_state(ctx, &st);
  release_ctx(ctx);
  return r;
}

Result* Engine::execute(const char* json, size_t len) {
  const double t0 = now_ms();
  std::shared_lock<std::shared_mutex> lk;
  QuerySpec search;
  std::shared_ptr<const Plan> plan_ptr = plan_for(json, len, &lk, &search);
  const QuerySpec& spec = plan_ptr ? plan_ptr->spec : search;
  RunningQuery running(this, spec.query_id);
  if (!plan_ptr) return search.type == QuerySpec::Select ? execute_select(search, t0) : execute_search(search, t0);
  check_cancel(spec.query_id);
  return execute_plan(*plan_ptr, t0);
}

std::string Engine::explain(const char* json, size_t len) {
  QuerySpec spec = parse_query_spec(json, len);
  auto ds = store_.get(spec.data_source);
  std::shared_lock<std::shared_mutex> lk(ds->mu);
  Plan plan = make_plan(spec, ds, opt_);
  return explain_json(plan);
}

// ---- sharded execution ---------------------------------------------------------------------------------
void Engine::comm_init(const void* id, int rank, int world) {
  if (device_ < 0) throw NoDevice("this engine was created without a CUDA device");
  if (world < 1 || rank < 0 || rank >= world) throw Invalid("sdo_engine_comm_init: rank / world out of range");
  std::lock_guard<std::mutex> lk(comm_mu_);
  if (comm_) throw Invalid("sdo_engine_comm_init: this engine already has a communicator");
  SDO_CUDA(cudaSetDevice(device_));
  comm_ = nccl_comm_init(id, rank, world);
  comm_rank_ = rank;
  comm_world_ = world;
}

// ---- a query's turn on the communicator ------------------------------------------------------------------------------
// Collectives of one communicator must be issued in the same order on every rank.  A sharded query holds the turn from
// its first collective to its last (its scan runs before that, concurrently with other queries).  Callers that run
// several sharded queries at once give each a turn number -- the same number on every rank, consecutive from 0 -- and
// the engine issues their collectives in that order; without a number (turn < 0) the caller keeps the ranks in step.
static bool turn_trace() {
  static const bool on = [] { const char* v = getenv("SDO_TURN_TRACE"); return v && v[0] == '1'; }();
  return on;
}
#define SDO_TRACE(...) do { if (turn_trace()) { fprintf(stderr, "[sdo rank %d %.3f] ", comm_rank_trace, now_ms()); fprintf(stderr, __VA_ARGS__); fprintf(stderr, "\n"); fflush(stderr); } } while (0)
static thread_local int comm_rank_trace = -1;

struct Engine::CommTurn {
  Engine* e;
  int64_t turn;
  bool held = false;
  CommTurn(Engine* e_, int64_t turn_) : e(e_), turn(turn_) { comm_rank_trace = e->comm_rank_; }
  void acquire() {
    SDO_TRACE("turn %lld wants the communicator (next %lld)", (long long)turn, (long long)e->next_turn_);
    std::unique_lock<std::mutex> lk(e->turn_mu_);
    auto ready = [&] { return !e->turn_busy_ && (turn < 0 || e->next_turn_ == turn); };
    if (turn >= 0 && turn < e->next_turn_) throw Invalid("sdo_query_execute_sharded_at: turn " + std::to_string((long long)turn) + " has already been taken");
    if (!e->turn_cv_.wait_for(lk, std::chrono::seconds(120), ready))
      throw std::runtime_error("sharded query: turn " + std::to_string((long long)turn) + " never came (the turn before it was not issued on this rank)");
    e->turn_busy_ = true;
    held = true;
    SDO_TRACE("turn %lld holds the communicator", (long long)turn);
  }
  // The order NCCL cares about is the order its calls are ISSUED in: once a query's last collective has been queued the next
  // turn may issue its own, while this query's stream still waits for the data and the host finishes the result.
  void release() {
    std::unique_lock<std::mutex> lk(e->turn_mu_);
    if (!held || released) return;
    released = true;
    e->turn_busy_ = false;
    if (turn >= 0) e->next_turn_ = turn + 1;
    SDO_TRACE("turn %lld released", (long long)turn);
    e->turn_cv_.notify_all();
  }
  bool released = false;
  ~CommTurn() {
    std::unique_lock<std::mutex> lk(e->turn_mu_);
    SDO_TRACE("turn %lld done (held %d)", (long long)turn, (int)held);
    if (held && released) {
    } else if (held) {
      e->turn_busy_ = false;
      if (turn >= 0) e->next_turn_ = turn + 1;
    } else if (turn >= 0 && turn >= e->next_turn_) {
      // the query failed before its first collective: give the turn up so that later ones do not wait for it
      e->turn_cv_.wait_for(lk, std::chrono::seconds(120), [&] { return !e->turn_busy_ && e->next_turn_ == turn; });
      if (e->next_turn_ == turn) e->next_turn_ = turn + 1;
    }
    e->turn_cv_.notify_all();
  }
};

// which merge a sharded query takes: decided from quantities that are the same on every rank (key space after dictionary
// alignment, slots, sketches, forced strategy) -- never from this rank's row counts
enum class ShardMerge { Gather, Reduce, Sparse };
static ShardMerge shard_merge_mode(const Plan& plan, const PlanOptions& opt, uint64_t blob_bytes) {
  static const uint64_t gather_max = [] { const char* v = getenv("SDO_GATHER_MAX_BYTES"); return v ? (uint64_t)atoll(v) : (256ull << 10); }();
  static const uint64_t reduce_max = [] { const char* v = getenv("SDO_REDUCE_MAX_BYTES"); return v ? (uint64_t)atoll(v) : (64ull << 20); }();
  const bool dense_everywhere = plan.spec.gpu_strategy != "hash" && plan.spec.gpu_strategy != "emit" &&
                                (plan.n_keys <= opt.global_max_keys || (plan.dev_dims.empty() && plan.n_keys <= (1ull << 31)));
  if (!dense_everywhere || plan.spec.gpu_shard_merge == "sparse") return ShardMerge::Sparse;
  if (plan.spec.gpu_shard_merge == "gather") return ShardMerge::Gather;
  if (plan.spec.gpu_shard_merge == "reduce") return ShardMerge::Reduce;
  if (blob_bytes <= gather_max) return ShardMerge::Gather;
  if (blob_bytes <= reduce_max) return ShardMerge::Reduce;
  return ShardMerge::Sparse;
}

// One sharded query: scan the local segments, merge the partial aggregates over NCCL, finish.  Everything is queued on
// one stream by this call.  Three merges (merge algebra of PostAggregate.scala:39-60):
//  * Gather: small dense tables -- one all-gather of the raw partial blobs (aggregate slots + HLL registers), folded in
//    rank order on `root` (bit-reproducible);
//  * Reduce: larger dense tables -- the table is transposed to [slot][key] so that every slot is one contiguous run with
//    one operator, then one ncclReduce per slot (int64 / float64 sum, int64 min / max) and one uint8 max over the HLL
//    registers, all in one NCCL group, to `root`;
//  * Sparse: hash tables and key spaces too large to reduce densely -- every rank turns its groups into (key, slots)
//    rows, splits them by owner = mix(key) mod world, exchanges them with grouped ncclSend / ncclRecv, merges what it
//    receives and owns a disjoint share of the groups.  A query that wants the top K by one aggregate takes the top K
//    of every share, gathers the K * world candidates and finishes on `root`; any other query returns every rank's share
//    on that rank (*out is non-NULL everywhere; the caller concatenates, as Spark unions partitions).
Result* Engine::execute_sharded(const char* json, size_t len, int root, int64_t turn_no) {
  if (!comm_) throw Invalid("sdo_query_execute_sharded: sdo_engine_comm_init has not been called on this engine");
  if (root < 0 || root >= comm_world_) throw Invalid("sdo_query_execute_sharded: root out of range");
  const double t0 = now_ms();
  CommTurn turn(this, turn_no);
  std::shared_lock<std::shared_mutex> lk;
  std::shared_ptr<const Plan> plan_ptr = plan_for(json, len, &lk);
  RunningQuery running(this, plan_ptr->spec.query_id);
  // every rank must produce the same table layout even when its own shard has no matching rows: scan zero units
  Plan variant;
  const Plan* plan = plan_ptr.get();
  if (plan->always_false) {
    variant = scan_nothing_variant(*plan);
    plan = &variant;
  }
  const uint64_t cells = plan->n_keys * (uint64_t)plan->n_slots;
  const uint64_t blob = round_up(std::max<uint64_t>(cells, 1) * 8, 256) + (plan->n_hll ? plan->n_keys * (uint64_t)plan->n_hll * HLL_M : 0);
  const ShardMerge mode = shard_merge_mode(*plan, opt_, blob);
  if (mode == ShardMerge::Sparse && plan->n_hll) throw Unsupported("cardinality aggregations over a sparse key space cannot be merged across GPUs");
  if (mode == ShardMerge::Sparse) {
    // Key spaces that are sparse on their own account (not forced for a test) have about as many groups as rows on every
    // rank: aggregating locally would pay a pass of atomics here and a second one at the owner for nothing, so the rows
    // themselves are exchanged (STRAT_EMIT) and aggregated once, where they are owned.
    bool emit = plan->spec.gpu_strategy == "emit" || (plan->spec.gpu_strategy.empty() && plan->spec.gpu_shard_merge.empty());
    std::vector<int> writers(plan->n_slots, 0);
    for (auto& a : plan->dev_aggs) writers[a.slot]++;
    for (int x = 1; x < plan->n_slots; x++) emit = emit && writers[x] == 1;
    if (emit || (plan->strategy != STRAT_HASH && plan->strategy != STRAT_GLOBAL)) {
      // (else: a forced sparse merge of a small key space: the groups come out of a hash table)
      variant = *plan;
      variant.strategy = emit ? STRAT_EMIT : STRAT_HASH;
      plan = &variant;
    }
  }
  QueryCtx* ctx = acquire_ctx();
  ScanState st;
  st.plan_ms = now_ms() - t0;
  Result* r = nullptr;
  std::vector<void*> temps;  // stream-ordered allocations of this call
  auto dev_alloc = [&](size_t bytes) {
    void* p = nullptr;
    SDO_CUDA(cudaMallocAsync(&p, std::max<size_t>(bytes, 16), ctx->stream));
    temps.push_back(p);
    return p;
  };
  try {
    cudaStream_t s = ctx->stream;
    if (plan->strategy == STRAT_EMIT && comm_world_ > 1 && comm_world_ <= 64) {
      // the scan writes every row straight into its owner's region.  Keys whose round-robin share (key mod world) fits a
      // direct-indexed table are dealt out that way -- the owner then merges by plain indexing; others by hash.
      const uint64_t per_rank = (plan->n_keys + (uint64_t)comm_world_ - 1) / (uint64_t)comm_world_;
      st.emit_owners = comm_world_;
      st.emit_interleave = plan->n_keys <= (1ull << 40) && per_rank * (uint64_t)plan->n_slots * 8 <= (16ull << 30) ? 1 : 0;
    }
    run_scan(ctx, *plan, &st);
    const int ns = plan->n_slots;
    st.shard_merge = mode == ShardMerge::Gather ? "gather" : mode == ShardMerge::Reduce ? "reduce" : "sparse";
    if (mode != ShardMerge::Sparse) st.merged = true;  // the table will hold other ranks' groups too
    if (mode == ShardMerge::Gather) {
      if (st.dense_blob_bytes != blob) throw std::runtime_error("internal: partial blob size differs from the plan's");
      void* gathered = take_scratch(ctx, blob * comm_world_);
      if (!gathered) gathered = dev_alloc(blob * comm_world_);
      turn.acquire();
      nccl_all_gather_bytes(comm_, st.table, gathered, blob, s);
      SDO_TRACE("turn %lld gather queued", (long long)turn_no);
      turn.release();
      if (comm_rank_ == root) {
        launch_merge_gathered((const unsigned long long*)gathered, comm_world_, st.dense_blob_bytes, st.dense_table_bytes, st.table_rows * (uint64_t)ns, ns, st.params, st.table, s);
        st.launches += 1;
        r = finalize(ctx, *plan, &st, t0);
      } else {
        SDO_CUDA(cudaStreamSynchronize(s));
      }
    } else if (mode == ShardMerge::Reduce) {
      if (st.dense_blob_bytes != blob) throw std::runtime_error("internal: partial blob size differs from the plan's");
      unsigned long long* tr = (unsigned long long*)dev_alloc(std::max<uint64_t>(cells, 1) * 8);
      launch_slot_transpose(st.table, tr, plan->n_keys, ns, true, s);
      turn.acquire();
      nccl_group_start();
      for (int x = 0; x < ns; x++) {
        const int op = plan->slot_op[x];
        nccl_reduce(comm_, tr + (uint64_t)x * plan->n_keys, tr + (uint64_t)x * plan->n_keys, plan->n_keys, op == SLOT_SUM_F64 ? NcclType::F64 : NcclType::I64,
                    op == SLOT_MIN_I64 ? NcclOp::Min : op == SLOT_MAX_I64 ? NcclOp::Max : NcclOp::Sum, root, s);
      }
      if (plan->n_hll) nccl_reduce(comm_, st.hll, st.hll, plan->n_keys * (uint64_t)plan->n_hll * HLL_M, NcclType::U8, NcclOp::Max, root, s);
      nccl_group_end();
      turn.release();
      st.launches += 2;
      if (comm_rank_ == root) {
        launch_slot_transpose(tr, st.table, plan->n_keys, ns, false, s);
        r = finalize(ctx, *plan, &st, t0);
      } else {
        SDO_CUDA(cudaStreamSynchronize(s));
      }
    } else {
      // ---- sparse: rows, split by owner, exchanged, merged ----
      const int W = comm_world_;
      if (W > 64) throw Unsupported("the sparse merge handles at most 64 ranks");
      // phase marks on the stream (scan | rows ready | split by owner | exchanged | merged | top-k): reported in the stats
      cudaEvent_t mark[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
      int n_marks = 0;
      auto mark_here = [&] { if (n_marks < 6) { SDO_CUDA(cudaEventCreate(&mark[n_marks])); SDO_CUDA(cudaEventRecord(mark[n_marks++], s)); } };
      struct MarkGuard { cudaEvent_t* m; ~MarkGuard() { for (int i = 0; i < 6; i++) if (m[i]) cudaEventDestroy(m[i]); } } mark_guard{mark};
      auto phases = [&] {
        std::string o = ",\"shardPhasesMs\":{";
        static const char* names[] = {"rows", "split", "exchange", "merge", "topk"};
        SDO_CUDA(cudaEventSynchronize(mark[n_marks - 1]));
        for (int i = 1; i < n_marks; i++) {
          float ms = 0;
          cudaEventElapsedTime(&ms, mark[i - 1], mark[i]);
          char b[64];
          snprintf(b, sizeof b, "%s\"%s\":%.4f", i > 1 ? "," : "", names[i], ms);
          o += b;
        }
        return o + "}";
      };
      SDO_CUDA(cudaEventRecord(ctx->ev[5], s));
      const bool emit = st.strategy == STRAT_EMIT;
      const int nsr = emit ? ns - 1 : ns;  // words per exchanged row (raw rows do not carry their count of 1)
      if (!emit && !st.rows_valid) compact_rows(ctx, *plan, &st);
      const uint64_t n = st.n_rows;
      mark_here();  // [0] rows ready
      // top K by one aggregate: K candidates per share, gathered, finished on root
      int tk_slot = -1, tk_f64 = 0, tk_desc = 1;
      int64_t k_want = 0;
      const bool want_topk = topk_request(*plan, &tk_slot, &tk_f64, &tk_desc, &k_want);
      unsigned long long* counts = (unsigned long long*)dev_alloc((2 * 64 + 64 * 64) * 8);  // [0,64) counts, [64,128) cursors, then every rank's counts
      unsigned long long* all_counts = counts + 128;
      SDO_CUDA(cudaMemsetAsync(counts, 0, 2 * 64 * 8, s));
      std::vector<unsigned long long> h((size_t)64 * W);
      unsigned long long send_off[65] = {0}, recv_off[65] = {0};
      std::vector<RowSegment> segments;  // what this rank merges: its own rows for itself + what it received
      if (emit && st.emit_regions) {
        // the scan already wrote every row into its owner's region: the regions are the send buffers
        SDO_CUDA(cudaMemcpyAsync(counts, st.emit_cursors_dev, (size_t)W * 8, cudaMemcpyDeviceToDevice, s));
        turn.acquire();
        nccl_all_gather_bytes(comm_, counts, all_counts, 64 * 8, s);
        SDO_CUDA(cudaMemcpyAsync(h.data(), all_counts, h.size() * 8, cudaMemcpyDeviceToHost, s));
        SDO_CUDA(cudaStreamSynchronize(s));
        st.d2h_bytes += h.size() * 8;
        for (int d = 0; d < W; d++)
          if (h[(size_t)comm_rank_ * 64 + d] != st.emit_counts[d]) throw std::runtime_error("internal: emitted row counts changed under the exchange");
        for (int src = 0; src < W; src++) recv_off[src + 1] = recv_off[src] + (src == comm_rank_ ? 0 : h[(size_t)src * 64 + comm_rank_]);
        const uint64_t m = recv_off[W];
        mark_here();  // [1] split by owner (nothing left to do)
        unsigned long long* rk = (unsigned long long*)dev_alloc(std::max<uint64_t>(m, 1) * 8);
        unsigned long long* rs = (unsigned long long*)dev_alloc(std::max<uint64_t>(m * nsr, 1) * 8);
        uint64_t sent = 0;
        nccl_group_start();
        for (int p = 0; p < W; p++) {
          if (p == comm_rank_) continue;
          const uint64_t sn = st.emit_counts[p], rn = recv_off[p + 1] - recv_off[p];
          if (sn) {
            nccl_send_bytes(comm_, st.rows_keys + (uint64_t)p * st.emit_cap, sn * 8, p, s);
            if (nsr) nccl_send_bytes(comm_, st.rows_slots + (uint64_t)p * st.emit_cap * nsr, sn * nsr * 8, p, s);
            sent += sn;
          }
          if (rn) {
            nccl_recv_bytes(comm_, rk + recv_off[p], rn * 8, p, s);
            if (nsr) nccl_recv_bytes(comm_, rs + recv_off[p] * nsr, rn * nsr * 8, p, s);
          }
        }
        nccl_group_end();
        mark_here();  // [2] exchanged
        SDO_TRACE("turn %lld exchange queued (%llu rows in)", (long long)turn_no, (unsigned long long)m);
        st.launches += 1;
        st.exchange_bytes = sent * (uint64_t)(nsr + 1) * 8;
        segments.push_back({st.rows_keys + (uint64_t)comm_rank_ * st.emit_cap, st.rows_slots + (uint64_t)comm_rank_ * st.emit_cap * nsr, st.emit_counts[comm_rank_]});
        segments.push_back({rk, rs, m});
      } else {
        // groups (or raw rows in one stream): count per owner, scatter into runs, exchange the runs
        launch_partition_rows(st.rows_keys, st.rows_slots, n, nsr, W, nullptr, nullptr, counts, nullptr, s);
        turn.acquire();
        nccl_all_gather_bytes(comm_, counts, all_counts, 64 * 8, s);
        SDO_CUDA(cudaMemcpyAsync(h.data(), all_counts, h.size() * 8, cudaMemcpyDeviceToHost, s));
        SDO_CUDA(cudaStreamSynchronize(s));
        st.d2h_bytes += h.size() * 8;
        unsigned long long cur[64] = {0};
        for (int d = 0; d < W; d++) send_off[d + 1] = send_off[d] + h[(size_t)comm_rank_ * 64 + d];   // rows this rank sends to d
        for (int src = 0; src < W; src++) recv_off[src + 1] = recv_off[src] + h[(size_t)src * 64 + comm_rank_];  // rows src sends here
        if (send_off[W] != n) throw std::runtime_error("internal: partition counts do not add up");
        for (int d = 0; d < W; d++) cur[d] = send_off[d];
        SDO_CUDA(cudaMemcpyAsync(counts + 64, cur, (size_t)W * 8, cudaMemcpyHostToDevice, s));
        unsigned long long* pk = (unsigned long long*)dev_alloc(std::max<uint64_t>(n, 1) * 8);
        unsigned long long* ps = (unsigned long long*)dev_alloc(std::max<uint64_t>(n * nsr, 1) * 8);
        launch_partition_rows(st.rows_keys, st.rows_slots, n, nsr, W, pk, ps, counts, counts + 64, s);
        const uint64_t m = recv_off[W];
        mark_here();  // [1] split by owner
        unsigned long long* rk = (unsigned long long*)dev_alloc(std::max<uint64_t>(m, 1) * 8);
        unsigned long long* rs = (unsigned long long*)dev_alloc(std::max<uint64_t>(m * nsr, 1) * 8);
        nccl_group_start();
        for (int p = 0; p < W; p++) {
          const uint64_t sn = send_off[p + 1] - send_off[p], rn = recv_off[p + 1] - recv_off[p];
          if (p == comm_rank_) {
            if (sn != rn) throw std::runtime_error("internal: self exchange counts differ");
            if (sn) {
              SDO_CUDA(cudaMemcpyAsync(rk + recv_off[p], pk + send_off[p], sn * 8, cudaMemcpyDeviceToDevice, s));
              if (nsr) SDO_CUDA(cudaMemcpyAsync(rs + recv_off[p] * nsr, ps + send_off[p] * nsr, sn * nsr * 8, cudaMemcpyDeviceToDevice, s));
            }
            continue;
          }
          if (sn) {
            nccl_send_bytes(comm_, pk + send_off[p], sn * 8, p, s);
            if (nsr) nccl_send_bytes(comm_, ps + send_off[p] * nsr, sn * nsr * 8, p, s);
          }
          if (rn) {
            nccl_recv_bytes(comm_, rk + recv_off[p], rn * 8, p, s);
            if (nsr) nccl_recv_bytes(comm_, rs + recv_off[p] * nsr, rn * nsr * 8, p, s);
          }
        }
        nccl_group_end();
        mark_here();  // [2] exchanged
        st.launches += 3;
        st.exchange_bytes = (n - (send_off[comm_rank_ + 1] - send_off[comm_rank_])) * (uint64_t)(nsr + 1) * 8;
        segments.push_back({rk, rs, m});
      }
      // This rank's share: rows, or (top-K) the merged table as it stands.  Keys dealt out round-robin (key mod world) leave
      // every rank a slice it can index directly -- no key storage, no probing; hashed keys go into a hash table.
      uint64_t rows_in = 0;
      for (auto& g : segments) rows_in += g.n;
      const uint64_t share_keys = (plan->n_keys + (uint64_t)W - 1) / (uint64_t)W;
      // (a direct table only pays when the rows fill a fair part of it: a filtered query with few survivors would spend its
      // time initialising and searching gigabytes of empty rows)
      if (emit && st.emit_regions && st.emit_interleave && rows_in * 8 >= share_keys)
        merge_received_rows_dense(ctx, *plan, &st, segments, (uint64_t)comm_rank_, (uint64_t)W, (plan->n_keys + (uint64_t)W - 1) / (uint64_t)W, emit, /*compact=*/!want_topk);
      else
        merge_received_rows(ctx, *plan, &st, segments, emit, /*compact=*/!want_topk);
      mark_here();  // [3] merged
      if (!want_topk) turn.release();  // the exchange was this query's last collective
      if (want_topk) {
        const int k = (int)k_want;
        unsigned long long* cand_k = (unsigned long long*)dev_alloc((size_t)k * 8 * W);
        unsigned long long* cand_s = (unsigned long long*)dev_alloc((size_t)k * ns * 8 * W);
        unsigned long long* mine_k = (unsigned long long*)dev_alloc((size_t)k * 8);
        unsigned long long* mine_s = (unsigned long long*)dev_alloc((size_t)k * ns * 8);
        bool selected = false;
        if (st.table_rows > 65536) {
          // radix select over the merged hash table in place; ties beyond the candidate list fall back to the tournament
          void* scratch = dev_alloc(select_scratch_bytes(k));
          launch_select_topk(select_view(st, ns, tk_slot, tk_f64, tk_desc), ns, k, scratch, mine_k, mine_s, s);
          st.launches += 16;
          SelState sel;
          SDO_CUDA(cudaMemcpyAsync(&sel, scratch, sizeof sel, cudaMemcpyDeviceToHost, s));
          SDO_CUDA(cudaStreamSynchronize(s));
          selected = !sel.overflow;
        }
        if (!selected) compact_rows(ctx, *plan, &st);
        if (selected) {
        } else if (st.n_rows > (uint64_t)k) {
          const size_t se = topk_scratch_elems(st.n_rows, k);
          unsigned long long *sa = (unsigned long long*)dev_alloc(se * 8), *sb = (unsigned long long*)dev_alloc(se * 8), *oi = (unsigned long long*)dev_alloc((size_t)k * 8);
          launch_topk(st.rows_keys, st.rows_slots, st.n_rows, ns, tk_slot, tk_f64, tk_desc, k, sa, sb, oi, s);
          launch_gather_rows(st.rows_keys, st.rows_slots, oi, (uint64_t)k, ns, mine_k, mine_s, s);
          st.launches += 2;
        } else {
          // fewer groups than K: all of them, padded with the empty key
          SDO_CUDA(cudaMemsetAsync(mine_k, 0xff, (size_t)k * 8, s));
          SDO_CUDA(cudaMemsetAsync(mine_s, 0, (size_t)k * ns * 8, s));
          if (st.n_rows) {
            SDO_CUDA(cudaMemcpyAsync(mine_k, st.rows_keys, st.n_rows * 8, cudaMemcpyDeviceToDevice, s));
            SDO_CUDA(cudaMemcpyAsync(mine_s, st.rows_slots, st.n_rows * ns * 8, cudaMemcpyDeviceToDevice, s));
          }
        }
        mark_here();  // [4] top-k of this share
        st.extra_stats = phases();
        nccl_group_start();
        nccl_all_gather_bytes(comm_, mine_k, cand_k, (size_t)k * 8, s);
        nccl_all_gather_bytes(comm_, mine_s, cand_s, (size_t)k * ns * 8, s);
        nccl_group_end();
        turn.release();
        if (comm_rank_ == root) {
          // the candidates replace the share: finalize orders them on the host and cuts at K
          if (st.rows_keys) SDO_CUDA(cudaFreeAsync(st.rows_keys, s));
          if (st.rows_slots) SDO_CUDA(cudaFreeAsync(st.rows_slots, s));
          st.rows_keys = st.rows_slots = nullptr;
          st.rows_valid = true;
          SDO_CUDA(cudaMallocAsync((void**)&st.rows_keys, (size_t)k * 8 * W, s));
          SDO_CUDA(cudaMallocAsync((void**)&st.rows_slots, (size_t)k * ns * 8 * W, s));
          SDO_CUDA(cudaMemcpyAsync(st.rows_keys, cand_k, (size_t)k * 8 * W, cudaMemcpyDeviceToDevice, s));
          SDO_CUDA(cudaMemcpyAsync(st.rows_slots, cand_s, (size_t)k * ns * 8 * W, cudaMemcpyDeviceToDevice, s));
          st.n_rows = (uint64_t)k * W;
          r = finalize(ctx, *plan, &st, t0);
        } else {
          SDO_CUDA(cudaStreamSynchronize(s));
        }
      } else {
        st.extra_stats = phases();
        r = finalize(ctx, *plan, &st, t0);  // this rank's share of the groups
      }
    }
    for (void* p : temps) cudaFreeAsync(p, s);
    temps.clear();
  } catch (...) {
    cudaStreamSynchronize(ctx->stream);
    for (void* p : temps) cudaFreeAsync(p, ctx->stream);
    free_state(ctx, &st);
    release_ctx(ctx);
    throw;
  }
  free_state(ctx, &st);
  release_ctx(ctx);
  return r;
}

// The same into a direct-indexed table over this rank's interleaved share of the key space: key = row * key_mul + key_off.
void Engine::merge_received_rows_dense(QueryCtx* ctx, const Plan& plan, ScanState* st, const std::vector<RowSegment>& segs, uint64_t key_off, uint64_t key_mul,
                                       uint64_t n_keys, bool implicit_count, bool compact) {
  cudaStream_t s = ctx->stream;
  const int ns = plan.n_slots;
  // (the segments may live in st->rows_*: those are released by the caller's free_state, not here)
  unsigned long long* old_keys = st->rows_keys;
  unsigned long long* old_slots = st->rows_slots;
  for (void* q : {(void*)st->table, (void*)st->hkeys})
    if (q && !((char*)q >= ctx->scratch && (char*)q < ctx->scratch + ctx->scratch_bytes)) SDO_CUDA(cudaFreeAsync(q, s));
  st->table = nullptr; st->hkeys = nullptr;
  st->table_preinit = false;
  SDO_CUDA(cudaMallocAsync((void**)&st->table, std::max<uint64_t>(n_keys, 1) * (uint64_t)ns * 8, s));
  st->params.table = st->table;
  st->params.n_keys = n_keys;
  st->table_rows = n_keys;
  st->dense_key_off = key_off;
  st->dense_key_mul = key_mul;
  st->strategy = STRAT_GLOBAL;
  SDO_CUDA(cudaMemsetAsync(ctx->counters, 0, COUNTER_BYTES, s));
  launch_fill_slots(st->table, n_keys, ns, nullptr, st->params, s);
  for (auto& g : segs) launch_merge_rows_dense(g.keys, g.slots, g.n, st->params, key_off, key_mul, implicit_count, s);
  st->launches += 1 + (int)segs.size();
  st->rows_keys = nullptr; st->rows_slots = nullptr;
  if (old_keys) SDO_CUDA(cudaFreeAsync(old_keys, s));   // stream-ordered: after the merges that read them
  if (old_slots) SDO_CUDA(cudaFreeAsync(old_slots, s));
  st->rows_valid = false;
  st->emit_regions = false;
  st->merged = true;
  if (compact) compact_rows(ctx, plan, st);
}

// Which single aggregate a query wants the top K groups of (TopN with a numeric metric, or a limitSpec over one
// aggregate column; K <= 1024, no having): the shapes a device top-k serves.
bool Engine::topk_request(const Plan& plan, int* slot, int* is_f64, int* desc_out, int64_t* k_want) {
  const QuerySpec& q = plan.spec;
  if (q.has_having || plan.n_hll) return false;
  std::string col;
  bool desc = true;
  int64_t k = 0;
  if (q.type == QuerySpec::TopN && q.metric.kind == TopNMetric::Numeric) { col = q.metric.metric; desc = !q.metric.inverted; k = q.threshold; }
  else if (q.type == QuerySpec::GroupBy && q.limit.present && q.limit.columns.size() == 1) { col = q.limit.columns[0].dimension; desc = !q.limit.columns[0].ascending; k = q.limit.limit; }
  if (col.empty() || k <= 0 || k > 1024) return false;
  for (auto& a : plan.aggs)
    if (a.name == col && a.hll_index < 0 && !a.const_times_count) {
      *slot = a.slot;
      *is_f64 = (a.kind == AggregationSpec::DoubleSum) ? 1 : 0;  // min / max doubles are already order-preserving ints
      *desc_out = desc ? 1 : 0;
      *k_want = k;
      return true;
    }
  return false;
}

// Fold (key, slots) rows -- this rank's own and those received from other ranks -- into one group per key: a fresh hash
// table sized for the rows, then (compact) compaction back to rows.  Leaves st->rows_* = this rank's share.
// implicit_count: raw rows of n_slots - 1 words, each counting 1.
void Engine::merge_received_rows(QueryCtx* ctx, const Plan& plan, ScanState* st, const std::vector<RowSegment>& segs, bool implicit_count, bool compact) {
  cudaStream_t s = ctx->stream;
  const int ns = plan.n_slots;
  unsigned long long* old_keys = st->rows_keys;
  unsigned long long* old_slots = st->rows_slots;
  for (void* q : {(void*)st->table, (void*)st->hkeys})
    if (q && !((char*)q >= ctx->scratch && (char*)q < ctx->scratch + ctx->scratch_bytes)) SDO_CUDA(cudaFreeAsync(q, s));
  st->table = nullptr; st->hkeys = nullptr;
  st->table_preinit = false;
  uint64_t n = 0;
  for (auto& g : segs) n += g.n;
  const uint64_t cap = pow2_ceil(std::max<uint64_t>(1024, n + n / 2));
  const uint32_t stride = (uint32_t)pow2_ceil((uint64_t)ns + 1);
  SDO_CUDA(cudaMallocAsync((void**)&st->table, cap * (uint64_t)stride * 8, s));
  st->params.hkeys = st->table;
  st->params.hmask = cap - 1;
  st->params.hstride = stride;
  st->params.table = st->table;
  st->table_rows = cap;
  st->strategy = STRAT_HASH;
  SDO_CUDA(cudaMemsetAsync(ctx->counters, 0, COUNTER_BYTES, s));
  launch_fill_hash(st->table, cap, stride, ns, st->params, s);
  for (auto& g : segs) launch_merge_rows(g.keys, g.slots, g.n, st->params, s, implicit_count);
  st->launches += 1 + (int)segs.size();
  st->rows_keys = nullptr; st->rows_slots = nullptr;
  if (old_keys) SDO_CUDA(cudaFreeAsync(old_keys, s));   // stream-ordered: after the merges that may read them
  if (old_slots) SDO_CUDA(cudaFreeAsync(old_slots, s));
  st->rows_valid = false;
  st->emit_regions = false;
  st->merged = true;
  if (compact) compact_rows(ctx, plan, st);
}

Partial* Engine::partial(const char* json, size_t len) {
  QuerySpec spec = parse_query_spec(json, len);
  auto ds = store_.get(spec.data_source);
  std::unique_ptr<Partial> p(new Partial());
  p->engine = this;
  p->ds = ds;
  ds->mu.lock_shared();
  try {
    p->plan = make_plan(spec, ds, opt_);
    if (device_ < 0) throw NoDevice("this engine was created without a CUDA device (planning only); there is no CPU execution path");
    p->ctx = acquire_ctx();
    // every rank must produce the same table layout even when its own shard has no matching rows:
    // scan zero units, so the tables are initialised and nothing is added
    if (p->plan.always_false) p->plan = scan_nothing_variant(p->plan);
    run_scan(p->ctx, p->plan, &p->st);
    if (p->st.strategy == STRAT_HASH) compact_rows(p->ctx, p->plan, &p->st);
    p->st.merged = true;  // from here on the caller may fold other ranks' state into the table
    for (int i = 0; i < p->plan.n_slots; i++) p->slot_ops.push_back(p->plan.slot_op[i]);
  } catch (...) {
    if (p->ctx) {
      cudaStreamSynchronize(p->ctx->stream);
      free_state(p->ctx, &p->st);
      release_ctx(p->ctx);
    }
    ds->mu.unlock_shared();
    throw;
  }
  return p.release();
}

void Engine::partial_view(Partial* p, void* out_) {
  sdo_partial_view* v = (sdo_partial_view*)out_;
  memset(v, 0, sizeof *v);
  v->n_slots = p->plan.n_slots;
  v->slot_ops = p->slot_ops.data();
  v->n_hll = p->plan.n_hll;
  if (p->st.strategy == STRAT_HASH || p->st.rows_valid) {
    v->dense = 0;
    v->n_keys = (int64_t)p->st.n_rows;
    v->keys = p->st.rows_keys;
    v->slots = p->st.rows_slots;
    v->hll = p->st.rows_hll;
  } else {
    v->dense = 1;
    v->n_keys = (int64_t)p->st.table_rows;
    v->keys = nullptr;
    v->slots = p->st.table;
    v->hll = p->st.hll;
    v->blob_bytes = (int64_t)p->st.dense_blob_bytes;
  }
}

void Engine::partial_export(Partial* p, void* dst_dev, int64_t dst_bytes) {
  if (p->st.strategy == STRAT_HASH || p->st.rows_valid || !p->st.dense_blob_bytes) throw Invalid("sdo_partial_export applies to dense partials only");
  if (dst_bytes < (int64_t)p->st.dense_blob_bytes) throw Invalid("sdo_partial_export: destination too small");
  SDO_CUDA(cudaSetDevice(device_));
  SDO_CUDA(cudaMemcpyAsync(dst_dev, p->st.table, p->st.dense_blob_bytes, cudaMemcpyDeviceToDevice, p->ctx->stream));
}

void Engine::partial_merge_gathered(Partial* p, const void* gathered_dev, int n_ranks) {
  if (p->st.strategy == STRAT_HASH || p->st.rows_valid || !p->st.dense_blob_bytes) throw Invalid("sdo_partial_merge_gathered applies to dense partials only");
  if (n_ranks < 1) throw Invalid("n_ranks out of range");
  SDO_CUDA(cudaSetDevice(device_));
  const uint64_t cells = p->st.table_rows * (uint64_t)p->plan.n_slots;
  launch_merge_gathered((const unsigned long long*)gathered_dev, n_ranks, p->st.dense_blob_bytes, p->st.dense_table_bytes, cells, p->plan.n_slots,
                        p->st.params, p->st.table, p->ctx->stream);
  p->st.launches += 1;
}

void Engine::partial_partition(Partial* p, int n_ranks, int64_t* counts_out) {
  if (!p->st.rows_valid) throw Invalid("sdo_partial_partition applies to sparse partials only");
  if (n_ranks < 1 || n_ranks > 64) throw Invalid("n_ranks out of range");
  SDO_CUDA(cudaSetDevice(device_));
  cudaStream_t s = p->ctx->stream;
  const int ns = p->plan.n_slots;
  const uint64_t n = p->st.n_rows;
  unsigned long long *counts = nullptr, *ok = nullptr, *os = nullptr;
  SDO_CUDA(cudaMallocAsync((void**)&counts, 2 * 64 * 8, s));
  SDO_CUDA(cudaMemsetAsync(counts, 0, 2 * 64 * 8, s));
  launch_partition_rows(p->st.rows_keys, p->st.rows_slots, n, ns, n_ranks, nullptr, nullptr, counts, nullptr, s);
  unsigned long long h[64] = {0}, cur[64] = {0};
  SDO_CUDA(cudaMemcpyAsync(h, counts, (size_t)n_ranks * 8, cudaMemcpyDeviceToHost, s));
  SDO_CUDA(cudaStreamSynchronize(s));
  unsigned long long run = 0;
  for (int r = 0; r < n_ranks; r++) { cur[r] = run; run += h[r]; counts_out[r] = (int64_t)h[r]; }
  SDO_CUDA(cudaMemcpyAsync(counts + 64, cur, (size_t)n_ranks * 8, cudaMemcpyHostToDevice, s));
  SDO_CUDA(cudaMallocAsync((void**)&ok, std::max<uint64_t>(n, 1) * 8, s));
  SDO_CUDA(cudaMallocAsync((void**)&os, std::max<uint64_t>(n, 1) * ns * 8, s));
  launch_partition_rows(p->st.rows_keys, p->st.rows_slots, n, ns, n_ranks, ok, os, counts, counts + 64, s);
  SDO_CUDA(cudaStreamSynchronize(s));
  SDO_CUDA(cudaFreeAsync(p->st.rows_keys, s));
  SDO_CUDA(cudaFreeAsync(p->st.rows_slots, s));
  SDO_CUDA(cudaFreeAsync(counts, s));
  p->st.rows_keys = ok;
  p->st.rows_slots = os;
}

void Engine::partial_merge_rows(Partial* p, const void* keys, const void* slots, int64_t n) {
  if (!p->st.rows_valid) throw Invalid("sdo_partial_merge_rows applies to sparse partials only");
  if (p->plan.n_hll) throw Unsupported("cardinality aggregations cannot be merged on the sparse path");
  SDO_CUDA(cudaSetDevice(device_));
  merge_received_rows(p->ctx, p->plan, &p->st, {{(const unsigned long long*)keys, (const unsigned long long*)slots, (uint64_t)n}}, false, true);
}

Result* Engine::partial_finish(Partial* p) {
  SDO_CUDA(cudaSetDevice(device_));
  Result* r = finalize(p->ctx, p->plan, &p->st, now_ms());
  p->finished = true;
  return r;
}

void Engine::partial_free(Partial* p) {
  if (!p) return;
  if (p->ctx) {
    cudaSetDevice(device_);
    cudaStreamSynchronize(p->ctx->stream);
    free_state(p->ctx, &p->st);
    release_ctx(p->ctx);
  }
  p->ds->mu.unlock_shared();
  delete p;
}

}  // namespace sdo
