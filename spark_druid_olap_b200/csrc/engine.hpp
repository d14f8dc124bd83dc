// Engine: owns the segment store and runs query plans on one GPU.
// Replaces the execution behind DruidRDD.compute (DruidRDD.scala:152-242): where the reference
// posts the query to a broker/historical and streams JSON rows back, the engine launches the scan
// kernels on its HBM-resident segments and returns a columnar result.
#pragma once
#include <atomic>
#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <shared_mutex>
#include <string>
#include <vector>

#include "comm.hpp"
#include "plan.hpp"
#include "store.hpp"

namespace sdo {

// Pinned host buffers for large result columns, recycled across queries (cudaMallocHost is slow).
class PinnedPool {
 public:
  ~PinnedPool();
  void* get(size_t bytes, size_t* got);
  void put(void* p, size_t bytes);

 private:
  std::mutex mu_;
  std::vector<std::pair<void*, size_t>> free_;
  size_t pooled_ = 0;
};

struct ResultColumn {
  std::string name;
  int type = 0;  // sdo_res_type
  std::vector<int64_t> i64;
  std::vector<double> f64;
  std::vector<uint32_t> ids;
  std::shared_ptr<Dictionary> dict;  // STRING
  std::vector<uint8_t> hll;          // cardinality columns: n_rows * 512 registers, gathered on first request from:
  const uint8_t* hll_rows = nullptr; // every group's sketches as they came off the device (Result::hll_backing)
  std::vector<uint64_t> hll_pick;    // per result row: index of its sketch in hll_rows
  void* raw = nullptr;               // large results: column values in a pooled pinned buffer instead of the vectors
  size_t raw_bytes = 0;
};

struct Result {
  int64_t n_rows = 0;
  std::vector<ResultColumn> columns;
  std::string stats;
  PinnedPool* pool = nullptr;  // owner of the raw buffers (the engine must outlive its results)
  void* hll_backing = nullptr;  // pooled pinned buffer behind the columns' hll_rows
  size_t hll_backing_bytes = 0;
  ~Result() {
    for (auto& c : columns)
      if (c.raw && pool) pool->put(c.raw, c.raw_bytes);
    if (hll_backing && pool) pool->put(hll_backing, hll_backing_bytes);
  }
};

// one in-flight query's device resources
struct QueryCtx {
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  void* pinned_in = nullptr;
  size_t pinned_in_bytes = 0;
  void* pinned_out = nullptr;
  size_t pinned_out_bytes = 0;
  unsigned int* counters = nullptr;  // device: [0] chunk ticket [1] overflow; +8 bytes: u64 compaction counter; zero between queries
  // Device scratch of the context for a query's small buffers (plan blob, small dense table, lane partials): every
  // stream-ordered allocation saved before the scan launch is a few microseconds of a 100-microsecond query.
  char* scratch = nullptr;
  size_t scratch_bytes = 0, scratch_used = 0;
};

// device state of a finished scan (before finalisation / merging)
struct ScanState {
  ScanParams params{};
  int strategy = 0;
  uint64_t table_rows = 0;   // dense: n_keys; hash: capacity
  unsigned long long* table = nullptr;
  unsigned long long* hkeys = nullptr;
  uint8_t* hll = nullptr;
  bool hll_in_table = false;      // dense: registers live behind the table in the same allocation
  size_t dense_blob_bytes = 0;    // dense: bytes of table (+ registers), contiguous
  size_t dense_table_bytes = 0;   // dense: bytes of the table part (256-aligned)
  void* blob = nullptr;      // per-query device blob (spans, coarse index, segptr, bitsets)
  unsigned long long* partials = nullptr;
  int launches = 0;
  size_t h2d_bytes = 0, d2h_bytes = 0;
  float scan_ms = 0;
  double plan_ms = 0;         // host: parse + plan
  bool jit = false;           // the scan ran on a run-time specialised kernel
  double jit_compile_ms = 0;  // > 0 when this query paid for the compilation
  // sparse rows after compaction / merging (multi-GPU)
  unsigned long long* rows_keys = nullptr;
  unsigned long long* rows_slots = nullptr;
  uint8_t* rows_hll = nullptr;
  bool table_preinit = false;     // small dense table: lives in the context scratch, initial image uploaded with the plan blob
  bool merged = false;            // the table may hold other ranks' groups too: local row counts no longer bound it
  double* hll_stats = nullptr;    // per sketch (sum of 2^-register, zero registers), computed on the device
  uint64_t n_rows = 0;
  bool rows_valid = false;
  uint64_t dense_key_mul = 1;         // ... and the key step between its rows (interleaved shares: the world size)
  // STRAT_EMIT output split by owner: region o = rows_keys + o * emit_cap (emit_counts[o] rows)
  int emit_owners = 0, emit_interleave = 0;
  bool emit_regions = false;
  bool emit_full_regions = false;     // retry after an owner's region overflowed: every region holds all rows
  uint64_t emit_cap = 0;
  std::vector<uint64_t> emit_counts;
  unsigned long long* emit_cursors_dev = nullptr;
  uint64_t dense_key_off = 0;         // dense table that holds a rank's slice of the key space: key of its row 0
  bool counted = false;               // this scan is in the engine's in-flight count
  int others_at_launch = 0;           // scans already in flight when this one was launched
  const char* shard_merge = nullptr;  // sharded execution: which merge ran (gather / reduce / sparse)
  std::string extra_stats;            // more stats JSON members (", \"name\": value" ...)
  uint64_t exchange_bytes = 0;        // sparse merge: bytes of rows this rank sent to other ranks
};

class Engine;
struct Partial {
  Engine* engine = nullptr;
  std::shared_ptr<DataSource> ds;
  Plan plan;
  ScanState st;
  QueryCtx* ctx = nullptr;
  std::vector<int> slot_ops;
  bool finished = false;
};

class Engine {
 public:
  explicit Engine(int device);
  ~Engine();
  Store& store() { return store_; }
  int device() const { return device_; }
  Result* execute(const char* json, size_t len);
  std::string explain(const char* json, size_t len);
  void cancel(const std::string& query_id);
  // sharded execution with the merge inside the library (dense key spaces, small tables)
  void comm_init(const void* id, int rank, int world);
  // turn >= 0: this query's position in the communicator's collective order (same on every rank, consecutive from 0)
  Result* execute_sharded(const char* json, size_t len, int root, int64_t turn = -1);
  int64_t next_turn() {
    std::lock_guard<std::mutex> lk(turn_mu_);
    return next_turn_;
  }
  Partial* partial(const char* json, size_t len);
  void partial_view(Partial* p, void* view_out);
  void partial_export(Partial* p, void* dst_dev, int64_t dst_bytes);
  void partial_merge_gathered(Partial* p, const void* gathered_dev, int n_ranks);
  void partial_partition(Partial* p, int n_ranks, int64_t* counts_out);
  void partial_merge_rows(Partial* p, const void* keys, const void* slots, int64_t n);
  Result* partial_finish(Partial* p);
  void partial_free(Partial* p);

 private:
  int device_;
  int sm_count_ = 148;
  std::atomic<int> scans_in_flight_{0};
  Store store_;
  PlanOptions opt_;
  std::mutex ctx_mu_;
  std::vector<QueryCtx*> free_ctx_;
  std::mutex cancel_mu_;
  std::map<std::string, double> cancelled_;      // query id -> time of the cancel request (ms)
  std::map<std::string, int> running_;           // query id -> queries running under it
  std::map<std::string, double> running_since_;  // query id -> start of the oldest of them
  struct RunningQuery {  // scope guard: registers a query id for the duration of its execution
    RunningQuery(Engine* e, const std::string& qid);
    ~RunningQuery();
    Engine* e_;
    std::string qid_;
  };
  std::mutex lut_mu_;
  PinnedPool pinned_pool_;
  void* comm_ = nullptr;  // ncclComm_t of this engine's rank (sdo_engine_comm_init)
  int comm_rank_ = 0, comm_world_ = 1;
  // collectives of one communicator must be issued in the same order on every rank: one query at a time holds the turn
  struct CommTurn;
  std::mutex turn_mu_;
  std::condition_variable turn_cv_;
  bool turn_busy_ = false;
  int64_t next_turn_ = 0;
  std::mutex comm_mu_;
  static bool topk_request(const Plan& plan, int* slot, int* is_f64, int* desc, int64_t* k_want);
  struct RowSegment { const unsigned long long* keys; const unsigned long long* slots; uint64_t n; };
  void merge_received_rows(QueryCtx* ctx, const Plan& plan, ScanState* st, const std::vector<RowSegment>& segs, bool implicit_count, bool compact);
  void merge_received_rows_dense(QueryCtx* ctx, const Plan& plan, ScanState* st, const std::vector<RowSegment>& segs, uint64_t key_off, uint64_t key_mul,
                                 uint64_t n_keys, bool implicit_count, bool compact);
  Result* finalize_columnar(QueryCtx* ctx, const Plan& plan, ScanState* st, double t_start_ms);

  QueryCtx* acquire_ctx();
  void release_ctx(QueryCtx* c);
  void ensure_pinned(void** p, size_t* cur, size_t need);
  void check_cancel(const std::string& qid);
  void prepare_params(QueryCtx* ctx, const Plan& plan, ScanState* st);
  void* take_scratch(QueryCtx* ctx, size_t bytes);  // nullptr when it does not fit
  void run_scan(QueryCtx* ctx, const Plan& plan, ScanState* st);
  Result* execute_select(const QuerySpec& spec, double t_start_ms);
  // search_out (optional): a search query has no plan of its own -- its spec is handed back and nullptr returned
  std::shared_ptr<const Plan> plan_for(const char* json, size_t len, std::shared_lock<std::shared_mutex>* lock_out, QuerySpec* search_out = nullptr);
  Result* execute_search(const QuerySpec& spec, double t_start_ms);
  Result* execute_plan(const Plan& plan, double t_start_ms);
  std::mutex plan_mu_;
  std::map<std::string, std::shared_ptr<const Plan>> plan_cache_;
  void compact_rows(QueryCtx* ctx, const Plan& plan, ScanState* st);
  Result* finalize(QueryCtx* ctx, const Plan& plan, ScanState* st, double t_start_ms);
  void free_state(QueryCtx* ctx, ScanState* st);
  const uint16_t* ensure_hll_lut(DataSource& ds, int di, cudaStream_t st);
  const void* ensure_hll_column(DataSource& ds, int di, Segment& sg, cudaStream_t st);
};

double hll_estimate(const uint8_t* regs);
double hll_estimate_from(double zinv, double zeros);
bool hll_needs_bias_tables(double zinv, double zeros);

}  // namespace sdo
