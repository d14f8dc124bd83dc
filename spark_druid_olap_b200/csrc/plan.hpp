// Query planner: QuerySpec + segment store -> device scan plan.
// Mirrors what a Druid historical decides per query (segment pruning by interval, bitmap-index
// filter evaluation on dictionary values, key layout) and what the reference's cost model
// decides about output cardinality (sources/druid/DruidQueryCostModel.scala:691-716: product of
// dimension cardinalities capped by row count) -- here that number picks the aggregation strategy.
#pragma once
#include <atomic>
#include <memory>
#include <string>
#include <vector>

#include "kernels.hpp"
#include "spec.hpp"
#include "store.hpp"

namespace sdo {

// One slot of a segptr row.  kind 0 = dimension ids, 1 = metric, 2 = the dimension's precomputed HLL++ sketch column
// (uint16 per row: register index << 6 | rho of the row's value), built on first use like a Druid hyperUnique metric.
struct PlanColumn { bool is_metric; int index; int width; bool is_hll = false; };

struct PlanDim {
  std::string output_name;
  bool is_time = false;   // derived from __time (timeFormat extraction)
  int ds_dim = -1;        // datasource dimension index (non-time)
  int time_fmt = -1;      // index into TimeKey::formatted (time)
  uint64_t card = 1;
  uint64_t stride = 1;
  // extraction function on a dimension: rows group on remap[source id], an id of `derived` (sorted, like a datasource
  // dictionary); both are built per query from the dictionary, not from rows
  std::shared_ptr<Dictionary> derived;
  std::vector<uint32_t> remap;
};

struct PlanAgg {
  AggregationSpec::Kind kind;
  std::string name;
  bool out_float = false;   // FLOAT64 result column (else INT64)
  int slot = 0;             // 8-byte slot (not for HLL)
  bool const_times_count = false;  // longSum over a constant metric = const * row count
  int64_t const_value = 1;
  bool slot_is_f64_ord = false;    // min/max of doubles kept as order-preserving int64
  int hll_index = -1;
};

struct TimeKey {
  int64_t bucket = INT64_MIN;           // granularity bucket start (ms) or INT64_MIN for "all"
  std::vector<std::string> formatted;   // one per time-format dimension
  bool operator<(const TimeKey& o) const { return bucket != o.bucket ? bucket < o.bucket : formatted < o.formatted; }
  bool operator==(const TimeKey& o) const { return bucket == o.bucket && formatted == o.formatted; }
};

struct HostLeaf {
  DevLeaf leaf;                  // bits pointer filled at upload
  std::vector<uint32_t> bitset;  // BITSET payload
};

// a value several queries of one cached plan may read and set; copies of the plan (per-query variants) share it
struct SharedCount {
  std::shared_ptr<std::atomic<int64_t>> p = std::make_shared<std::atomic<int64_t>>(-1);
  int64_t load() const { return p->load(); }
  void store(int64_t x) const { p->store(x); }
};

struct Plan {
  QuerySpec spec;
  // rows that pass the filter, counted by the first hash-strategy run of this plan (sizes the hash table): later runs of
  // the cached plan skip the count pass.  The plan cache drops a plan when its datasource changes, and the count with it.
  mutable SharedCount passing_rows;
  std::shared_ptr<DataSource> ds;
  uint64_t ds_version = 0;          // datasource version this plan was made against
  std::vector<int> seg_index;       // used segments -> index in ds->segments
  std::vector<DevSpan> spans;       // + sentinel appended at upload
  std::vector<uint32_t> coarse;     // coarse span index (COARSE units per entry)
  uint32_t n_units = 0;
  int64_t rows_scanned = 0;
  std::vector<PlanColumn> columns;
  bool always_false = false;        // filter folded to FALSE or no span: empty result without a launch
  std::vector<HostLeaf> leaves;
  std::vector<uint8_t> op_code, op_arg;
  std::vector<PlanDim> dims;        // in spec order (output order)
  std::vector<DevDim> dev_dims;     // non-time dims
  std::vector<PlanAgg> aggs;        // in spec order
  std::vector<DevAgg> dev_aggs;
  std::vector<DevExpr> exprs;       // argument programs of AGG_*_EXPR aggregations (DevAgg::col indexes this)
  std::vector<int> hll_ds_dim;      // per dev agg with AGG_HLL and HLL_LUT: datasource dim whose LUT is needed (-1 otherwise)
  std::vector<TimeKey> time_keys;   // index = time key id
  bool zero_fill = false;           // timeseries with context.skipEmptyBuckets = false: groups without rows are reported too
  bool has_time_key = false;
  bool has_bucket = false;          // granularity != all
  int n_time_fmt = 0;
  uint64_t n_tk = 1;
  uint64_t tk_stride = 1;           // packed key = time key * tk_stride + dimension part
  bool key_overflow = false;        // key space does not fit 64 bits
  uint64_t n_keys = 1;              // dense key-space size (if !key_overflow)
  int n_slots = 1;
  int n_hll = 0;
  int32_t slot_op[MAX_SLOTS];
  uint64_t slot_init[MAX_SLOTS];
  int strategy = STRAT_LANE;
  std::string strategy_reason;
  int64_t algorithmic_bytes = 0;
  int bytes_per_row = 0;
};

struct PlanOptions {
  uint64_t lane_max_cells = 48;                 // n_keys * n_slots for lane-private accumulators (96 KB of shared memory)
  uint64_t block_max_bytes = 160 * 1024;        // shared-memory table
  uint64_t global_max_keys = 1ull << 26;        // direct-indexed HBM table
  uint64_t global_big_max_keys = 1ull << 31;    // ... and beyond that while the key space stays dense relative to the rows
  uint64_t global_big_max_bytes = 24ull << 30;
  uint64_t hll_max_bytes = 8ull << 30;
  int64_t jit_min_rows = 2000000;               // scans at least this long get a shape-specialised kernel
};

// dictionary the ids of a non-time output dimension refer to; null for integer dimensions (id = value)
inline std::shared_ptr<Dictionary> dim_dictionary(const Plan& p, const PlanDim& pd) {
  if (pd.derived) return pd.derived;
  const DimMeta& dm = p.ds->dims[pd.ds_dim];
  return dm.enc == DimEncoding::Dict ? dm.dict : nullptr;
}

Plan make_plan(const QuerySpec& spec, std::shared_ptr<DataSource> ds, const PlanOptions& opt);
std::string explain_json(const Plan& p);
const char* strategy_name(int s);

}  // namespace sdo
