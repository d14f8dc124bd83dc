// HBM-resident segment store: time-partitioned, dictionary-encoded columns.
// Schema follows the reference index spec (src/test/resources/tpch_index_task.json.template:
// timestamp column, string dimensions, long/double metrics, segmentGranularity MONTH,
// queryGranularity DAY) and is exposed to the planner in the shape of DruidDataSource /
// DruidColumn / DruidSegmentInfo (metadata/DruidDataSource.scala:42-153,
// metadata/DruidMetadataCache.scala:64-76).
//
// Layout decisions (B200-first):
//  * every column of a segment is one contiguous, 256-byte aligned HBM array;
//  * dimension ids are 1, 2 or 4 bytes wide by datasource-wide cardinality; string
//    dictionaries are datasource-global and sorted in UTF-8 byte order (= Spark UTF8String
//    order), so `bound` is an id range and key packing is identical across segments;
//  * integer-typed dimensions are stored directly (id = value), no dictionary;
//  * rows inside a segment are sorted by (truncated) time and the host keeps a run index
//    (distinct time -> first row), so interval pruning and time bucketing become row spans
//    and the scan kernels never read a time column.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <vector>

#include "common.hpp"
#include "json.hpp"

namespace sdo {

struct Dictionary {
  std::vector<int64_t> offsets{0};
  std::string bytes;
  int64_t size() const { return (int64_t)offsets.size() - 1; }
  const char* data(int64_t i) const { return bytes.data() + offsets[i]; }
  size_t len(int64_t i) const { return (size_t)(offsets[i + 1] - offsets[i]); }
  std::string get(int64_t i) const { return std::string(data(i), len(i)); }
  void push(const char* p, size_t n) { bytes.append(p, n); offsets.push_back((int64_t)bytes.size()); }
  // first id whose value is >= s (unsigned byte order)
  int64_t lower_bound(const std::string& s) const;
  int64_t upper_bound(const std::string& s) const;
  int64_t find(const std::string& s) const;  // -1 if absent
};
int utf8_cmp(const char* a, size_t na, const char* b, size_t nb);

enum class DimEncoding { Dict, Direct };

struct DimMeta {
  std::string name;
  DimEncoding enc = DimEncoding::Dict;
  int src_type = 0;          // sdo_src_type
  int width = 1;             // bytes per id in HBM
  int64_t cardinality = 0;   // Dict: dictionary size; Direct: max value + 1
  std::shared_ptr<Dictionary> dict = std::make_shared<Dictionary>();  // Dict only; replaced (never mutated) on merge so results can keep a snapshot
  // lazily built device LUT id -> (register index << 6 | rho) for HLL (Dict dims)
  void* hll_lut = nullptr;
  int64_t hll_lut_card = 0;
};
struct MetricMeta {
  std::string name;
  bool is_float = false;
  bool constant = true;   // every loaded row has the same int64 value (e.g. the rollup `count` = 1)
  int64_t const_value = 0;
  bool seen = false;
};

struct Segment {
  int64_t start_ms = 0, end_ms = 0;  // segment-granularity interval
  int partition = 0;
  int64_t n_rows = 0;
  std::vector<void*> dim_data;     // device, per datasource dim (nullptr if absent)
  std::vector<void*> dim_hll;      // device uint16 sketch column per dim (HLL++ idx << 6 | rho of each row), built on first use
  std::vector<int> dim_width;      // width the data is currently encoded with
  std::vector<void*> metric_data;  // device, 8 bytes per row
  // run index over the sorted, truncated time column
  std::vector<int64_t> time_vals;
  std::vector<int64_t> time_rows;  // size = time_vals.size() + 1
  size_t device_bytes = 0;
};

struct DataSource {
  std::string name;
  std::string segment_granularity = "MONTH";
  std::string query_granularity = "DAY";
  std::vector<DimMeta> dims;
  std::vector<MetricMeta> metrics;
  std::vector<std::unique_ptr<Segment>> segments;  // sorted by (start_ms, partition)
  int64_t n_rows = 0;
  mutable std::shared_mutex mu;  // loads take it exclusively, queries shared
  uint64_t version = 0;          // bumped by every load / dimension change: cached plans of older versions are stale

  int dim_index(const std::string& n) const;
  int metric_index(const std::string& n) const;
};

// One column of an sdo_segment_load call, normalised.
struct LoadColumn {
  std::string name;
  int kind = 0, type = 0, src_type = 0;
  const void* data = nullptr;
  const char* bytes = nullptr;
  const int64_t* dict_offsets = nullptr;
  int64_t n_dict = 0;
};

class Store {
 public:
  explicit Store(int device) : device_(device) {}
  ~Store();
  void create(const std::string& name, const std::string& options_json);
  void drop(const std::string& name);
  void load(const std::string& name, int64_t n_rows, const std::vector<LoadColumn>& cols, int partition);
  std::shared_ptr<DataSource> get(const std::string& name) const;
  std::string metadata_json(const std::string& name) const;
  std::string segments_json(const std::string& name) const;
  void extend_dimension(const std::string& name, const std::string& dim, const Dictionary* values, int64_t min_cardinality);
  std::shared_ptr<Dictionary> dictionary(const std::string& name, const std::string& dim) const;
  // encoded segments + dictionaries + time index to / from one file (restart without re-encoding)
  void save(const std::string& name, const std::string& path) const;
  void restore(const std::string& name, const std::string& path);
  int device() const { return device_; }

 private:
  int device_;
  mutable std::mutex mu_;
  std::map<std::string, std::shared_ptr<DataSource>> sources_;
  void free_segment(Segment& s);
  void reencode_segments(DataSource& ds, int di, const std::vector<uint32_t>* lut, int new_width);
  void merge_dictionary(DataSource& ds, int di, const Dictionary& bd, std::vector<uint32_t>* code_to_id);
};

// Druid-style segment identifier: <datasource>_<interval>_<version>_<partition> (DruidSegmentInfo.identifier)
std::string segment_identifier(const DataSource& ds, const Segment& s);
int64_t truncate_time(const std::string& gran, int64_t t);
int64_t segment_start(const std::string& gran, int64_t t);
int64_t segment_end(const std::string& gran, int64_t start);

}  // namespace sdo
