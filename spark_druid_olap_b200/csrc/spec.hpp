// Query model the engine executes: the GroupBy / Timeseries / TopN vocabulary of
//   /root/reference/src/main/scala/org/sparklinedata/druid/DruidQuerySpec.scala
// parsed from the json4s wire format (jsonClass hints are accepted and ignored;
// dispatch is on "queryType" / "type", which is what Druid itself reads).
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "json.hpp"
#include "minijs.hpp"
#include "timeutil.hpp"

namespace sdo {

// ---- filters (DruidQuerySpec.scala:152-281) -----------------------------------------------
struct FilterNode {
  enum Kind { Selector, Bound, In, Regex, Contains, Javascript, And, Or, Not, True, False } kind = True;
  std::string dimension;
  std::string value;              // Selector / Regex pattern / Contains
  std::vector<std::string> values;  // In
  bool has_lower = false, has_upper = false, lower_strict = false, upper_strict = false, alpha_numeric = false;
  std::string lower, upper;       // Bound
  std::vector<FilterNode> fields; // And / Or / Not(1)
  bool ignore_case = false;       // Contains: ASCII case-insensitive (search query "insensitive_contains")
  std::shared_ptr<const minijs::Function> js;  // Javascript: function(value) -> truthy keeps the row
};

// ---- dimensions (DruidQuerySpec.scala:108-138) --------------------------------------------
// timeFormat applies to __time (key = formatted bucket) and to string dimensions holding ISO dates (q8.json groups by
// the year of o_orderdate this way).  On a dimension every function is a string -> string map, applied once per
// dictionary entry at plan time; rows group on the id of the mapped value.  A null result is the empty value.
struct ExtractionFn {
  enum Kind { None, TimeFormat, TimeParse, Javascript, Lookup, Regex, Substring, Upper, Lower, Strlen } kind = None;
  std::string format;       // TimeFormat: Joda output pattern; TimeParse: resultFormat
  std::string time_zone;    // empty = UTC
  std::string locale;
  std::string parse_format; // TimeParse: input pattern (TimeParsingExtractionFunctionSpec.timeFormat)
  std::shared_ptr<const minijs::Function> js;  // Javascript (JavaScriptExtractionFunctionSpec)
  std::vector<std::pair<std::string, std::string>> lookup;  // Lookup: map entries
  bool retain_missing = false, has_replace_missing = false;
  std::string replace_missing;  // Lookup / Regex: replaceMissingValueWith
  std::string expr;         // Regex: pattern, first capture group (whole match if it has none)
  int64_t index = 0, length = -1;  // Substring
};
struct DimensionSpec {
  std::string dimension, output_name;
  ExtractionFn fn;
};

// ---- aggregations (DruidQuerySpec.scala:283-377) ------------------------------------------
struct AggregationSpec {
  enum Kind { Count, LongSum, DoubleSum, LongMin, LongMax, DoubleMin, DoubleMax, Cardinality, HyperUnique } kind = Count;
  std::string name;
  std::string field;                 // metric (Function / HyperUnique)
  std::vector<std::string> fields;   // Cardinality fieldNames
  bool by_row = true;
  bool has_filter = false;           // FilteredAggregationSpec wraps a selector
  FilterNode filter;
  // DoubleSum / DoubleMin / DoubleMax of an arithmetic expression over metrics instead of one metric: from a
  // JavascriptAggregationSpec (:350-360) or an `expression` attribute.  N_PUSH_PARAM's arg indexes `fields`.
  bool has_expr = false;
  minijs::NumProgram expr;
};

// ---- post-aggregations (DruidQuerySpec.scala:379-430) -------------------------------------
struct PostAggSpec {
  enum Kind { FieldAccess, Constant, Arithmetic, HyperUniqueCardinality } kind = FieldAccess;
  std::string name, field, fn;
  double value = 0;
  std::vector<PostAggSpec> fields;
};

// ---- limit / having / topN metric (DruidQuerySpec.scala:437-506) ---------------------------
struct OrderByColumn { std::string dimension; bool ascending = true; };
struct LimitSpec { bool present = false; int64_t limit = 0; std::vector<OrderByColumn> columns; };
struct HavingNode {
  enum Kind { GreaterThan, EqualTo, LessThan, And, Or, Not } kind = EqualTo;
  std::string aggregation;
  double value = 0;
  std::vector<HavingNode> specs;
};
struct TopNMetric {
  enum Kind { Numeric, Lexicographic, AlphaNumeric } kind = Numeric;
  bool inverted = false;
  std::string metric;         // Numeric
  std::string previous_stop;  // Lexicographic / AlphaNumeric
};

struct SegmentIntervalRef { Interval itvl; std::string version; int part = -1; };

struct QuerySpec {
  enum Type { GroupBy, Timeseries, TopN, Search, Select } type = GroupBy;
  // Select (SelectSpecWithIntervals / ...WithSegmentIntervals, DruidQuerySpec.scala:1021-1098): the rows themselves, a
  // page at a time, in segment order then row order.  paging_ids: segment identifier -> offset among the segment's
  // matching rows where the previous page stopped (PagingSpec, :976-978).
  std::vector<std::string> select_dims, select_metrics;   // empty = all
  std::vector<std::pair<std::string, int64_t>> paging_ids;
  int64_t paging_threshold = 0;
  bool paging_from_next = false;
  // Search (SearchQuerySpec, DruidQuerySpec.scala:870-927): which values of `search_dimensions` occur, with row counts.
  // The planner turns a group-by on one dimension without aggregations into this (QuerySpecTransforms.scala:224-278).
  std::vector<std::string> search_dimensions;       // empty = every dimension
  std::vector<FilterNode> search_terms;             // per term a Contains on a placeholder dimension; all must match
  bool search_sort_strlen = false;                  // sort: lexicographic (default) or strlen
  int64_t search_limit = 1000;
  std::string data_source;
  std::vector<DimensionSpec> dimensions;  // TopN: exactly one
  Granularity granularity;
  bool has_filter = false;
  FilterNode filter;
  std::vector<AggregationSpec> aggregations;
  std::vector<PostAggSpec> post_aggregations;
  LimitSpec limit;
  bool has_having = false;
  HavingNode having;
  int64_t threshold = 0;  // TopN
  TopNMetric metric;      // TopN
  std::vector<Interval> intervals;
  bool by_segments = false;  // intervals given as {"type":"segments", ...}
  std::vector<SegmentIntervalRef> segments;
  std::string query_id;
  bool skip_empty_buckets = true;
  std::string gpu_strategy;  // context.gpuStrategy: force lane|block|global|hash (tests)
  int gpu_ticket_rows = 0;   // context.gpuTicketRows: 512 | 1024 | 2048 | 4096 rows per warp ticket (tests, profiling); 0 = by shape
  std::string gpu_hll_place; // context.gpuHllPlace: force the HLL registers into global|shared memory (tests, profiling)
  std::string gpu_shard_merge;  // context.gpuShardMerge: force the cross-GPU merge of a sharded query: gather | reduce | sparse (tests)
  int gpu_jit = -1;          // context.gpuJit: true / false force the run-time specialised kernel on / off (-1 = engine decides)
  // context.gpuTimeBoundary = [minMs, maxMs]: the datasource's time range over ALL shards, set by the
  // sharded host layer so that every rank numbers time-derived keys the same way
  bool has_time_boundary = false;
  int64_t tb_min = 0, tb_max = 0;
};

// Throws Invalid (malformed spec), Unsupported (javascript / spatial / search / select ...).
QuerySpec parse_query_spec(const Json& j);
QuerySpec parse_query_spec(const char* json, size_t len);

}  // namespace sdo
