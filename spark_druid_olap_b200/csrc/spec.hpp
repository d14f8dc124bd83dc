// Query model the engine executes: the GroupBy / Timeseries / TopN vocabulary of
//   /root/reference/src/main/scala/org/sparklinedata/druid/DruidQuerySpec.scala
// parsed from the json4s wire format (jsonClass hints are accepted and ignored;
// dispatch is on "queryType" / "type", which is what Druid itself reads).
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "json.hpp"
#include "timeutil.hpp"

namespace sdo {

// ---- filters (DruidQuerySpec.scala:152-281) -----------------------------------------------
struct FilterNode {
  enum Kind { Selector, Bound, In, Regex, Contains, And, Or, Not, True, False } kind = True;
  std::string dimension;
  std::string value;              // Selector / Regex pattern / Contains
  std::vector<std::string> values;  // In
  bool has_lower = false, has_upper = false, lower_strict = false, upper_strict = false, alpha_numeric = false;
  std::string lower, upper;       // Bound
  std::vector<FilterNode> fields; // And / Or / Not(1)
};

// ---- dimensions (DruidQuerySpec.scala:108-138) --------------------------------------------
struct ExtractionFn {
  enum Kind { None, TimeFormat } kind = None;
  std::string format;
  std::string time_zone;  // empty = UTC
  std::string locale;
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
  enum Type { GroupBy, Timeseries, TopN } type = GroupBy;
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
