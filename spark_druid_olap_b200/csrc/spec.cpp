#include "spec.hpp"

#include <algorithm>
#include <limits>

namespace sdo {
namespace {

std::string req_str(const Json& j, const char* key, const char* what) {
  const Json* v = j.find(key);
  if (!v) throw Invalid(std::string(what) + ": missing '" + key + "'");
  return j.get_str(key);
}

FilterNode parse_filter(const Json& j) {
  if (!j.is_obj()) throw Invalid("filter must be an object");
  std::string t = req_str(j, "type", "filter");
  FilterNode f;
  if (t == "selector") {
    // SelectorFilterSpec, DruidQuerySpec.scala:164-169.  A JSON null value means "" (null == "").
    f.kind = FilterNode::Selector;
    f.dimension = req_str(j, "dimension", "selector filter");
    f.value = j.get_str("value", "");
  } else if (t == "bound") {
    // BoundFilterSpec, DruidQuerySpec.scala:238-253
    f.kind = FilterNode::Bound;
    f.dimension = req_str(j, "dimension", "bound filter");
    if (j.has("lower")) { f.has_lower = true; f.lower = j.get_str("lower"); }
    if (j.has("upper")) { f.has_upper = true; f.upper = j.get_str("upper"); }
    f.lower_strict = j.get_bool("lowerStrict", false);
    f.upper_strict = j.get_bool("upperStrict", false);
    f.alpha_numeric = j.get_bool("alphaNumeric", false);
  } else if (t == "extraction") {
    // ExtractionFilterSpec + InExtractionFnSpec (SQL IN), DruidQuerySpec.scala:97-103,202-208
    const Json* fn = j.find("extractionFn");
    if (!fn || fn->get_str("type") != "lookup") throw Unsupported("extraction filter with a non-lookup extractionFn");
    const Json* lk = fn->find("lookup");
    const Json* mp = lk ? lk->find("map") : nullptr;
    if (!mp || !mp->is_obj()) throw Invalid("extraction filter: lookup.map missing");
    std::string want = j.get_str("value", "true");
    f.kind = FilterNode::In;
    f.dimension = req_str(j, "dimension", "extraction filter");
    for (auto& kv : mp->obj) {
      std::string mapped = kv.second.is_str() ? kv.second.str : (kv.second.is_bool() ? (kv.second.b ? "true" : "false") : "");
      if (mapped == want) f.values.push_back(kv.first);
    }
  } else if (t == "in") {
    f.kind = FilterNode::In;
    f.dimension = req_str(j, "dimension", "in filter");
    const Json* vs = j.find("values");
    if (!vs || !vs->is_arr()) throw Invalid("in filter: 'values' missing");
    for (auto& v : vs->arr) f.values.push_back(v.is_null() ? "" : (v.is_str() ? v.str : (v.is_int ? std::to_string(v.inum) : std::to_string(v.num))));
  } else if (t == "regex") {
    f.kind = FilterNode::Regex;  // RegexFilterSpec, DruidQuerySpec.scala:171-176
    f.dimension = req_str(j, "dimension", "regex filter");
    f.value = req_str(j, "pattern", "regex filter");
  } else if (t == "contains" || t == "search") {
    if (t == "search") {
      const Json* q = j.find("query");
      if (!q || (q->get_str("type") != "contains" && q->get_str("type") != "insensitive_contains")) throw Unsupported("search filter query type");
      f.ignore_case = q->get_str("type") == "insensitive_contains" || !q->get_bool("caseSensitive", true);
      f.value = q->get_str("value");
    } else {
      f.value = req_str(j, "value", "contains filter");  // StringContainsSpec, DruidQuerySpec.scala:178-183
    }
    f.kind = FilterNode::Contains;
    f.dimension = req_str(j, "dimension", "contains filter");
  } else if (t == "and" || t == "or") {
    f.kind = t == "and" ? FilterNode::And : FilterNode::Or;  // LogicalFilterSpec, :185-197
    const Json* fs = j.find("fields");
    if (!fs || !fs->is_arr()) throw Invalid(t + " filter: 'fields' missing");
    for (auto& c : fs->arr) f.fields.push_back(parse_filter(c));
    if (f.fields.empty()) f.kind = t == "and" ? FilterNode::True : FilterNode::False;
  } else if (t == "not") {
    f.kind = FilterNode::Not;  // NotFilterSpec, :199-200
    const Json* c = j.find("field");
    if (!c) throw Invalid("not filter: 'field' missing");
    f.fields.push_back(parse_filter(*c));
  } else if (t == "noop") {
    f.kind = FilterNode::True;
  } else if (t == "javascript") {
    // JavascriptFilterSpec :218-236.  The expression subset of minijs.hpp is evaluated once per dictionary entry at
    // plan time; functions outside it raise Unsupported and the caller falls back to Spark.
    f.kind = FilterNode::Javascript;
    f.dimension = req_str(j, "dimension", "javascript filter");
    f.value = req_str(j, "function", "javascript filter");
    auto fn = std::make_shared<minijs::Function>(minijs::parse_function(f.value));
    if (fn->params.size() != 1) throw Invalid("javascript filter: the function must take exactly one parameter");
    f.js = fn;
  } else if (t == "spatial") {
    // SpatialFilterSpec :270-281 -- out of scope; the caller falls back to Spark
    throw Unsupported("filter type '" + t + "' is not supported by the GPU engine");
  } else {
    throw Unsupported("unknown filter type '" + t + "'");
  }
  return f;
}

DimensionSpec parse_dimension(const Json& j) {
  DimensionSpec d;
  if (j.is_str()) { d.dimension = d.output_name = j.str; return d; }
  if (!j.is_obj()) throw Invalid("dimension spec must be a string or object");
  std::string t = j.get_str("type", "default");
  d.dimension = req_str(j, "dimension", "dimension spec");
  d.output_name = j.get_str("outputName", d.dimension);
  if (t == "default") return d;
  if (t == "extraction") {
    // ExtractionDimensionSpec + TimeFormatExtractionFunctionSpec, DruidQuerySpec.scala:72-80,130-138
    const Json* fn = j.find("extractionFn");
    if (!fn) throw Invalid("extraction dimension: 'extractionFn' missing");
    std::string ft = fn->get_str("type");
    if (ft == "timeFormat") {
      d.fn.kind = ExtractionFn::TimeFormat;
      d.fn.format = fn->get_str("format", "yyyy-MM-dd'T'HH:mm:ss.SSS'Z'");
      d.fn.time_zone = fn->get_str("timeZone", "");
      d.fn.locale = fn->get_str("locale", "");
    } else if (ft == "time") {
      // TimeParsingExtractionFunctionSpec, DruidQuerySpec.scala:82-88
      d.fn.kind = ExtractionFn::TimeParse;
      d.fn.parse_format = req_str(*fn, "timeFormat", "time extraction");
      d.fn.format = req_str(*fn, "resultFormat", "time extraction");
    } else if (ft == "javascript") {
      // JavaScriptExtractionFunctionSpec, DruidQuerySpec.scala:90-96
      d.fn.kind = ExtractionFn::Javascript;
      auto f = std::make_shared<minijs::Function>(minijs::parse_function(req_str(*fn, "function", "javascript extraction")));
      if (f->params.size() != 1) throw Invalid("javascript extraction: the function must take exactly one parameter");
      d.fn.js = f;
    } else if (ft == "lookup") {
      const Json* lk = fn->find("lookup");
      const Json* mp = lk ? lk->find("map") : nullptr;
      if (!mp || !mp->is_obj()) throw Invalid("lookup extraction: lookup.map missing");
      d.fn.kind = ExtractionFn::Lookup;
      for (auto& kv : mp->obj) d.fn.lookup.emplace_back(kv.first, kv.second.is_str() ? kv.second.str : std::string());
      d.fn.retain_missing = fn->get_bool("retainMissingValue", false);
      if (fn->has("replaceMissingValueWith") && !fn->find("replaceMissingValueWith")->is_null()) {
        d.fn.has_replace_missing = true;
        d.fn.replace_missing = fn->get_str("replaceMissingValueWith");
      }
      if (d.fn.retain_missing && d.fn.has_replace_missing) throw Invalid("lookup extraction: retainMissingValue and replaceMissingValueWith exclude each other");
    } else if (ft == "regex") {
      d.fn.kind = ExtractionFn::Regex;
      d.fn.expr = req_str(*fn, "expr", "regex extraction");
      if (fn->get_bool("replaceMissingValue", false)) {
        d.fn.has_replace_missing = true;
        d.fn.replace_missing = fn->get_str("replaceMissingValueWith", "");
      }
    } else if (ft == "substring") {
      d.fn.kind = ExtractionFn::Substring;
      d.fn.index = (int64_t)fn->get_num("index", 0);
      d.fn.length = fn->has("length") && !fn->find("length")->is_null() ? (int64_t)fn->get_num("length", 0) : -1;
      if (d.fn.index < 0 || (fn->has("length") && d.fn.length < 0 && !fn->find("length")->is_null())) throw Invalid("substring extraction: negative index or length");
    } else if (ft == "upper" || ft == "lower" || ft == "strlen") {
      d.fn.kind = ft == "upper" ? ExtractionFn::Upper : ft == "lower" ? ExtractionFn::Lower : ExtractionFn::Strlen;
    } else {
      throw Unsupported("extraction function '" + ft + "' is not supported");
    }
    return d;
  }
  throw Unsupported("dimension spec type '" + t + "' is not supported");
}

AggregationSpec parse_aggregation(const Json& j) {
  if (!j.is_obj()) throw Invalid("aggregation must be an object");
  std::string t = req_str(j, "type", "aggregation");
  AggregationSpec a;
  a.name = j.get_str("name", "");
  if (t == "filtered") {
    // FilteredAggregationSpec, DruidQuerySpec.scala:370-377
    const Json* ag = j.find("aggregator");
    const Json* fl = j.find("filter");
    if (!ag || !fl) throw Invalid("filtered aggregation needs 'filter' and 'aggregator'");
    a = parse_aggregation(*ag);
    if (a.has_filter) throw Unsupported("nested filtered aggregation");
    if (j.has("name")) a.name = j.get_str("name");
    a.has_filter = true;
    a.filter = parse_filter(*fl);
    return a;
  }
  if (a.name.empty()) throw Invalid("aggregation: missing 'name'");
  // FunctionAggregationSpec types, DruidQuerySpec.scala:298-306; AggregateTransform.scala:505-515
  if (t == "count") { a.kind = AggregationSpec::Count; a.field = j.get_str("fieldName", ""); return a; }
  struct { const char* n; AggregationSpec::Kind k; } fn[] = {
      {"longSum", AggregationSpec::LongSum}, {"doubleSum", AggregationSpec::DoubleSum},
      {"longMin", AggregationSpec::LongMin}, {"longMax", AggregationSpec::LongMax},
      {"doubleMin", AggregationSpec::DoubleMin}, {"doubleMax", AggregationSpec::DoubleMax},
      {"min", AggregationSpec::DoubleMin}, {"max", AggregationSpec::DoubleMax}};
  for (auto& f : fn)
    if (t == f.n) {
      a.kind = f.k;
      if (j.has("expression")) {
        // sum / min / max of an arithmetic expression over metrics (the attribute later Druid versions give these
        // aggregators); same numeric program as a javascript aggregator
        if (f.k != AggregationSpec::DoubleSum && f.k != AggregationSpec::DoubleMin && f.k != AggregationSpec::DoubleMax)
          throw Unsupported("aggregation '" + a.name + "': 'expression' is supported on doubleSum / doubleMin / doubleMax");
        minijs::Function fn = minijs::parse_expression(req_str(j, "expression", "aggregation"));
        if (fn.params.empty()) throw Unsupported("aggregation '" + a.name + "': the expression mentions no column");
        a.has_expr = true;
        a.expr = minijs::compile_numeric(*fn.body);
        a.fields = fn.params;
        return a;
      }
      a.field = req_str(j, "fieldName", "aggregation");
      return a;
    }
  if (t == "javascript") {
    // JavascriptAggregationSpec, DruidQuerySpec.scala:350-360; what JSAggGenerator.scala:36-119 emits for
    // SUM / MIN / MAX of an expression: fnAggregate(current, f1, f2, ...) folds an arithmetic expression of the fields
    // into `current`, fnCombine and fnReset are the matching merge and identity.
    const Json* fs = j.find("fieldNames");
    if (!fs || !fs->is_arr() || fs->arr.empty()) throw Invalid("javascript aggregation: 'fieldNames' missing");
    minijs::Function agg = minijs::parse_function(req_str(j, "fnAggregate", "javascript aggregation"));
    if (agg.params.size() != fs->arr.size() + 1) throw Invalid("javascript aggregation '" + a.name + "': fnAggregate must take `current` plus one parameter per field");
    minijs::Fold fold;
    std::unique_ptr<minijs::Node> arg = minijs::split_fold(&agg, &fold);
    a.kind = fold == minijs::Fold::Sum ? AggregationSpec::DoubleSum : fold == minijs::Fold::Min ? AggregationSpec::DoubleMin : AggregationSpec::DoubleMax;
    // fnCombine / fnReset must be the merge and identity of the same fold
    minijs::Function comb = minijs::parse_function(req_str(j, "fnCombine", "javascript aggregation"));
    minijs::Function reset = minijs::parse_function(req_str(j, "fnReset", "javascript aggregation"));
    if (comb.params.size() != 2 || !reset.params.empty()) throw Invalid("javascript aggregation '" + a.name + "': fnCombine takes two partials, fnReset none");
    const double inf = std::numeric_limits<double>::infinity();
    struct Probe { double a, b; } probes[] = {{3.5, -2.25}, {-7.0, 11.0}};
    for (auto& pr : probes) {
      double got = minijs::to_number(minijs::eval(*comb.body, {minijs::Value::number(pr.a), minijs::Value::number(pr.b)}));
      double want = fold == minijs::Fold::Sum ? pr.a + pr.b : fold == minijs::Fold::Min ? std::min(pr.a, pr.b) : std::max(pr.a, pr.b);
      if (got != want) throw Unsupported("javascript aggregation '" + a.name + "': fnCombine is not the merge of fnAggregate's fold");
    }
    double r0 = minijs::to_number(minijs::eval(*reset.body, {}));
    if (r0 != (fold == minijs::Fold::Sum ? 0.0 : fold == minijs::Fold::Min ? inf : -inf))
      throw Unsupported("javascript aggregation '" + a.name + "': fnReset is not the identity of fnAggregate's fold");
    a.has_expr = true;
    a.expr = minijs::compile_numeric(*arg);
    for (auto& f : fs->arr) {
      if (!f.is_str()) throw Invalid("javascript aggregation: fieldNames must be strings");
      a.fields.push_back(f.str);
    }
    for (size_t k = 0; k < a.expr.code.size(); k++)
      if (a.expr.code[k] == minijs::N_PUSH_PARAM) a.expr.arg[k] = (uint8_t)(a.expr.arg[k] - 1);  // parameter 0 is `current`
    return a;
  }
  if (t == "cardinality") {
    // CardinalityAggregationSpec, DruidQuerySpec.scala:316-325
    a.kind = AggregationSpec::Cardinality;
    const Json* fs = j.find("fieldNames");
    if (!fs || !fs->is_arr() || fs->arr.empty()) throw Invalid("cardinality aggregation: 'fieldNames' missing");
    for (auto& f : fs->arr) a.fields.push_back(f.str);
    a.by_row = j.get_bool("byRow", false);
    if (a.fields.size() != 1) throw Unsupported("cardinality over more than one field");
    return a;
  }
  if (t == "hyperUnique") {
    a.kind = AggregationSpec::HyperUnique;  // :327-336
    a.field = req_str(j, "fieldName", "hyperUnique aggregation");
    return a;
  }
  throw Unsupported("aggregation type '" + t + "' is not supported by the GPU engine");
}

PostAggSpec parse_post_agg(const Json& j) {
  if (!j.is_obj()) throw Invalid("postAggregation must be an object");
  std::string t = req_str(j, "type", "postAggregation");
  PostAggSpec p;
  p.name = j.get_str("name", "");
  if (t == "fieldAccess") {
    p.kind = PostAggSpec::FieldAccess;
    p.field = req_str(j, "fieldName", "fieldAccess");
    if (p.name.empty()) p.name = p.field;
  } else if (t == "constant") {
    p.kind = PostAggSpec::Constant;
    p.value = j.get_num("value", 0);
  } else if (t == "hyperUniqueCardinality") {
    p.kind = PostAggSpec::HyperUniqueCardinality;
    p.field = req_str(j, "fieldName", "hyperUniqueCardinality");
    if (p.name.empty()) p.name = p.field;
  } else if (t == "arithmetic") {
    p.kind = PostAggSpec::Arithmetic;
    p.fn = req_str(j, "fn", "arithmetic postAggregation");
    if (p.fn != "+" && p.fn != "-" && p.fn != "*" && p.fn != "/" && p.fn != "quotient") throw Invalid("arithmetic fn '" + p.fn + "'");
    const Json* fs = j.find("fields");
    if (!fs || !fs->is_arr() || fs->arr.empty()) throw Invalid("arithmetic postAggregation: 'fields' missing");
    for (auto& f : fs->arr) p.fields.push_back(parse_post_agg(f));
  } else {
    throw Unsupported("postAggregation type '" + t + "' is not supported");
  }
  return p;
}

HavingNode parse_having(const Json& j) {
  std::string t = req_str(j, "type", "having");
  HavingNode h;
  if (t == "greaterThan" || t == "equalTo" || t == "lessThan") {
    h.kind = t == "greaterThan" ? HavingNode::GreaterThan : (t == "equalTo" ? HavingNode::EqualTo : HavingNode::LessThan);
    h.aggregation = req_str(j, "aggregation", "having");
    h.value = j.get_num("value", 0);
  } else if (t == "and" || t == "or") {
    h.kind = t == "and" ? HavingNode::And : HavingNode::Or;
    const Json* hs = j.find("havingSpecs");
    if (!hs || !hs->is_arr()) throw Invalid("having: 'havingSpecs' missing");
    for (auto& c : hs->arr) h.specs.push_back(parse_having(c));
  } else if (t == "not") {
    h.kind = HavingNode::Not;
    const Json* c = j.find("havingSpec");
    if (!c) throw Invalid("having: 'havingSpec' missing");
    h.specs.push_back(parse_having(*c));
  } else {
    throw Unsupported("having type '" + t + "'");
  }
  return h;
}

Granularity parse_granularity(const Json* j) {
  if (!j) return Granularity();
  if (j->is_str()) return simple_granularity(j->str);
  if (!j->is_obj()) throw Invalid("granularity must be a string or object");
  std::string t = j->get_str("type", "");
  Granularity g;
  if (t == "all") return g;
  if (t == "none") { g.kind = Granularity::None; return g; }
  if (t == "duration") {
    g.kind = Granularity::Duration;  // DurationGranularitySpec, DruidQuerySpec.scala:142
    g.duration_ms = j->get_int("duration", 0);
    if (g.duration_ms <= 0) throw Invalid("duration granularity needs a positive 'duration'");
    if (j->has("origin")) {
      const Json* o = j->find("origin");
      g.origin_ms = o->is_num() ? j->get_int("origin") : parse_iso_instant(o->str);
      g.has_origin = true;
    }
    return g;
  }
  if (t == "period") {
    g.kind = Granularity::Period;  // PeriodGranularitySpec, DruidQuerySpec.scala:144-150
    parse_period(req_str(*j, "period", "period granularity"), &g.months, &g.period_ms);
    g.zone = load_zone(j->get_str("timeZone", "UTC"));
    if (j->has("origin")) { g.origin_ms = parse_iso_instant(j->get_str("origin")); g.has_origin = true; }
    return g;
  }
  throw Invalid("unknown granularity type '" + t + "'");
}

void parse_intervals(const Json* j, QuerySpec* q) {
  if (!j) return;  // default = whole datasource, QueryIntervals.scala:45
  if (j->is_arr()) {
    for (auto& s : j->arr) {
      if (!s.is_str()) throw Invalid("interval must be a string");
      q->intervals.push_back(parse_iso_interval(s.str));
    }
    return;
  }
  if (j->is_str()) { q->intervals.push_back(parse_iso_interval(j->str)); return; }
  if (j->is_obj()) {
    std::string t = j->get_str("type", "");
    if (t == "intervals") { parse_intervals(j->find("intervals"), q); return; }
    if (t == "segments") {
      // SegmentIntervals, DruidQuerySpec.scala:509-541: [{itvl, ver, part}]
      q->by_segments = true;
      const Json* segs = j->find("segments");
      if (!segs || !segs->is_arr()) throw Invalid("segments intervals: 'segments' missing");
      for (auto& s : segs->arr) {
        SegmentIntervalRef r;
        r.itvl = parse_iso_interval(req_str(s, "itvl", "segment interval"));
        r.version = s.get_str("ver", "");
        r.part = (int)s.get_int("part", -1);
        q->segments.push_back(r);
        q->intervals.push_back(r.itvl);
      }
      return;
    }
  }
  throw Invalid("unrecognised 'intervals'");
}

TopNMetric parse_topn_metric(const Json& j) {
  TopNMetric m;
  if (j.is_str()) { m.metric = j.str; return m; }
  std::string t = req_str(j, "type", "topN metric");
  if (t == "numeric") { m.metric = req_str(j, "metric", "numeric topN metric"); return m; }
  if (t == "lexicographic" || t == "alphaNumeric") {
    m.kind = t == "lexicographic" ? TopNMetric::Lexicographic : TopNMetric::AlphaNumeric;
    m.previous_stop = j.get_str("previousStop", "");
    return m;
  }
  if (t == "inverted") {
    const Json* in = j.find("metric");
    if (!in) throw Invalid("inverted topN metric: 'metric' missing");
    m = parse_topn_metric(*in);
    m.inverted = !m.inverted;
    return m;
  }
  throw Unsupported("topN metric type '" + t + "'");
}

}  // namespace

QuerySpec parse_query_spec(const Json& j) {
  if (!j.is_obj()) throw Invalid("query must be a JSON object");
  std::string qt = req_str(j, "queryType", "query");
  QuerySpec q;
  if (qt == "groupBy") q.type = QuerySpec::GroupBy;
  else if (qt == "timeseries") q.type = QuerySpec::Timeseries;
  else if (qt == "topN") q.type = QuerySpec::TopN;
  else if (qt == "search") q.type = QuerySpec::Search;
  else if (qt == "select") q.type = QuerySpec::Select;
  else throw Unsupported("queryType '" + qt + "' is not supported by the GPU engine (groupBy, timeseries, topN, search, select)");
  const Json* ds = j.find("dataSource");
  if (!ds) throw Invalid("query: missing 'dataSource'");
  q.data_source = ds->is_obj() ? ds->get_str("name") : j.get_str("dataSource");
  q.granularity = parse_granularity(j.find("granularity"));
  if (const Json* f = j.find("filter")) { q.has_filter = true; q.filter = parse_filter(*f); }
  if (const Json* a = j.find("aggregations")) {
    if (!a->is_arr()) throw Invalid("'aggregations' must be an array");
    for (auto& x : a->arr) q.aggregations.push_back(parse_aggregation(x));
  }
  if (const Json* a = j.find("postAggregations")) {
    if (!a->is_arr()) throw Invalid("'postAggregations' must be an array");
    for (auto& x : a->arr) {
      q.post_aggregations.push_back(parse_post_agg(x));
      if (q.post_aggregations.back().name.empty()) throw Invalid("postAggregation: missing 'name'");
    }
  }
  parse_intervals(j.find("intervals"), &q);
  if (const Json* c = j.find("context")) {
    q.query_id = c->get_str("queryId", "");
    q.skip_empty_buckets = c->get_bool("skipEmptyBuckets", true);
    q.gpu_strategy = c->get_str("gpuStrategy", "");
    q.gpu_ticket_rows = (int)c->get_num("gpuTicketRows", 0);
    if (q.gpu_ticket_rows != 0 && q.gpu_ticket_rows != 512 && q.gpu_ticket_rows != 1024 && q.gpu_ticket_rows != 2048 && q.gpu_ticket_rows != 4096)
      throw Invalid("context.gpuTicketRows must be 512, 1024, 2048 or 4096");
    q.gpu_hll_place = c->get_str("gpuHllPlace", "");
    if (!q.gpu_hll_place.empty() && q.gpu_hll_place != "global" && q.gpu_hll_place != "shared") throw Invalid("context.gpuHllPlace must be 'global' or 'shared'");
    q.gpu_shard_merge = c->get_str("gpuShardMerge", "");
    if (!q.gpu_shard_merge.empty() && q.gpu_shard_merge != "gather" && q.gpu_shard_merge != "reduce" && q.gpu_shard_merge != "sparse")
      throw Invalid("context.gpuShardMerge must be 'gather', 'reduce' or 'sparse'");
    if (c->has("gpuJit")) q.gpu_jit = c->get_bool("gpuJit", false) ? 1 : 0;
    if (const Json* tb = c->find("gpuTimeBoundary")) {
      if (!tb->is_arr() || tb->arr.size() != 2 || !tb->arr[0].is_num() || !tb->arr[1].is_num()) throw Invalid("context.gpuTimeBoundary must be [minMs, maxMs]");
      q.has_time_boundary = true;
      q.tb_min = tb->arr[0].is_int ? tb->arr[0].inum : (int64_t)tb->arr[0].num;
      q.tb_max = tb->arr[1].is_int ? tb->arr[1].inum : (int64_t)tb->arr[1].num;
      if (q.tb_max < q.tb_min) throw Invalid("context.gpuTimeBoundary: max < min");
    }
  }
  if (q.type == QuerySpec::GroupBy) {
    if (const Json* d = j.find("dimensions")) {
      if (!d->is_arr()) throw Invalid("'dimensions' must be an array");
      for (auto& x : d->arr) q.dimensions.push_back(parse_dimension(x));
    }
    if (const Json* l = j.find("limitSpec")) {
      // LimitSpec / OrderByColumnSpec, DruidQuerySpec.scala:437-456
      q.limit.present = true;
      q.limit.limit = l->get_int("limit", 2147483647);
      if (const Json* cs = l->find("columns")) {
        if (!cs->is_arr()) throw Invalid("limitSpec.columns must be an array");
        for (auto& c : cs->arr) {
          OrderByColumn o;
          if (c.is_str()) o.dimension = c.str;
          else {
            o.dimension = req_str(c, "dimension", "orderBy column");
            std::string dir = c.get_str("direction", "ascending");
            for (auto& ch : dir) ch = (char)tolower((unsigned char)ch);
            o.ascending = !(dir == "descending" || dir == "desc");
          }
          q.limit.columns.push_back(o);
        }
      }
    }
    if (const Json* h = j.find("having")) { q.has_having = true; q.having = parse_having(*h); }
  } else if (q.type == QuerySpec::Select) {
    if (q.granularity.kind != Granularity::All) throw Unsupported("select query with a granularity other than 'all'");
    if (j.get_bool("descending", false)) throw Unsupported("select query with descending = true");
    for (const char* key : {"dimensions", "metrics"}) {
      if (const Json* d = j.find(key)) {
        if (!d->is_arr()) throw Invalid(std::string("'") + key + "' must be an array");
        for (auto& x : d->arr) {
          if (x.is_obj() && x.get_str("type", "default") != "default") throw Unsupported("select query with an extraction dimension");
          (key[0] == 'd' ? q.select_dims : q.select_metrics).push_back(x.is_obj() ? x.get_str("dimension") : x.str);
        }
      }
    }
    const Json* ps = j.find("pagingSpec");
    if (!ps || !ps->is_obj()) throw Invalid("select: missing 'pagingSpec'");
    q.paging_threshold = ps->get_int("threshold", 0);
    if (q.paging_threshold <= 0) throw Invalid("select: pagingSpec.threshold must be positive");
    q.paging_from_next = ps->get_bool("fromNext", false);
    if (const Json* ids = ps->find("pagingIdentifiers")) {
      if (!ids->is_obj()) throw Invalid("select: pagingSpec.pagingIdentifiers must be an object");
      for (auto& kv : ids->obj) {
        if (!kv.second.is_num()) throw Invalid("select: paging offsets must be numbers");
        q.paging_ids.emplace_back(kv.first, kv.second.is_int ? kv.second.inum : (int64_t)kv.second.num);
      }
    }
  } else if (q.type == QuerySpec::Search) {
    if (q.granularity.kind != Granularity::All) throw Unsupported("search query with a granularity other than 'all'");
    if (const Json* d = j.find("searchDimensions")) {
      if (!d->is_arr()) throw Invalid("'searchDimensions' must be an array");
      for (auto& x : d->arr) q.search_dimensions.push_back(x.is_obj() ? x.get_str("dimension") : x.str);
    }
    const Json* sq = j.find("query");
    if (!sq || !sq->is_obj()) throw Invalid("search: missing 'query'");
    // SearchQueryQuerySpec (:543-552): insensitive_contains (what the planner emits, with value ""), contains, fragment
    const std::string st = sq->get_str("type");
    auto term = [&](const std::string& v, bool ignore_case) {
      FilterNode f;
      f.kind = FilterNode::Contains;
      f.value = v;
      f.ignore_case = ignore_case;
      q.search_terms.push_back(f);
    };
    if (st == "insensitive_contains") term(sq->get_str("value", ""), true);
    else if (st == "contains") term(sq->get_str("value", ""), !sq->get_bool("caseSensitive", true));
    else if (st == "fragment") {
      const Json* vs = sq->find("values");
      if (!vs || !vs->is_arr()) throw Invalid("search fragment query: 'values' missing");
      for (auto& v : vs->arr) term(v.str, !sq->get_bool("caseSensitive", false));
    } else throw Unsupported("search query type '" + st + "'");
    q.search_limit = j.get_int("limit", 1000);
    if (q.search_limit < 0) throw Invalid("search: negative 'limit'");
    if (const Json* so = j.find("sort")) {
      const std::string t = so->is_obj() ? so->get_str("type", "lexicographic") : so->str;
      if (t == "strlen") q.search_sort_strlen = true;
      else if (t != "lexicographic") throw Unsupported("search sort '" + t + "'");
    }
  } else if (q.type == QuerySpec::TopN) {
    const Json* d = j.find("dimension");
    if (!d) throw Invalid("topN: missing 'dimension'");
    q.dimensions.push_back(parse_dimension(*d));
    q.threshold = j.get_int("threshold", 0);
    if (q.threshold <= 0) throw Invalid("topN: 'threshold' must be positive");
    const Json* m = j.find("metric");
    if (!m) throw Invalid("topN: missing 'metric'");
    q.metric = parse_topn_metric(*m);
  }
  return q;
}

QuerySpec parse_query_spec(const char* json, size_t len) {
  try {
    return parse_query_spec(json_parse(json, len));
  } catch (const JsonError& e) {
    throw Invalid(e.what());
  }
}

}  // namespace sdo
