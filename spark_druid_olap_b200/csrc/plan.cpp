#include "plan.hpp"

#include <algorithm>
#include <cmath>
#include <functional>
#include <map>
#include <numeric>
#include <regex>
#include <set>

namespace sdo {

const char* strategy_name(int s) {
  switch (s) {
    case STRAT_LANE: return "lane";
    case STRAT_BLOCK: return "block";
    case STRAT_GLOBAL: return "global";
    case STRAT_SPAN: return "span";
    case STRAT_HASH: return "hash";
    case STRAT_EMIT: return "emit";
  }
  return "?";
}

namespace {

// ---- filter compilation ---------------------------------------------------------------------
// Filters are evaluated on dictionary values once, on the host, and reach the device as id
// ranges or id bitsets (a Druid historical does the same against its bitmap index).
// Semantics: ProjectFilterTransfom.scala:198-265 (=,<,<=,>,>=), :361-371 (IN), :383-398
// (IS [NOT] NULL -> selector ""), null == "" ; bound compares strings unless alphaNumeric
// (DruidQuerySpec.scala:238-253).
struct CNode {
  enum Kind { Const, Leaf, And, Or, Not } kind = Const;
  bool value = true;
  int leaf = -1;
  std::vector<CNode> kids;
};

struct Compiler {
  const DataSource& ds;
  Plan& plan;
  std::map<std::pair<int, int>, int> col_slot;  // (kind, index) -> slot

  int column_slot(bool is_metric, int index, bool is_hll = false) {
    auto key = std::make_pair(is_hll ? 2 : (is_metric ? 1 : 0), index);
    auto it = col_slot.find(key);
    if (it != col_slot.end()) return it->second;
    int slot = (int)plan.columns.size();
    if (slot >= 250) throw Unsupported("query references too many columns");
    plan.columns.push_back({is_metric, index, is_hll ? 2 : (is_metric ? 8 : ds.dims[index].width), is_hll});
    col_slot[key] = slot;
    return slot;
  }

  static bool canonical_uint(const std::string& s, uint64_t* v) {
    if (s.empty() || s.size() > 10) return false;
    if (s.size() > 1 && s[0] == '0') return false;
    uint64_t x = 0;
    for (char c : s) {
      if (c < '0' || c > '9') return false;
      x = x * 10 + (uint64_t)(c - '0');
    }
    *v = x;
    return true;
  }

  int add_range_leaf(int di, uint64_t lo, uint64_t hi) {
    HostLeaf hl;
    hl.leaf.col = column_slot(false, di);
    hl.leaf.width = ds.dims[di].width;
    hl.leaf.kind = LEAF_RANGE;
    hl.leaf.lo = (uint32_t)lo;
    hl.leaf.hi = (uint32_t)hi;
    hl.leaf.bits = nullptr;
    plan.leaves.push_back(std::move(hl));
    return (int)plan.leaves.size() - 1;
  }
  CNode range_node(int di, uint64_t lo, uint64_t hi) {
    const uint64_t card = (uint64_t)ds.dims[di].cardinality;
    hi = std::min(hi, card);
    CNode n;
    if (lo >= hi) { n.value = false; return n; }
    if (lo == 0 && hi >= card) { n.value = true; return n; }
    n.kind = CNode::Leaf;
    n.leaf = add_range_leaf(di, lo, hi);
    return n;
  }
  CNode bitset_node(int di, std::vector<uint32_t>&& bits, uint64_t n_set, uint64_t first, uint64_t last) {
    const uint64_t card = (uint64_t)ds.dims[di].cardinality;
    CNode n;
    if (n_set == 0) { n.value = false; return n; }
    if (n_set == card) { n.value = true; return n; }
    if (n_set == last - first + 1) return range_node(di, first, last + 1);  // contiguous ids
    HostLeaf hl;
    hl.leaf.col = column_slot(false, di);
    hl.leaf.width = ds.dims[di].width;
    hl.leaf.kind = LEAF_BITSET;
    hl.leaf.lo = hl.leaf.hi = 0;
    hl.leaf.bits = nullptr;
    hl.bitset = std::move(bits);
    plan.leaves.push_back(std::move(hl));
    n.kind = CNode::Leaf;
    n.leaf = (int)plan.leaves.size() - 1;
    return n;
  }
  // evaluate `pred` on the string value of every id
  CNode predicate_node(int di, const std::function<bool(const char*, size_t)>& pred) {
    const DimMeta& dm = ds.dims[di];
    const uint64_t card = (uint64_t)dm.cardinality;
    if (dm.enc == DimEncoding::Direct && card > (1ull << 27))
      throw Unsupported("string predicate on integer dimension '" + dm.name + "' with " + std::to_string(card) + " ids");
    std::vector<uint32_t> bits((card + 31) / 32 + 1, 0);
    uint64_t n_set = 0, first = 0, last = 0;
    char buf[24];
    for (uint64_t id = 0; id < card; id++) {
      bool ok;
      if (dm.enc == DimEncoding::Dict) ok = pred(dm.dict->data((int64_t)id), dm.dict->len((int64_t)id));
      else {
        int n = snprintf(buf, sizeof buf, "%llu", (unsigned long long)id);
        ok = pred(buf, (size_t)n);
      }
      if (ok) {
        bits[id >> 5] |= 1u << (id & 31);
        if (!n_set) first = id;
        last = id;
        n_set++;
      }
    }
    return bitset_node(di, std::move(bits), n_set, first, last);
  }
  static CNode constant(bool v) { CNode n; n.value = v; return n; }

  static bool looks_numeric(const std::string& s, double* v) {
    if (s.empty()) return false;
    char* end = nullptr;
    *v = strtod(s.c_str(), &end);
    return end && *end == 0;
  }
  static bool bound_pred(const FilterNode& f, const char* p, size_t n) {
    std::string s(p, n);
    if (f.alpha_numeric) {
      double v, lo, hi;
      bool vn = looks_numeric(s, &v);
      if (f.has_lower) {
        int c;
        if (vn && looks_numeric(f.lower, &lo)) c = v < lo ? -1 : (v > lo ? 1 : 0);
        else c = utf8_cmp(s.data(), s.size(), f.lower.data(), f.lower.size());
        if (c < 0 || (c == 0 && f.lower_strict)) return false;
      }
      if (f.has_upper) {
        int c;
        if (vn && looks_numeric(f.upper, &hi)) c = v < hi ? -1 : (v > hi ? 1 : 0);
        else c = utf8_cmp(s.data(), s.size(), f.upper.data(), f.upper.size());
        if (c > 0 || (c == 0 && f.upper_strict)) return false;
      }
      return true;
    }
    if (f.has_lower) {
      int c = utf8_cmp(s.data(), s.size(), f.lower.data(), f.lower.size());
      if (c < 0 || (c == 0 && f.lower_strict)) return false;
    }
    if (f.has_upper) {
      int c = utf8_cmp(s.data(), s.size(), f.upper.data(), f.upper.size());
      if (c > 0 || (c == 0 && f.upper_strict)) return false;
    }
    return true;
  }

  CNode compile(const FilterNode& f) {
    switch (f.kind) {
      case FilterNode::True: return constant(true);
      case FilterNode::False: return constant(false);
      case FilterNode::And: case FilterNode::Or: {
        const bool is_and = f.kind == FilterNode::And;
        CNode n;
        n.kind = is_and ? CNode::And : CNode::Or;
        for (auto& c : f.fields) {
          CNode k = compile(c);
          if (k.kind == CNode::Const) {
            if (k.value != is_and) return constant(!is_and);  // FALSE in AND / TRUE in OR
            continue;
          }
          if (k.kind == n.kind) for (auto& g : k.kids) n.kids.push_back(std::move(g));
          else n.kids.push_back(std::move(k));
        }
        if (n.kids.empty()) return constant(is_and);
        if (n.kids.size() == 1) return std::move(n.kids[0]);
        return n;
      }
      case FilterNode::Not: {
        CNode k = compile(f.fields.at(0));
        if (k.kind == CNode::Const) return constant(!k.value);
        if (k.kind == CNode::Not) return std::move(k.kids[0]);
        CNode n;
        n.kind = CNode::Not;
        n.kids.push_back(std::move(k));
        return n;
      }
      default: break;
    }
    // leaf predicates on one dimension
    if (f.dimension == "__time") {
      // Literal(null) comparisons become selector(__time, "") = never true
      // (ProjectFilterTransfom.scala:405; FilterPlanningTest "timestamp1" wraps it in not()).
      if (f.kind == FilterNode::Selector && f.value.empty()) return constant(false);
      throw Unsupported("filters on __time other than the null test; use query intervals");
    }
    int di = ds.dim_index(f.dimension);
    if (di < 0) {
      // Druid treats an unknown column as all-null ("")
      switch (f.kind) {
        case FilterNode::Selector: return constant(f.value.empty());
        case FilterNode::In: return constant(std::find(f.values.begin(), f.values.end(), std::string()) != f.values.end());
        case FilterNode::Bound: return constant(bound_pred(f, "", 0));
        case FilterNode::Contains: return constant(f.value.empty());
        case FilterNode::Javascript: return constant(minijs::truthy(minijs::eval(*f.js->body, {minijs::Value::null()})));
        case FilterNode::Regex: {
          std::regex re;
          try { re = std::regex(f.value, std::regex::ECMAScript); } catch (const std::regex_error&) { throw Invalid("bad regex '" + f.value + "'"); }
          return constant(std::regex_search(std::string(), re));
        }
        default: return constant(false);
      }
    }
    const DimMeta& dm = ds.dims[di];
    const uint64_t card = (uint64_t)dm.cardinality;
    switch (f.kind) {
      case FilterNode::Selector: {
        if (dm.enc == DimEncoding::Dict) {
          int64_t id = dm.dict->find(f.value);
          return id < 0 ? constant(false) : range_node(di, (uint64_t)id, (uint64_t)id + 1);
        }
        uint64_t v;
        if (!canonical_uint(f.value, &v) || v >= card) return constant(false);
        return range_node(di, v, v + 1);
      }
      case FilterNode::In: {
        std::vector<uint64_t> ids;
        for (auto& s : f.values) {
          if (dm.enc == DimEncoding::Dict) {
            int64_t id = dm.dict->find(s);
            if (id >= 0) ids.push_back((uint64_t)id);
          } else {
            uint64_t v;
            if (canonical_uint(s, &v) && v < card) ids.push_back(v);
          }
        }
        std::sort(ids.begin(), ids.end());
        ids.erase(std::unique(ids.begin(), ids.end()), ids.end());
        if (ids.empty()) return constant(false);
        std::vector<uint32_t> bits((card + 31) / 32 + 1, 0);
        for (uint64_t id : ids) bits[id >> 5] |= 1u << (id & 31);
        return bitset_node(di, std::move(bits), ids.size(), ids.front(), ids.back());
      }
      case FilterNode::Bound: {
        if (dm.enc == DimEncoding::Dict && !f.alpha_numeric) {
          uint64_t lo = 0, hi = card;
          if (f.has_lower) lo = (uint64_t)(f.lower_strict ? dm.dict->upper_bound(f.lower) : dm.dict->lower_bound(f.lower));
          if (f.has_upper) hi = (uint64_t)(f.upper_strict ? dm.dict->lower_bound(f.upper) : dm.dict->upper_bound(f.upper));
          return range_node(di, lo, hi);
        }
        if (dm.enc == DimEncoding::Direct && f.alpha_numeric) {
          double l = 0, u = 0;
          bool ln = !f.has_lower || looks_numeric(f.lower, &l), un = !f.has_upper || looks_numeric(f.upper, &u);
          if (ln && un) {
            double lo = 0, hi = (double)card;
            if (f.has_lower) lo = f.lower_strict ? std::floor(l) + 1 : std::ceil(l);
            if (f.has_upper) hi = f.upper_strict ? std::ceil(u) : std::floor(u) + 1;
            lo = std::max(lo, 0.0);
            hi = std::min(std::max(hi, 0.0), (double)card);
            return range_node(di, (uint64_t)lo, (uint64_t)hi);
          }
        }
        const FilterNode* fp = &f;
        return predicate_node(di, [fp](const char* p, size_t n) { return bound_pred(*fp, p, n); });
      }
      case FilterNode::Contains: {
        const std::string& needle = f.value;
        if (f.ignore_case) {
          // search "insensitive_contains": ASCII letters compare without case
          auto lower = [](char c) { return (char)(c >= 'A' && c <= 'Z' ? c + 32 : c); };
          return predicate_node(di, [&needle, lower](const char* p, size_t n) {
            if (needle.empty()) return true;
            return std::search(p, p + n, needle.begin(), needle.end(), [lower](char a, char b) { return lower(a) == lower(b); }) != p + n;
          });
        }
        return predicate_node(di, [&needle](const char* p, size_t n) {
          if (needle.empty()) return true;
          return std::search(p, p + n, needle.begin(), needle.end()) != p + n;
        });
      }
      case FilterNode::Javascript: {
        // Druid hands the function the dimension value, null for the empty value (JavascriptFilterSpec,
        // DruidQuerySpec.scala:218-236); one evaluation per dictionary entry, rows then test an id bitset
        const minijs::Node* body = f.js->body.get();
        return predicate_node(di, [body](const char* p, size_t n) {
          return minijs::truthy(minijs::eval(*body, {n ? minijs::Value::string(std::string(p, n)) : minijs::Value::null()}));
        });
      }
      case FilterNode::Regex: {
        // Druid's regex filter is java.util.regex Pattern.matcher(v).find(); std::regex ECMAScript
        // covers the common subset (classes, anchors, alternation, quantifiers).
        std::regex re;
        try { re = std::regex(f.value, std::regex::ECMAScript | std::regex::optimize); } catch (const std::regex_error&) { throw Invalid("bad regex '" + f.value + "'"); }
        return predicate_node(di, [&re](const char* p, size_t n) { return std::regex_search(p, p + n, re); });
      }
      default: break;
    }
    throw Unsupported("filter kind");
  }

  // stack entries the postfix form of `n` needs when and/or children are emitted deepest first (and/or commute, so the
  // order is free): the device evaluator (scan_kernel.cuh eval_filter16) holds 8 masks, and with at most MAX_LEAVES
  // leaves the deepest-first order never needs more than 4
  static int stack_need(const CNode& n) {
    if (n.kind == CNode::Const || n.kind == CNode::Leaf) return 1;
    if (n.kind == CNode::Not) return stack_need(n.kids[0]);
    std::vector<int> need;
    for (auto& k : n.kids) need.push_back(stack_need(k));
    std::sort(need.begin(), need.end(), std::greater<int>());
    return need.size() > 1 ? std::max(need[0], need[1] + 1) : need[0];
  }
  void emit(const CNode& n, int* depth, int* max_depth) {
    switch (n.kind) {
      case CNode::Const:
        plan.op_code.push_back(n.value ? FOP_TRUE : FOP_FALSE);
        plan.op_arg.push_back(0);
        ++*depth;
        break;
      case CNode::Leaf:
        plan.op_code.push_back(FOP_LEAF);
        plan.op_arg.push_back((uint8_t)n.leaf);
        ++*depth;
        break;
      case CNode::Not:
        emit(n.kids[0], depth, max_depth);
        plan.op_code.push_back(FOP_NOT);
        plan.op_arg.push_back(0);
        break;
      case CNode::And: case CNode::Or: {
        std::vector<const CNode*> kids;
        for (auto& k : n.kids) kids.push_back(&k);
        std::stable_sort(kids.begin(), kids.end(), [](const CNode* a, const CNode* b) { return stack_need(*a) > stack_need(*b); });
        for (size_t i = 0; i < kids.size(); i++) {
          emit(*kids[i], depth, max_depth);
          if (i) {
            plan.op_code.push_back(n.kind == CNode::And ? FOP_AND : FOP_OR);
            plan.op_arg.push_back(0);
            --*depth;
          }
        }
        break;
      }
    }
    *max_depth = std::max(*max_depth, *depth);
  }
};

std::vector<Interval> normalise(std::vector<Interval> v) {
  std::sort(v.begin(), v.end(), [](const Interval& a, const Interval& b) { return a.start < b.start; });
  std::vector<Interval> o;
  for (auto& i : v) {
    if (i.end <= i.start) continue;
    if (!o.empty() && i.start <= o.back().end) o.back().end = std::max(o.back().end, i.end);
    else o.push_back(i);
  }
  return o;
}

// ---- extraction functions on dimensions ----------------------------------------------------------------------------
// Parse `v` by a SimpleDateFormat-style pattern of numeric fields (y M d H m s S), quoted text and literal characters
// (TimeParsingExtractionFunctionSpec.timeFormat, DruidQuerySpec.scala:82-88).  A run of n pattern letters reads n
// digits, or as many as there are when the next pattern character is not a letter.
bool parse_by_pattern(const std::string& pat, const char* v, size_t nv, int64_t* out_ms) {
  int64_t f[7] = {1970, 1, 1, 0, 0, 0, 0};  // y M d H m s S
  size_t i = 0, k = 0;
  while (i < pat.size()) {
    const char c = pat[i];
    if (c == '\'') {
      size_t e = pat.find('\'', i + 1);
      if (e == std::string::npos) throw Invalid("time extraction: unterminated quote in '" + pat + "'");
      const size_t n = e - i - 1;
      if (n == 0) { if (k >= nv || v[k] != '\'') return false; k++; }
      else { if (k + n > nv || memcmp(v + k, pat.data() + i + 1, n) != 0) return false; k += n; }
      i = e + 1;
      continue;
    }
    if (!isalpha((unsigned char)c)) {
      if (k >= nv || v[k] != c) return false;
      i++; k++;
      continue;
    }
    size_t run = 1;
    while (i + run < pat.size() && pat[i + run] == c) run++;
    const char* which = strchr("yMdHmsS", c);
    if (!which) throw Unsupported(std::string("time extraction: pattern letter '") + c + "'");
    const bool next_is_field = i + run < pat.size() && isalpha((unsigned char)pat[i + run]);
    size_t digits = 0;
    int64_t val = 0;
    while (k < nv && isdigit((unsigned char)v[k]) && (!next_is_field || digits < run)) { val = val * 10 + (v[k] - '0'); k++; digits++; }
    if (!digits) return false;
    if (c == 'y' && run == 2) val += val < 70 ? 2000 : 1900;
    f[which - "yMdHmsS"] = val;
    i += run;
  }
  if (k != nv || f[1] < 1 || f[1] > 12 || f[2] < 1 || f[2] > 31 || f[3] > 23 || f[4] > 59 || f[5] > 59) return false;
  *out_ms = days_from_civil(f[0], (int)f[1], (int)f[2]) * MS_DAY + f[3] * 3600000ll + f[4] * 60000ll + f[5] * 1000ll + f[6];
  return true;
}

// the value a dimension's extraction function yields for one dictionary value ("" = null)
struct Extractor {
  const ExtractionFn& fn;
  Zone zone;
  std::regex re;
  explicit Extractor(const ExtractionFn& f) : fn(f) {
    if (fn.kind == ExtractionFn::TimeFormat) zone = load_zone(fn.time_zone);
    if (fn.kind == ExtractionFn::Regex) {
      try { re = std::regex(fn.expr, std::regex::ECMAScript | std::regex::optimize); } catch (const std::regex_error&) { throw Invalid("bad regex '" + fn.expr + "'"); }
    }
  }
  std::string operator()(const char* p, size_t n) const {
    switch (fn.kind) {
      case ExtractionFn::TimeFormat: {
        // Druid reads the dimension value as an ISO-8601 instant; null stays null
        if (!n) return std::string();
        int64_t ms;
        try { ms = parse_iso_instant(std::string(p, n)); } catch (const Invalid&) { throw Invalid("timeFormat extraction: dimension value '" + std::string(p, n) + "' is not an ISO-8601 instant"); }
        return joda_format(fn.format, ms, zone.offset_at(ms));
      }
      case ExtractionFn::TimeParse: {
        int64_t ms;
        if (!n) return std::string();
        if (!parse_by_pattern(fn.parse_format, p, n, &ms)) return std::string(p, n);  // Druid passes unparseable values through
        return joda_format(fn.format, ms, 0);
      }
      case ExtractionFn::Javascript: {
        minijs::Value v = minijs::eval(*fn.js->body, {n ? minijs::Value::string(std::string(p, n)) : minijs::Value::null()});
        return v.kind == minijs::Value::Null ? std::string() : minijs::to_string(v);
      }
      case ExtractionFn::Lookup: {
        for (auto& kv : fn.lookup)
          if (kv.first.size() == n && memcmp(kv.first.data(), p, n) == 0) return kv.second;
        if (fn.retain_missing) return std::string(p, n);
        return fn.has_replace_missing ? fn.replace_missing : std::string();
      }
      case ExtractionFn::Regex: {
        std::cmatch m;
        if (std::regex_search(p, p + n, m, re)) return m.size() > 1 ? m[1].str() : m[0].str();
        return fn.has_replace_missing ? fn.replace_missing : std::string(p, n);
      }
      case ExtractionFn::Substring: {
        if ((size_t)fn.index >= n) return std::string();
        size_t len = fn.length < 0 ? n - (size_t)fn.index : std::min<size_t>((size_t)fn.length, n - (size_t)fn.index);
        return std::string(p + fn.index, len);
      }
      case ExtractionFn::Upper: case ExtractionFn::Lower: {
        std::string o(p, n);  // ASCII letters
        for (char& c : o) c = (char)(fn.kind == ExtractionFn::Upper ? (c >= 'a' && c <= 'z' ? c - 32 : c) : (c >= 'A' && c <= 'Z' ? c + 32 : c));
        return o;
      }
      case ExtractionFn::Strlen: return std::to_string(n);
      default: break;
    }
    return std::string(p, n);
  }
};

void build_extraction(const DimMeta& dm, const ExtractionFn& fn, PlanDim* pd) {
  const uint64_t card = (uint64_t)std::max<int64_t>(1, dm.cardinality);
  if (dm.enc == DimEncoding::Direct && card > (1ull << 27))
    throw Unsupported("extraction function on integer dimension '" + dm.name + "' with " + std::to_string(card) + " ids");
  Extractor ex(fn);
  std::vector<std::string> vals(card);
  char buf[24];
  for (uint64_t id = 0; id < card; id++) {
    if (dm.enc == DimEncoding::Dict) vals[id] = id < (uint64_t)dm.dict->size() ? ex(dm.dict->data((int64_t)id), dm.dict->len((int64_t)id)) : std::string();
    else {
      int n = snprintf(buf, sizeof buf, "%llu", (unsigned long long)id);
      vals[id] = ex(buf, (size_t)n);
    }
  }
  std::vector<uint32_t> order(card);
  std::iota(order.begin(), order.end(), 0u);
  std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
    int c = utf8_cmp(vals[a].data(), vals[a].size(), vals[b].data(), vals[b].size());
    return c ? c < 0 : a < b;
  });
  pd->derived = std::make_shared<Dictionary>();
  pd->remap.resize(card);
  for (uint64_t k = 0; k < card; k++) {
    const std::string& v = vals[order[k]];
    if (k == 0 || v != vals[order[k - 1]]) pd->derived->push(v.data(), v.size());
    pd->remap[order[k]] = (uint32_t)(pd->derived->size() - 1);
  }
  pd->card = (uint64_t)pd->derived->size();
}

}  // namespace

Plan make_plan(const QuerySpec& spec, std::shared_ptr<DataSource> ds_ptr, const PlanOptions& opt) {
  if (spec.type == QuerySpec::Search)
    throw Unsupported("a search query has no single plan: sdo_query_execute runs it as one group-by per search dimension");
  if (spec.type == QuerySpec::Select) throw Unsupported("a select query returns rows, not groups: sdo_query_execute pages through it; there is no aggregation plan");
  Plan plan;
  plan.spec = spec;
  plan.ds = ds_ptr;
  const DataSource& ds = *ds_ptr;
  Compiler cc{ds, plan, {}};

  // ---- dimensions --------------------------------------------------------------------------
  // AggregateTransform.scala:48-96: grouping expressions become DefaultDimensionSpec, or
  // ExtractionDimensionSpec(__time, timeFormat) for date-part expressions on the time column.
  std::vector<std::pair<std::string, Zone>> time_fmts;  // (format, zone)
  for (auto& d : spec.dimensions) {
    PlanDim pd;
    pd.output_name = d.output_name;
    if (d.fn.kind != ExtractionFn::None && d.dimension != "__time") {
      // extraction function on a dimension (q8.json: timeFormat "yyyy" of o_orderdate): a string -> string map applied
      // to the dictionary; rows then group on the id of the mapped value
      pd.ds_dim = ds.dim_index(d.dimension);
      if (pd.ds_dim < 0) throw NotFound("unknown dimension '" + d.dimension + "' in datasource '" + ds.name + "'");
      build_extraction(ds.dims[pd.ds_dim], d.fn, &pd);
      plan.dims.push_back(pd);
      continue;
    }
    if (d.fn.kind != ExtractionFn::None && d.fn.kind != ExtractionFn::TimeFormat)
      throw Unsupported("only the timeFormat extraction applies to __time");
    if (d.fn.kind == ExtractionFn::TimeFormat) {
      pd.is_time = true;
      pd.time_fmt = (int)time_fmts.size();
      time_fmts.emplace_back(d.fn.format, load_zone(d.fn.time_zone));
    } else {
      if (d.dimension == "__time") throw Unsupported("grouping on raw __time; use a granularity or a timeFormat extraction");
      pd.ds_dim = ds.dim_index(d.dimension);
      if (pd.ds_dim < 0) {
        if (ds.metric_index(d.dimension) >= 0) throw Unsupported("grouping on metric column '" + d.dimension + "'");
        throw NotFound("unknown dimension '" + d.dimension + "' in datasource '" + ds.name + "'");
      }
      pd.card = (uint64_t)std::max<int64_t>(1, ds.dims[pd.ds_dim].cardinality);
    }
    plan.dims.push_back(pd);
  }
  plan.n_time_fmt = (int)time_fmts.size();
  plan.has_bucket = spec.granularity.kind != Granularity::All;
  plan.has_time_key = plan.has_bucket || plan.n_time_fmt > 0;

  // ---- filter -------------------------------------------------------------------------------
  if (spec.has_filter) {
    CNode root = cc.compile(spec.filter);
    if (root.kind == CNode::Const) {
      if (!root.value) plan.always_false = true;
      plan.leaves.clear();  // nothing references them
    } else {
      int depth = 0, max_depth = 0;
      cc.emit(root, &depth, &max_depth);
      if (max_depth > FILTER_STACK) throw Unsupported("filter expression nests too deep for the device evaluator");
      if ((int)plan.op_code.size() > MAX_OPS) throw Unsupported("filter has too many operators (" + std::to_string(plan.op_code.size()) + ")");
    }
  }

  // ---- aggregations ---------------------------------------------------------------------------
  // AggregateTransform.scala:331-347 (count), :505-515 (sum/min/max mapping), :454-479 (HLL).
  plan.n_slots = 1;
  plan.slot_op[0] = SLOT_SUM_I64;
  plan.slot_init[0] = 0;
  for (auto& a : spec.aggregations) {
    PlanAgg pa;
    pa.kind = a.kind;
    pa.name = a.name;
    int leaf = -1;
    bool agg_never = false;
    if (a.has_filter) {
      size_t before = plan.leaves.size();
      CNode n = cc.compile(a.filter);
      if (n.kind == CNode::Const) {
        agg_never = !n.value;
      } else if (n.kind == CNode::Leaf) {
        leaf = n.leaf;
      } else {
        plan.leaves.resize(before);
        throw Unsupported("filtered aggregation '" + a.name + "' needs a single-dimension filter");
      }
      if (agg_never) {
        // a leaf that never matches: empty id range on any dimension
        if (ds.dims.empty()) throw Unsupported("filtered aggregation on a datasource without dimensions");
        leaf = cc.add_range_leaf(0, 0, 0);
      }
    }
    auto new_slot = [&](int op, uint64_t init) {
      if (plan.n_slots >= MAX_SLOTS) throw Unsupported("too many aggregations");
      plan.slot_op[plan.n_slots] = op;
      plan.slot_init[plan.n_slots] = init;
      return plan.n_slots++;
    };
    DevAgg da{};
    da.leaf = leaf;
    da.hll_index = -1;
    int hll_dim = -1;
    bool emit_dev = true;
    switch (a.kind) {
      case AggregationSpec::Count:
        if (leaf < 0) { pa.slot = 0; emit_dev = false; }
        else { da.op = AGG_COUNT_IF; da.col = 0; pa.slot = da.slot = new_slot(SLOT_SUM_I64, 0); }
        break;
      case AggregationSpec::LongSum: case AggregationSpec::DoubleSum:
      case AggregationSpec::LongMin: case AggregationSpec::LongMax:
      case AggregationSpec::DoubleMin: case AggregationSpec::DoubleMax: {
        if (a.has_expr) {
          // argument = arithmetic over metric columns (JSAggGenerator.scala:36-62 / `expression`): its postfix program
          // goes to the kernel with parameters resolved to column slots
          if (plan.exprs.size() >= (size_t)MAX_EXPRS) throw Unsupported("more than " + std::to_string(MAX_EXPRS) + " expression aggregations");
          if (a.expr.code.size() > (size_t)MAX_EXPR_OPS || a.expr.consts.size() > (size_t)MAX_EXPR_CONSTS || a.expr.max_depth > EXPR_STACK)
            throw Unsupported("aggregation '" + a.name + "': expression too long for the device evaluator");
          DevExpr ex{};
          ex.n_ops = (int32_t)a.expr.code.size();
          for (size_t k = 0; k < a.expr.consts.size(); k++) ex.consts[k] = a.expr.consts[k];
          for (size_t k = 0; k < a.expr.code.size(); k++) {
            const uint8_t code = a.expr.code[k], arg = a.expr.arg[k];
            if (code == minijs::N_PUSH_PARAM) {
              if (arg >= a.fields.size()) throw Invalid("aggregation '" + a.name + "': expression parameter without a field");
              const std::string& fld = a.fields[arg];
              int mi = ds.metric_index(fld);
              if (mi < 0) {
                if (ds.dim_index(fld) >= 0) throw Unsupported("aggregation '" + a.name + "': dimension '" + fld + "' in an arithmetic expression");
                throw NotFound("unknown metric '" + fld + "' in datasource '" + ds.name + "'");
              }
              ex.code[k] = ds.metrics[mi].is_float ? EX_COL_F64 : EX_COL_I64;
              ex.arg[k] = (uint8_t)cc.column_slot(true, mi);
            } else if (code == minijs::N_PUSH_CONST) {
              ex.code[k] = EX_CONST;
              ex.arg[k] = arg;
            } else {
              static const uint8_t map[] = {0, 0, EX_ADD, EX_SUB, EX_MUL, EX_DIV, EX_NEG, EX_ABS, EX_MIN, EX_MAX, EX_FLOOR, EX_CEIL, EX_SQRT, EX_MOD};
              ex.code[k] = map[code];
              ex.arg[k] = 0;
            }
          }
          pa.out_float = true;
          da.col = (int32_t)plan.exprs.size();
          plan.exprs.push_back(ex);
          if (a.kind == AggregationSpec::DoubleSum) {
            da.op = AGG_SUM_EXPR;
            pa.slot = da.slot = new_slot(SLOT_SUM_F64, 0);
          } else {
            const bool is_min = a.kind == AggregationSpec::DoubleMin;
            da.op = is_min ? AGG_MIN_EXPR : AGG_MAX_EXPR;
            pa.slot_is_f64_ord = true;
            pa.slot = da.slot = new_slot(is_min ? SLOT_MIN_I64 : SLOT_MAX_I64, is_min ? (uint64_t)INT64_MAX : (uint64_t)INT64_MIN);
          }
          break;
        }
        int mi = ds.metric_index(a.field);
        if (mi < 0) throw NotFound("unknown metric '" + a.field + "' in datasource '" + ds.name + "'");
        const MetricMeta& mm = ds.metrics[mi];
        const bool want_f = a.kind == AggregationSpec::DoubleSum || a.kind == AggregationSpec::DoubleMin || a.kind == AggregationSpec::DoubleMax;
        pa.out_float = want_f;
        if (a.kind == AggregationSpec::LongSum && !mm.is_float && mm.constant && leaf < 0) {
          // the rollup `count` metric of a non-rolled-up index is 1 on every row: longSum(count) = count(*)
          // (AggregateTransform.scala:339-346)
          pa.const_times_count = true;
          pa.const_value = mm.const_value;
          pa.slot = 0;
          emit_dev = false;
          break;
        }
        da.col = cc.column_slot(true, mi);
        if (a.kind == AggregationSpec::LongSum) {
          da.op = mm.is_float ? AGG_SUM_I64_F64 : AGG_SUM_I64;
          pa.slot = da.slot = new_slot(SLOT_SUM_I64, 0);
        } else if (a.kind == AggregationSpec::DoubleSum) {
          da.op = mm.is_float ? AGG_SUM_F64 : AGG_SUM_F64_I64;
          pa.slot = da.slot = new_slot(SLOT_SUM_F64, 0);
        } else {
          const bool is_min = a.kind == AggregationSpec::LongMin || a.kind == AggregationSpec::DoubleMin;
          if (want_f) {
            da.op = mm.is_float ? (is_min ? AGG_MIN_F64 : AGG_MAX_F64) : (is_min ? AGG_MIN_F64_I64 : AGG_MAX_F64_I64);
            pa.slot_is_f64_ord = true;
          } else {
            da.op = mm.is_float ? (is_min ? AGG_MIN_I64_F64 : AGG_MAX_I64_F64) : (is_min ? AGG_MIN_I64 : AGG_MAX_I64);
          }
          pa.slot = da.slot = new_slot(is_min ? SLOT_MIN_I64 : SLOT_MAX_I64, is_min ? (uint64_t)INT64_MAX : (uint64_t)INT64_MIN);
        }
        break;
      }
      case AggregationSpec::Cardinality: case AggregationSpec::HyperUnique: {
        const std::string& fld = a.kind == AggregationSpec::Cardinality ? a.fields[0] : a.field;
        int di = ds.dim_index(fld);
        if (di < 0) throw NotFound("cardinality field '" + fld + "' is not a dimension of '" + ds.name + "'");
        const DimMeta& dm = ds.dims[di];
        pa.out_float = true;
        pa.hll_index = da.hll_index = plan.n_hll++;
        da.op = AGG_HLL;
        // the scan reads the dimension's precomputed sketch column (2 bytes per row), not its ids
        da.col = cc.column_slot(false, di, true);
        da.width = 2;
        da.slot = 0;
        da.hll_mode = HLL_PACKED;
        hll_dim = di;
        (void)dm;
        break;
      }
    }
    plan.aggs.push_back(pa);
    if (emit_dev) {
      if ((int)plan.dev_aggs.size() >= MAX_AGGS) throw Unsupported("too many aggregations");
      plan.dev_aggs.push_back(da);
      plan.hll_ds_dim.push_back(hll_dim);
    }
  }
  if ((int)plan.leaves.size() > MAX_LEAVES) throw Unsupported("filter has too many leaf predicates (" + std::to_string(plan.leaves.size()) + ")");
  if (plan.n_hll > 4) throw Unsupported("more than 4 cardinality aggregations");

  // ---- segments, spans, time keys ---------------------------------------------------------------
  // QueryIntervals.scala:40-130 (half-open ms intervals), DruidMetadataCache.scala:94,131-143 (overlap).
  std::vector<Interval> ivs = spec.intervals.empty() ? std::vector<Interval>{{INT64_MIN / 2, INT64_MAX / 2}} : normalise(spec.intervals);
  struct RawSpan { int seg; int64_t rb, re; TimeKey tk; };
  std::vector<RawSpan> raw;
  std::map<TimeKey, uint64_t> tk_ids;
  for (size_t si = 0; si < ds.segments.size(); si++) {
    const Segment& sg = *ds.segments[si];
    if (sg.n_rows == 0) continue;
    if (spec.by_segments) {
      bool hit = false;
      for (auto& r : spec.segments)
        if (r.itvl.start < sg.end_ms && sg.start_ms < r.itvl.end && (r.part < 0 || r.part == sg.partition)) { hit = true; break; }
      if (!hit) continue;
    }
    bool used = false;
    for (auto& iv : ivs) {
      if (iv.end <= sg.time_vals.front() || iv.start > sg.time_vals.back()) continue;
      size_t i0 = std::lower_bound(sg.time_vals.begin(), sg.time_vals.end(), iv.start) - sg.time_vals.begin();
      size_t i1 = std::lower_bound(sg.time_vals.begin(), sg.time_vals.end(), iv.end) - sg.time_vals.begin();
      if (i0 >= i1) continue;
      if (!plan.has_time_key) {
        raw.push_back({(int)plan.seg_index.size(), sg.time_rows[i0], sg.time_rows[i1], TimeKey()});
        used = true;
        continue;
      }
      size_t run0 = i0;
      TimeKey cur;
      for (size_t i = i0; i <= i1; i++) {
        TimeKey tk;
        if (i < i1) {
          int64_t t = sg.time_vals[i];
          if (plan.has_bucket) tk.bucket = bucket_start(spec.granularity, t);
          for (auto& tf : time_fmts) tk.formatted.push_back(joda_format(tf.first, t, tf.second.offset_at(t)));
        }
        if (i == i0) { cur = tk; continue; }
        if (i == i1 || !(tk == cur)) {
          raw.push_back({(int)plan.seg_index.size(), sg.time_rows[run0], sg.time_rows[i], cur});
          tk_ids[cur] = 0;
          run0 = i;
          cur = tk;
        }
      }
      used = true;
    }
    if (used) plan.seg_index.push_back((int)si);
  }
  if (plan.has_time_key && spec.has_time_boundary) {
    // Sharded execution: every rank must number the time keys identically, whatever days its own
    // segments happen to hold.  Enumerate the datasource's time grid (query granularity) over the
    // globally agreed boundary and the query intervals.
    Granularity qg;
    if (ds.query_granularity != "NONE" && ds.query_granularity != "ALL") qg = simple_granularity(ds.query_granularity);
    if (qg.kind != Granularity::Period || qg.months || qg.period_ms <= 0)
      throw Unsupported("sharded time-keyed queries need a fixed-period queryGranularity (e.g. DAY, HOUR) on the datasource");
    const int64_t unit = qg.period_ms;
    const int64_t g0 = bucket_start(qg, spec.tb_min), g1 = bucket_start(qg, spec.tb_max);
    if ((g1 - g0) / unit > (1ll << 22)) throw Unsupported("time grid too fine for a sharded time-keyed query");
    for (int64_t t = g0; t <= g1; t += unit) {
      bool in = false;
      for (auto& iv : ivs)
        if (t >= iv.start && t < iv.end) { in = true; break; }
      if (!in) continue;
      TimeKey tk;
      if (plan.has_bucket) tk.bucket = bucket_start(spec.granularity, t);
      for (auto& tf : time_fmts) tk.formatted.push_back(joda_format(tf.first, t, tf.second.offset_at(t)));
      tk_ids[tk] = 0;
    }
  }
  if (spec.type == QuerySpec::Timeseries && plan.has_bucket && !spec.skip_empty_buckets && plan.n_time_fmt == 0) {
    // context.skipEmptyBuckets = false (Druid's own default for timeseries): every bucket of the granularity between the
    // datasource's first and last timestamp inside the query intervals gets a row, empty ones with the aggregators'
    // identities.  The buckets become time keys here; finalize then keeps groups without rows.
    if (spec.granularity.kind == Granularity::None) throw Unsupported("skipEmptyBuckets = false with granularity none");
    int64_t lo = INT64_MAX, hi = INT64_MIN;
    for (auto& sgp : ds.segments)
      if (!sgp->time_vals.empty()) { lo = std::min(lo, sgp->time_vals.front()); hi = std::max(hi, sgp->time_vals.back()); }
    const Granularity& g = spec.granularity;
    for (auto& iv : ivs) {
      if (lo > hi) break;
      const int64_t a = std::max(iv.start, lo), b = std::min(iv.end, hi + 1);
      for (int64_t t = a; t < b;) {
        TimeKey tk;
        tk.bucket = bucket_start(g, t);
        tk_ids[tk] = 0;
        if (tk_ids.size() > (1u << 22)) throw Unsupported("skipEmptyBuckets = false: more than 4 M buckets in the query intervals");
        int64_t next;
        if (g.kind == Granularity::Duration) next = tk.bucket + g.duration_ms;
        else if (g.months) next = bucket_start(g, tk.bucket + (int64_t)g.months * 31 * MS_DAY + MS_DAY);
        else if (g.period_ms < MS_DAY) next = bucket_start(g, tk.bucket + g.period_ms);
        else next = bucket_start(g, tk.bucket + g.period_ms + 2 * 3600 * 1000);  // local-time aligned: step over a daylight-saving shift
        if (next <= tk.bucket) throw std::runtime_error("internal: granularity buckets do not advance");
        t = next;
      }
    }
    plan.zero_fill = true;
  }
  if (plan.has_time_key) {
    uint64_t id = 0;
    for (auto& kv : tk_ids) { kv.second = id++; plan.time_keys.push_back(kv.first); }
    plan.n_tk = std::max<uint64_t>(1, id);
  }

  // ---- key layout --------------------------------------------------------------------------------
  // Packed key = tk * tk_stride + sum(id_d * stride_d), first dimension most significant, so the
  // numeric order of keys is the lexicographic order of (time key, dim 0, dim 1, ...).
  __uint128_t K = 1;
  for (size_t d = plan.dims.size(); d-- > 0;) {
    PlanDim& pd = plan.dims[d];
    if (pd.is_time) continue;
    pd.stride = (uint64_t)K;
    K *= pd.card;
    if (K >> 63) { plan.key_overflow = true; break; }
  }
  plan.tk_stride = (uint64_t)K;
  K *= plan.n_tk;
  if (K >> 63) plan.key_overflow = true;
  if (plan.key_overflow) throw Unsupported("group key space exceeds 63 bits");
  plan.n_keys = (uint64_t)K;
  for (auto& pd : plan.dims) {
    if (pd.is_time) continue;
    DevDim dd;
    dd.col = cc.column_slot(false, pd.ds_dim);
    dd.width = ds.dims[pd.ds_dim].width;
    dd.stride = pd.stride;
    dd.remap = nullptr;  // run_scan points it at the uploaded copy of pd.remap
    if ((int)plan.dev_dims.size() >= MAX_DIMS) throw Unsupported("more than 8 grouping dimensions");
    plan.dev_dims.push_back(dd);
  }

  // ---- spans -> device records ---------------------------------------------------------------------
  uint64_t units = 0;
  for (auto& r : raw) {
    if (r.re <= r.rb) continue;
    DevSpan s;
    s.seg = (uint32_t)r.seg;
    s.row_begin = (uint32_t)r.rb;
    s.row_end = (uint32_t)r.re;
    s.unit_begin = (uint32_t)units;
    s.key_base = plan.has_time_key ? tk_ids[r.tk] * plan.tk_stride : 0;
    uint64_t al = (uint64_t)r.rb & ~3ull;
    units += ((uint64_t)r.re - al + UNIT_ROWS - 1) / UNIT_ROWS;
    if (units >= 0xFFFFFFF0ull) throw Unsupported("query spans too many rows for one launch");
    plan.rows_scanned += r.re - r.rb;
    plan.spans.push_back(s);
  }
  plan.n_units = (uint32_t)units;
  if (plan.spans.empty()) plan.always_false = true;
  {
    // coarse[i] = span that contains unit i * COARSE (64); one extra entry so [i], [i+1] is always a valid window
    const uint32_t C = 64;
    size_t n = (size_t)(plan.n_units / C) + 2;
    plan.coarse.resize(n);
    size_t s = 0;
    for (size_t i = 0; i < n; i++) {
      uint64_t u = (uint64_t)i * C;
      while (s + 1 < plan.spans.size() && plan.spans[s + 1].unit_begin <= u) s++;
      plan.coarse[i] = (uint32_t)s;
    }
  }

  // ---- algorithmic bytes: rows x sum of encoded widths of referenced columns ---------
  plan.bytes_per_row = 0;
  for (auto& c : plan.columns) plan.bytes_per_row += c.width;
  plan.algorithmic_bytes = plan.rows_scanned * plan.bytes_per_row;

  // ---- strategy ---------------------------------------------------------------------------------
  const uint64_t cells = plan.n_keys * (uint64_t)plan.n_slots;
  const uint64_t hll_bytes = plan.n_hll ? plan.n_keys * (uint64_t)plan.n_hll * 512ull : 0;
  int strat;
  std::string why;
  if (plan.n_keys <= (1ull << 20) && cells <= opt.lane_max_cells) { strat = STRAT_LANE; why = "key space x slots = " + std::to_string(cells) + " cells fit lane-private shared memory"; }
  else if (plan.dev_dims.empty() && plan.n_keys <= (1ull << 31)) {
    // timeseries with more buckets than lane-private memory holds: the key never changes inside a span
    strat = STRAT_SPAN;
    why = "no grouping dimension: private accumulators for the current time bucket, merged into a direct table of " + std::to_string(plan.n_keys) + " rows";
  }
  else if (plan.n_keys <= (1ull << 24) && cells * 8 <= opt.block_max_bytes) { strat = STRAT_BLOCK; why = "table of " + std::to_string(cells * 8) + " bytes fits shared memory"; }
  else if (plan.n_keys <= opt.global_max_keys) { strat = STRAT_GLOBAL; why = "dense key space of " + std::to_string(plan.n_keys) + " keys indexed directly in HBM"; }
  else if (plan.n_hll == 0 && plan.n_keys <= opt.global_big_max_keys && plan.n_keys <= 4 * (uint64_t)std::max<int64_t>(plan.rows_scanned, 1) &&
           cells * 8 <= opt.global_big_max_bytes) {
    // a key space no sparser than a quarter of the scanned rows (group by o_orderkey at SF100: 150 M keys, 600 M rows):
    // direct indexing costs memory HBM has, and runs 2.4x faster than hashing (SF10: 1.75 ms vs 4.29 ms)
    strat = STRAT_GLOBAL;
    why = "key space of " + std::to_string(plan.n_keys) + " keys is dense relative to " + std::to_string(plan.rows_scanned) + " rows: direct table of " + std::to_string(cells * 8) + " bytes in HBM";
  }
  else { strat = STRAT_HASH; why = "key space of " + std::to_string(plan.n_keys) + " keys is sparse: open-addressing hash table"; }
  // "emit" is a request to the sharded layer (exchange raw rows instead of local groups), not a local strategy
  if (!spec.gpu_strategy.empty() && spec.gpu_strategy != "emit") {
    const std::string& g = spec.gpu_strategy;
    int forced = g == "lane" ? STRAT_LANE : g == "block" ? STRAT_BLOCK : g == "global" ? STRAT_GLOBAL : g == "hash" ? STRAT_HASH : g == "span" ? STRAT_SPAN : -1;
    if (forced == STRAT_SPAN && !plan.dev_dims.empty()) throw Invalid("gpuStrategy span: the query has grouping dimensions");
    if (forced < 0) throw Invalid("context.gpuStrategy '" + g + "'");
    if (forced == STRAT_LANE && (cells > 104 || plan.n_keys > (1ull << 20))) throw Invalid("gpuStrategy lane: key space too large");
    if (forced == STRAT_BLOCK && (cells * 8 > 200 * 1024 || plan.n_keys > (1ull << 24))) throw Invalid("gpuStrategy block: key space too large");
    if (forced == STRAT_GLOBAL && plan.n_keys > (1ull << 31)) throw Invalid("gpuStrategy global: key space too large");
    strat = forced;
    why = "forced by context.gpuStrategy";
  }
  if (strat != STRAT_HASH && hll_bytes > opt.hll_max_bytes) throw Unsupported("cardinality aggregation over " + std::to_string(plan.n_keys) + " groups needs too much register memory");
  plan.strategy = strat;
  plan.strategy_reason = why;
  return plan;
}

std::string explain_json(const Plan& p) {
  std::string o = "{";
  o += "\"dataSource\":";
  json_escape(o, p.ds->name);
  o += ",\"queryType\":\"" + std::string(p.spec.type == QuerySpec::GroupBy ? "groupBy" : p.spec.type == QuerySpec::Timeseries ? "timeseries" : "topN") + "\"";
  o += ",\"strategy\":\"" + std::string(strategy_name(p.strategy)) + "\",\"strategyReason\":";
  json_escape(o, p.strategy_reason);
  o += ",\"alwaysFalse\":" + std::string(p.always_false ? "true" : "false");
  o += ",\"segments\":" + std::to_string(p.seg_index.size());
  o += ",\"spans\":" + std::to_string(p.spans.size());
  o += ",\"units\":" + std::to_string(p.n_units);
  o += ",\"rowsScanned\":" + std::to_string((long long)p.rows_scanned);
  o += ",\"bytesPerRow\":" + std::to_string(p.bytes_per_row);
  o += ",\"algorithmicBytes\":" + std::to_string((long long)p.algorithmic_bytes);
  o += ",\"keySpace\":" + std::to_string((unsigned long long)p.n_keys);
  o += ",\"timeKeys\":" + std::to_string((unsigned long long)(p.has_time_key ? p.time_keys.size() : 0));
  o += ",\"slots\":" + std::to_string(p.n_slots) + ",\"hll\":" + std::to_string(p.n_hll);
  o += ",\"columns\":[";
  for (size_t i = 0; i < p.columns.size(); i++) {
    if (i) o += ",";
    const auto& c = p.columns[i];
    o += "{\"name\":";
    json_escape(o, c.is_metric ? p.ds->metrics[c.index].name : (c.is_hll ? "hll16(" + p.ds->dims[c.index].name + ")" : p.ds->dims[c.index].name));
    o += ",\"width\":" + std::to_string(c.width) + "}";
  }
  o += "],\"filter\":{\"ops\":[";
  for (size_t i = 0; i < p.op_code.size(); i++) {
    if (i) o += ",";
    static const char* N[] = {"leaf", "and", "or", "not", "true", "false"};
    o += "\"" + std::string(N[p.op_code[i]]);
    if (p.op_code[i] == FOP_LEAF) o += ":" + std::to_string(p.op_arg[i]);
    o += "\"";
  }
  o += "],\"leaves\":[";
  for (size_t i = 0; i < p.leaves.size(); i++) {
    if (i) o += ",";
    const auto& l = p.leaves[i];
    const auto& c = p.columns[l.leaf.col];
    o += "{\"dimension\":";
    json_escape(o, p.ds->dims[c.index].name);
    if (l.leaf.kind == LEAF_RANGE) {
      o += ",\"kind\":\"range\",\"lo\":" + std::to_string(l.leaf.lo) + ",\"hi\":" + std::to_string(l.leaf.hi);
    } else {
      uint64_t n = 0;
      for (uint32_t w : l.bitset) n += (uint64_t)__builtin_popcount(w);
      o += ",\"kind\":\"bitset\",\"set\":" + std::to_string((unsigned long long)n);
    }
    o += "}";
  }
  o += "]}}";
  return o;
}

}  // namespace sdo
