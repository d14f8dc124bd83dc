"""Query corpus: the shapes the reference plans / benchmarks, built with the spec mirror.
Sources: tools/TpchBenchMark.scala:135-292 (benchmark SQL and the specs it plans to),
docs/benchmark/druid/queries/*.json (serialised specs), client/test/TPCHQueries.scala:30-140 (raw
specs with granularity / limit / post-aggregations), FilterCTest.scala:28-50 (in / not / null)."""
from spark_druid_olap_b200.specs import *  # noqa: F401,F403

DS = "tpch"
ALL_TIME = ["1992-01-01T00:00:00.000Z/1999-01-01T00:00:00.000Z"]


def agg(t, name, field):
    return FunctionAggregationSpec(t, name, field)


def ctx(strategy=None, qid="q"):
    return QuerySpecContext(qid, gpuStrategy=strategy) if strategy else None


def q1_style(strategy=None):
    """BASELINE config 1: group by l_returnflag, l_linestatus, l_shipdate <= 1998-09-02, sum/avg/count."""
    return GroupByQuerySpec(
        DS, [DefaultDimensionSpec("l_returnflag"), DefaultDimensionSpec("l_linestatus")], None, None, "all", None,
        [agg("count", "count_order", "count"), agg("doubleSum", "sum_qty", "l_quantity"),
         agg("doubleSum", "sum_base_price", "l_extendedprice"), agg("doubleSum", "sum_disc", "l_discount")],
        [ArithmeticPostAggregationSpec("avg_qty", "/", [FieldAccessPostAggregationSpec("sum_qty"), FieldAccessPostAggregationSpec("count_order")]),
         ArithmeticPostAggregationSpec("avg_price", "/", [FieldAccessPostAggregationSpec("sum_base_price"), FieldAccessPostAggregationSpec("count_order")]),
         ArithmeticPostAggregationSpec("avg_disc", "/", [FieldAccessPostAggregationSpec("sum_disc"), FieldAccessPostAggregationSpec("count_order")])],
        ["1992-01-01T00:00:00.000Z/1998-09-02T00:00:00.001Z"], ctx(strategy))


def basic_agg(strategy=None):
    """docs/benchmark/druid/queries/q1.json / basicAgg.json shape (count, doubleSum, doubleMax, longSum, HLL, post-agg)."""
    return GroupByQuerySpec(
        DS, [DefaultDimensionSpec("l_returnflag"), DefaultDimensionSpec("l_linestatus")], None, None, "all", None,
        [agg("count", "alias-1", "count"), agg("doubleSum", "alias-2", "l_extendedprice"), agg("doubleMax", "alias-3", "ps_supplycost"),
         agg("longSum", "alias-5", "ps_availqty"), agg("longSum", "alias-6", "count"),
         CardinalityAggregationSpec("alias-7", ["o_orderkey"])],
        [ArithmeticPostAggregationSpec("alias-4", "/", [FieldAccessPostAggregationSpec("alias-5"), FieldAccessPostAggregationSpec("alias-6")])],
        ALL_TIME, ctx(strategy))


def ship_date_range(strategy=None):
    """shipDteRange.json: interval (1995-12-01, 1997-09-02], count by flag/status."""
    return GroupByQuerySpec(
        DS, [DefaultDimensionSpec("l_returnflag", "f"), DefaultDimensionSpec("l_linestatus", "s")], None, None, "all", None,
        [agg("count", "count_order", "count")], None, ["1995-12-01T00:00:00.001Z/1997-09-02T00:00:00.001Z"], ctx(strategy))


def _nations():
    return Or(And(SelectorFilterSpec("s_nation", "FRANCE"), SelectorFilterSpec("c_nation", "GERMANY")),
              And(SelectorFilterSpec("c_nation", "FRANCE"), SelectorFilterSpec("s_nation", "GERMANY")))


def proj_filt_range(strategy=None):
    """projFiltRange.json: p_type selector AND nation pairs, ship-date interval, group by s_nation."""
    return GroupByQuerySpec(
        DS, [DefaultDimensionSpec("s_nation")], None, None, "all",
        And(SelectorFilterSpec("p_type", "ECONOMY ANODIZED STEEL"), _nations()),
        [agg("count", "alias-1", "count"), agg("doubleSum", "alias-2", "l_extendedprice"), agg("doubleMax", "alias-3", "ps_supplycost"),
         agg("longSum", "alias-5", "ps_availqty")], None,
        ["1995-12-01T00:00:00.001Z/1997-09-02T00:00:00.001Z"], ctx(strategy))


def q3(limit=None, strategy=None):
    """TPCH Q3 shape (q3.json with the javascript date test as the `bound` the newer planner emits,
    FilterPlanningTest.scala:67-81); limit=10 gives BASELINE config 2's Q3L."""
    ls = LimitSpec(limit, [OrderByColumnSpec("revenue", "descending")]) if limit else None
    return GroupByQuerySpec(
        DS, [DefaultDimensionSpec("o_orderkey"), DefaultDimensionSpec("o_orderdate"), DefaultDimensionSpec("o_shippriority")], ls, None, "all",
        And(SelectorFilterSpec("c_mktsegment", "BUILDING"), BoundFilterSpec("o_orderdate", upper="1995-03-15", upperStrict=True)),
        [agg("doubleSum", "revenue", "l_extendedprice")], None,
        ["1995-03-15T00:00:00.001Z/1999-01-01T00:00:00.000Z"], ctx(strategy))


def _q5_filter():
    return And(SelectorFilterSpec("s_region", "ASIA"),
               BoundFilterSpec("o_orderdate", lower="1994-01-01", lowerStrict=False, upper="1995-01-01", upperStrict=True))


def q5(strategy=None):
    return GroupByQuerySpec(DS, [DefaultDimensionSpec("s_nation")], None, None, "all", _q5_filter(),
                            [agg("doubleSum", "revenue", "l_extendedprice")], None, ALL_TIME, ctx(strategy))


def q5_topn(threshold=10, strategy=None):
    """BASELINE config 2's Q5T: TopN s_nation by revenue."""
    return TopNQuerySpec(DS, ALL_TIME, "all", _q5_filter(), [agg("doubleSum", "revenue", "l_extendedprice")], None,
                         DefaultDimensionSpec("s_nation"), threshold, NumericTopNMetricSpec("revenue"), ctx(strategy))


def q7(strategy=None):
    """q7.json: nation pairs, dims s_nation, c_nation, year of ship date (timeFormat on __time)."""
    return GroupByQuerySpec(
        DS, [DefaultDimensionSpec("s_nation"), DefaultDimensionSpec("c_nation"),
             ExtractionDimensionSpec("__time", "l_year", TimeFormatExtractionFunctionSpec("yyyy"))], None, None, "all", _nations(),
        [agg("doubleSum", "alias-1", "l_extendedprice")], None, ALL_TIME, ctx(strategy))


def q8(strategy=None):
    return GroupByQuerySpec(
        DS, [DefaultDimensionSpec("order_year", "o_year")], None, None, "all",
        And(SelectorFilterSpec("c_region", "AMERICA"), SelectorFilterSpec("p_type", "ECONOMY ANODIZED STEEL"),
            BoundFilterSpec("o_orderdate", lower="1995-01-01", lowerStrict=False, upper="1996-12-31", upperStrict=False)),
        [agg("doubleSum", "price", "l_extendedprice")], None, ALL_TIME, ctx(strategy))


def timeseries(period="P1D", with_sum=True, strategy=None, intervals=None):
    """BASELINE config 3: Timeseries at day / month granularity with HLL count-distinct of o_custkey."""
    aggs = [agg("count", "rows", "count"), CardinalityAggregationSpec("custs", ["o_custkey"])]
    if with_sum:
        aggs.append(agg("doubleSum", "price", "l_extendedprice"))
    return TimeSeriesQuerySpec(DS, intervals or ALL_TIME, PeriodGranularitySpec(period), None, aggs, None, ctx(strategy))


def by_orderkey(strategy=None):
    """BASELINE config 4: group by o_orderkey (~1.5M x SF groups)."""
    return GroupByQuerySpec(DS, [DefaultDimensionSpec("o_orderkey")], None, None, "all", None,
                            [agg("count", "n", "count"), agg("doubleSum", "price", "l_extendedprice")], None, ALL_TIME, ctx(strategy))


def q1_month_grain(strategy=None):
    """TPCHQueries.q1MonthGrain shape: P1M granularity + dims + timeFormat dim + limit ordered by dims."""
    return GroupByQuerySpec(
        DS, [DefaultDimensionSpec("l_returnflag"), DefaultDimensionSpec("l_linestatus"),
             ExtractionDimensionSpec("__time", "month", TimeFormatExtractionFunctionSpec("yyyy-MMM", "UTC"))],
        LimitSpec(10, [OrderByColumnSpec("l_returnflag"), OrderByColumnSpec("l_linestatus"), OrderByColumnSpec("month")]), None,
        PeriodGranularitySpec("P1M"), None,
        [agg("longSum", "count", "count"), agg("doubleSum", "sum_quantity", "l_quantity"), agg("doubleSum", "sum_base_price", "l_extendedprice"),
         agg("doubleSum", "l_discount", "l_discount"), agg("doubleSum", "l_tax", "l_tax"), CardinalityAggregationSpec("count_order", ["o_orderkey"])],
        [ArithmeticPostAggregationSpec("sum_disc_price", "-", [FieldAccessPostAggregationSpec("sum_base_price"), FieldAccessPostAggregationSpec("l_discount")]),
         ArithmeticPostAggregationSpec("avg_qty", "/", [FieldAccessPostAggregationSpec("sum_quantity"), FieldAccessPostAggregationSpec("count")])],
        ["1993-01-01T00:00:00.000/1993-03-01T00:00:00.000"], ctx(strategy))


def filter_cases():
    """(name, FilterSpec) pairs covering selector / bound / in / regex / contains / not / null."""
    return [
        ("in2", ExtractionFilterSpec("c_mktsegment", ["MACHINERY", "HOUSEHOLD"])),                       # FilterPlanningTest "in"
        ("in_many", ExtractionFilterSpec("p_type", ["ECONOMY ANODIZED STEEL", "PROMO PLATED TIN", "SMALL BRUSHED COPPER", "nope"])),
        ("not_in", NotFilterSpec("not", ExtractionFilterSpec("l_shipmode", ["AIR", "RAIL"]))),
        ("is_null", SelectorFilterSpec("l_shipmode", "")),
        ("is_not_null", NotFilterSpec("not", SelectorFilterSpec("l_shipmode", ""))),
        ("ts_null_guard", And(NotFilterSpec("not", SelectorFilterSpec("__time", "")), SelectorFilterSpec("l_returnflag", "R"))),
        ("bound_ge", And(NotFilterSpec("not", SelectorFilterSpec("o_orderdate", "")),                    # FilterPlanningTest "timestamp4"
                         BoundFilterSpec("o_orderdate", "1995-12-30", False, None, None, False))),
        ("bound_between_strict", BoundFilterSpec("s_nation", "CANADA", True, "JAPAN", True, False)),
        ("bound_int_lex", BoundFilterSpec("p_size", "2", False, "35", True, False)),                    # lexicographic on an integer dimension
        ("bound_int_num", BoundFilterSpec("p_size", "2", False, "35", True, True)),                     # alphaNumeric
        ("selector_int", SelectorFilterSpec("l_linenumber", "3")),
        ("selector_int_noncanonical", SelectorFilterSpec("l_linenumber", "03")),
        ("selector_absent_value", SelectorFilterSpec("s_nation", "ATLANTIS")),
        ("regex", RegexFilterSpec("p_type", "^(ECONOMY|PROMO) .* (TIN|STEEL)$")),
        ("contains", StringContainsSpec("p_type", "BRUSHED")),
        ("or_mix", Or(SelectorFilterSpec("l_returnflag", "A"), And(SelectorFilterSpec("l_linestatus", "O"), StringContainsSpec("l_shipmode", "AI")))),
        ("unknown_dim_null", SelectorFilterSpec("no_such_dim", "")),
        ("unknown_dim_value", SelectorFilterSpec("no_such_dim", "x")),
    ] + js_filter_cases()


def search_cases():
    """SearchQuerySpec shapes: what SearchQuerySpecTransform emits for `select distinct dim` / `group by dim order by dim
    limit n` (QuerySpecTransforms.scala:224-278) plus Druid's other search query types."""
    return [
        ("distinct_nation", SearchQuerySpec(DS, ALL_TIME, "all", None, ["s_nation"], InsensitiveContainsSearchQuerySpec(), 2147483647)),
        ("distinct_limit_sorted", SearchQuerySpec(DS, ALL_TIME, "all", SelectorFilterSpec("c_mktsegment", "BUILDING"), ["p_brand"],
                                                  InsensitiveContainsSearchQuerySpec(), 7, SortSearchQuerySpec("lexicographic"))),
        ("insensitive_two_dims", SearchQuerySpec(DS, ["1994-01-01T00:00:00.000Z/1995-01-01T00:00:00.000Z"], "all", None, ["s_nation", "c_nation", "l_shipmode"],
                                                 InsensitiveContainsSearchQuerySpec("ai"), 1000)),
        ("contains_case", SearchQuerySpec(DS, ALL_TIME, "all", None, ["p_type"], {"type": "contains", "value": "BRUSHED", "caseSensitive": True}, 1000,
                                          SortSearchQuerySpec("strlen"))),
        ("fragment", SearchQuerySpec(DS, ALL_TIME, "all", None, ["p_type"], {"type": "fragment", "values": ["promo", "TIN"]}, 1000)),
        ("integer_dim", SearchQuerySpec(DS, ALL_TIME, "all", None, ["p_size"], InsensitiveContainsSearchQuerySpec("4"), 1000)),
        ("no_match", SearchQuerySpec(DS, ALL_TIME, "all", None, ["s_nation"], InsensitiveContainsSearchQuerySpec("zzz"), 1000)),
        ("unknown_dim", SearchQuerySpec(DS, ALL_TIME, "all", None, ["no_such_dim", "l_returnflag"], InsensitiveContainsSearchQuerySpec(), 1000)),
    ]


def js_filter_cases():
    """JavascriptFilterSpec functions in the expression subset the engine evaluates per dictionary entry."""
    return [
        # docs/benchmark/druid/queries/q3.json, TPCHQueries.scala:119-121
        ("js_lt_date", JavascriptFilterSpec("o_orderdate", "function(x) { return(x < '1995-03-15') }")),
        # docs/benchmark/druid/queries/q5.json: two javascript bounds under an and
        ("js_date_window", And(JavascriptFilterSpec("o_orderdate", "function(x) { return(x >= '1994-01-01') }"),
                               JavascriptFilterSpec("o_orderdate", "function(x) { return(x < '1995-01-01') }"))),
        # the text JSCodeGenerator emits: FilterPlanningTest.scala "jsUpper" / "jsCoalesce" (:131-157)
        ("js_lower", JavascriptFilterSpec("c_mktsegment", "function (c_mktsegment) {\n            \n            \n            \n\n            "
                                                          "return(((c_mktsegment).toLowerCase())  ==  (\"building\"));\n            }")),
        ("js_coalesce", JavascriptFilterSpec("l_shipmode", "function (l_shipmode) {\n\n            return((((l_shipmode) || (\"no-mode\")))  ==  (\"AIR\"));\n            }")),
        ("js_numeric", JavascriptFilterSpec("p_size", "function(x) { return parseInt(x) % 7 == 3 || x > 45; }")),
        ("js_string_ops", JavascriptFilterSpec("p_brand", "function(b) { return b.substr(0, 6).concat('x') != 'Brand#x' ? false : (b.charAt(6) >= '3' && b.length == 8) }")),
        ("js_unknown_dim", JavascriptFilterSpec("no_such_dim", "function(x) { return x == null }")),
    ]


def extraction_dims():
    """(name, ExtractionDimensionSpec) on dimensions other than __time."""
    return [
        # docs/benchmark/druid/queries/q8.json: year of the o_orderdate string dimension
        ("year_of_orderdate", ExtractionDimensionSpec("o_orderdate", "o_year", TimeFormatExtractionFunctionSpec("yyyy"))),
        ("month_name_tz", ExtractionDimensionSpec("o_orderdate", "m", TimeFormatExtractionFunctionSpec("MMM yyyy", "+05:30"))),
        ("time_parse", ExtractionDimensionSpec("o_orderdate", "q", TimeParsingExtractionFunctionSpec("yyyy-MM-dd", "MM/yy"))),
        ("js_brand_group", ExtractionDimensionSpec("p_brand", "brand_group", JavaScriptExtractionFunctionSpec("function(b) { return b.substr(0, 7).toLowerCase() }"))),
        ("js_int_bucket", ExtractionDimensionSpec("p_size", "size_bucket", JavaScriptExtractionFunctionSpec("function(x) { return Math.floor(x / 10) * 10 }"))),
        ("js_null", ExtractionDimensionSpec("l_shipmode", "air_or_null", JavaScriptExtractionFunctionSpec("function(x) { return x == 'AIR' || x == 'REG AIR' ? 'by air' : null }"))),
        ("lookup", ExtractionDimensionSpec("l_returnflag", "flag", {"type": "lookup", "lookup": {"type": "map", "map": {"A": "accepted", "R": "returned"}}, "retainMissingValue": True})),
        ("lookup_replace", ExtractionDimensionSpec("c_mktsegment", "seg", {"type": "lookup", "lookup": {"type": "map", "map": {"BUILDING": "b"}}, "replaceMissingValueWith": "other"})),
        ("regex", ExtractionDimensionSpec("p_type", "material", {"type": "regex", "expr": " ([A-Z]+)$"})),
        ("substring", ExtractionDimensionSpec("s_nation", "n3", {"type": "substring", "index": 1, "length": 3})),
        ("strlen", ExtractionDimensionSpec("p_type", "len", {"type": "strlen"})),
        ("upper", ExtractionDimensionSpec("l_shipmode", "u", {"type": "lower"})),
    ]


def with_extraction(dim, other="l_linestatus", strategy=None, f=None):
    dims = [dim] + ([DefaultDimensionSpec(other)] if other else [])
    return GroupByQuerySpec(DS, dims, None, None, "all", f, [agg("count", "n", "count"), agg("doubleSum", "price", "l_extendedprice"),
                                                             CardinalityAggregationSpec("custs", ["o_custkey"])], None, ALL_TIME, ctx(strategy))


REVENUE_JS = ("function(current, l_extendedprice, l_discount) { return current + (l_extendedprice * (1 - l_discount)); }",
              "function(partialA, partialB) { return partialA + partialB; }", "function()                   { return 0; }")


def js_aggs():
    """Aggregations whose argument is arithmetic over metrics: the javascript form the reference generates
    (JSAggGenerator.scala:36-119, tpch_index_task.json.template:138-156) and the `expression` attribute."""
    return [
        agg("count", "n", "count"),
        JavascriptAggregationSpec("javascript", "revenue", ["l_extendedprice", "l_discount"], *REVENUE_JS),
        JavascriptAggregationSpec("javascript", "charge", ["l_extendedprice", "l_discount", "l_tax"],
                                  "function(current, l_extendedprice, l_discount, l_tax) { return current + (l_extendedprice *(1 - l_discount) * (1 + l_tax)); }",
                                  "function(partialA, partialB) { return (partialA + partialB); }", "function() { return 0; }"),
        # what JSAggGenerator emits for MIN / MAX over ((l) op (r)) trees, an integer metric on one side
        JavascriptAggregationSpec("javascript", "min_unit", ["l_extendedprice", "l_quantity"],
                                  "function (current, l_extendedprice, l_quantity) { return Math.min(current, ((l_extendedprice) / (l_quantity)));}",
                                  "function (partialA, partialB) {return Math.min(partialA, partialB);}", "function() { return Number.POSITIVE_INFINITY; }"),
        JavascriptAggregationSpec("javascript", "max_stock_value", ["ps_availqty", "ps_supplycost"],
                                  "function (current, ps_availqty, ps_supplycost) { return Math.max(current, ((ps_availqty) * (ps_supplycost)));}",
                                  "function (partialA, partialB) {return Math.max(partialA, partialB);}", "function() { return Number.NEGATIVE_INFINITY; }"),
        {"type": "doubleSum", "name": "disc_amount", "expression": "l_extendedprice * l_discount"},
        {"type": "doubleMax", "name": "max_abs_margin", "expression": "Math.abs(l_extendedprice - ps_supplycost * l_quantity) / 2"},
    ]


def js_agg_query(dims=("l_returnflag", "l_linestatus"), f=None, strategy=None, granularity="all"):
    return GroupByQuerySpec(DS, [DefaultDimensionSpec(d) for d in dims], None, None, granularity, f, js_aggs(), None, ALL_TIME, ctx(strategy))


def with_filter(f, strategy=None):
    return GroupByQuerySpec(DS, [DefaultDimensionSpec("l_shipmode"), DefaultDimensionSpec("l_returnflag")], None, None, "all", f,
                            [agg("count", "n", "count"), agg("longSum", "qty", "ps_availqty"), agg("doubleSum", "price", "l_extendedprice")],
                            None, ALL_TIME, ctx(strategy))


def agg_zoo(strategy=None):
    """Every aggregator type incl. min/max, cross-type sums and a filtered aggregation."""
    return GroupByQuerySpec(
        DS, [DefaultDimensionSpec("c_region")], None, None, "all", None,
        [agg("count", "n", "count"), agg("longSum", "lsum", "ps_availqty"), agg("longMin", "lmin", "ps_availqty"), agg("longMax", "lmax", "ps_availqty"),
         agg("doubleMin", "dmin", "l_discount"), agg("doubleMax", "dmax", "ps_supplycost"), agg("doubleSum", "dsum_of_long", "ps_availqty"),
         agg("longSum", "lsum_of_double", "l_quantity"), agg("doubleMin", "dmin_of_long", "ps_availqty"), agg("longMax", "lmax_of_double", "l_extendedprice"),
         FilteredAggregationSpec("n_air", SelectorFilterSpec("l_shipmode", "AIR"), agg("count", "n_air", "count")),
         FilteredAggregationSpec("price_r", SelectorFilterSpec("l_returnflag", "R"), agg("doubleSum", "price_r", "l_extendedprice")),
         CardinalityAggregationSpec("d_nation", ["s_nation"]), CardinalityAggregationSpec("d_cust", ["o_custkey"])],
        None, ALL_TIME, ctx(strategy))


def having_query(strategy=None):
    return GroupByQuerySpec(
        DS, [DefaultDimensionSpec("s_nation")], LimitSpec(2147483647, [OrderByColumnSpec("n", "descending"), OrderByColumnSpec("s_nation")]),
        LogicalBinaryOpHavingSpec("and", [ComparisonHavingSpec("greaterThan", "n", 10.0), NotOpHavingSpec("not", ComparisonHavingSpec("lessThan", "price", 1000.0))]),
        "all", None, [agg("count", "n", "count"), agg("doubleSum", "price", "l_extendedprice")], None, ALL_TIME, ctx(strategy))


def topn_variants():
    a = [agg("doubleSum", "revenue", "l_extendedprice"), agg("count", "n", "count")]
    return [
        ("numeric_desc", TopNQuerySpec(DS, ALL_TIME, "all", None, a, None, DefaultDimensionSpec("p_brand"), 3, NumericTopNMetricSpec("revenue"))),
        ("numeric_inverted", TopNQuerySpec(DS, ALL_TIME, "all", None, a, None, DefaultDimensionSpec("p_brand"), 3, InvertedTopNMetricSpec(NumericTopNMetricSpec("revenue")))),
        ("lexicographic", TopNQuerySpec(DS, ALL_TIME, "all", None, a, None, DefaultDimensionSpec("p_brand"), 4, LexiographicTopNMetricSpec("Brand#23"))),
        ("int_count", TopNQuerySpec(DS, ALL_TIME, "all", None, a, None, DefaultDimensionSpec("c_nation"), 200, NumericTopNMetricSpec("n"))),
    ]
