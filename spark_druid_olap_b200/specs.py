"""Host-side mirror of the reference's query model, same names and argument meaning:
/root/reference/src/main/scala/org/sparklinedata/druid/DruidQuerySpec.scala (case classes) serialised
as /root/reference/src/main/scala/org/sparklinedata/druid/Utils.scala:38-105 does it (json4s
ShortTypeHints: every polymorphic node carries "jsonClass"; Option = absent key).

In a Spark deployment these objects are the Scala ones and stay untouched; this mirror exists so
that tests and the bench build specs the way the reference's tests do (TPCHQueries.scala:30-140)
and hand the engine byte-compatible JSON.
"""
import json


class _Spec:
    _fields = ()
    _defaults = {}

    def __init__(self, *args, **kw):
        names = list(self._fields)
        if len(args) > len(names):
            raise TypeError(f"{type(self).__name__} takes at most {len(names)} arguments")
        vals = dict(self._defaults)
        vals.update(zip(names, args))
        for k, v in kw.items():
            if k not in names:
                raise TypeError(f"{type(self).__name__} has no field {k}")
            vals[k] = v
        for n in names:
            if n not in vals:
                raise TypeError(f"{type(self).__name__} missing field {n}")
            setattr(self, n, vals[n])

    def to_dict(self):
        d = {"jsonClass": type(self).__name__}
        for n in self._fields:
            v = getattr(self, n)
            if v is None:
                continue  # Option = None -> key absent
            d[n.rstrip("_")] = _ser(v)
        return d

    def to_json(self):
        return json.dumps(self.to_dict())

    def __repr__(self):
        return f"{type(self).__name__}({', '.join(f'{n}={getattr(self, n)!r}' for n in self._fields)})"


def _ser(v):
    if isinstance(v, _Spec):
        return v.to_dict()
    if isinstance(v, (list, tuple)):
        return [_ser(x) for x in v]
    if isinstance(v, dict):
        return {k: _ser(x) for k, x in v.items()}
    return v


# ---- extraction functions / dimensions (DruidQuerySpec.scala:47-138) ---------------------------
class TimeFormatExtractionFunctionSpec(_Spec):
    _fields = ("type", "format", "timeZone", "locale")

    def __init__(self, format, timeZone=None, locale=None):
        super().__init__("timeFormat", format, timeZone, locale)


class TimeParsingExtractionFunctionSpec(_Spec):
    """DruidQuerySpec.scala:82-88"""
    _fields = ("type", "timeFormat", "resultFormat")

    def __init__(self, timeFormat, resultFormat):
        super().__init__("time", timeFormat, resultFormat)


class JavaScriptExtractionFunctionSpec(_Spec):
    """DruidQuerySpec.scala:90-96"""
    _fields = ("type", "function", "injective")

    def __init__(self, function, injective=False):
        super().__init__("javascript", function, injective)


class LookUpMap(_Spec):
    _fields = ("type", "map")

    def to_dict(self):  # LookUpMap carries no jsonClass hint need, but json4s emits one for every case class in the hint list
        return {"type": self.type, "map": dict(self.map)}


class InExtractionFnSpec(_Spec):
    _fields = ("type", "lookup")

    def __init__(self, valLst):
        super().__init__("lookup", LookUpMap("map", {v: "true" for v in valLst}))


class DefaultDimensionSpec(_Spec):
    _fields = ("type", "dimension", "outputName")

    def __init__(self, dimension, outputName=None):
        super().__init__("default", dimension, outputName or dimension)


class ExtractionDimensionSpec(_Spec):
    _fields = ("type", "dimension", "outputName", "extractionFn")

    def __init__(self, dimension, outputName, extractionFn):
        super().__init__("extraction", dimension, outputName, extractionFn)


# ---- granularity (DruidQuerySpec.scala:140-150) ---------------------------------------------------
class DurationGranularitySpec(_Spec):
    _fields = ("type", "duration")

    def __init__(self, duration):
        super().__init__("duration", duration)


class PeriodGranularitySpec(_Spec):
    _fields = ("type", "period", "timeZone", "origin")

    def __init__(self, period, timeZone=None, origin=None):
        super().__init__("period", period, timeZone, origin)


# ---- filters (DruidQuerySpec.scala:152-281) -------------------------------------------------------
class SelectorFilterSpec(_Spec):
    _fields = ("type", "dimension", "value")

    def __init__(self, dimension, value):
        super().__init__("selector", dimension, value)


class RegexFilterSpec(_Spec):
    _fields = ("type", "dimension", "pattern")

    def __init__(self, dimension, pattern):
        super().__init__("regex", dimension, pattern)


class StringContainsSpec(_Spec):
    _fields = ("type", "dimension", "value")

    def __init__(self, dimension, value):
        super().__init__("contains", dimension, value)


class LogicalFilterSpec(_Spec):
    _fields = ("type", "fields")


class NotFilterSpec(_Spec):
    _fields = ("type", "field")


class ExtractionFilterSpec(_Spec):
    """SQL IN: extraction filter whose lookup maps every listed value to "true"."""
    _fields = ("type", "dimension", "value", "extractionFn")

    def __init__(self, dimension, valList):
        super().__init__("extraction", dimension, "true", InExtractionFnSpec(valList))


class JavascriptFilterSpec(_Spec):
    _fields = ("type", "dimension", "function")

    def __init__(self, dimension, function):
        super().__init__("javascript", dimension, function)


class BoundFilterSpec(_Spec):
    _fields = ("type", "dimension", "lower", "lowerStrict", "upper", "upperStrict", "alphaNumeric")

    def __init__(self, dimension, lower=None, lowerStrict=None, upper=None, upperStrict=None, alphaNumeric=False):
        super().__init__("bound", dimension, lower, lowerStrict, upper, upperStrict, alphaNumeric)


def And(*fields):
    return LogicalFilterSpec("and", list(fields))


def Or(*fields):
    return LogicalFilterSpec("or", list(fields))


def Not(field):
    return NotFilterSpec("not", field)


# ---- aggregations (DruidQuerySpec.scala:283-377) --------------------------------------------------
class FunctionAggregationSpec(_Spec):
    _fields = ("type", "name", "fieldName")


class CardinalityAggregationSpec(_Spec):
    _fields = ("type", "name", "fieldNames", "byRow")

    def __init__(self, name, fieldNames, byRow=True):
        super().__init__("cardinality", name, list(fieldNames), byRow)


class HyperUniqueAggregationSpec(_Spec):
    _fields = ("type", "name", "fieldName")

    def __init__(self, name, fieldName):
        super().__init__("hyperUnique", name, fieldName)


class FilteredAggregationSpec(_Spec):
    _fields = ("type", "name", "filter", "aggregator")

    def __init__(self, name, filter, aggregator):
        super().__init__("filtered", name, filter, aggregator)


class JavascriptAggregationSpec(_Spec):
    _fields = ("type", "name", "fieldNames", "fnAggregate", "fnCombine", "fnReset")


# ---- post-aggregations (DruidQuerySpec.scala:379-430) ---------------------------------------------
class FieldAccessPostAggregationSpec(_Spec):
    _fields = ("type", "fieldName")

    def __init__(self, fieldName):
        super().__init__("fieldAccess", fieldName)


class ConstantPostAggregationSpec(_Spec):
    _fields = ("type", "name", "value")

    def __init__(self, name, value):
        super().__init__("constant", name, value)


class HyperUniqueCardinalityPostAggregationSpec(_Spec):
    _fields = ("type", "name", "fieldName")

    def __init__(self, name, fieldName):
        super().__init__("hyperUniqueCardinality", name, fieldName)


class ArithmeticPostAggregationSpec(_Spec):
    _fields = ("type", "name", "fn", "fields", "ordering")

    def __init__(self, name, fn, fields, ordering=None):
        super().__init__("arithmetic", name, fn, list(fields), ordering)


# ---- limit / having / topN metric (DruidQuerySpec.scala:432-506) ----------------------------------
class OrderByColumnSpec(_Spec):
    _fields = ("dimension", "direction")

    def __init__(self, dimension, direction="ascending"):
        if isinstance(direction, bool):
            direction = "ascending" if direction else "descending"
        super().__init__(dimension, direction)

    def to_dict(self):
        return {"dimension": self.dimension, "direction": self.direction}


class LimitSpec(_Spec):
    _fields = ("type", "limit", "columns")

    def __init__(self, limit, columns):
        super().__init__("default", limit, list(columns))

    def to_dict(self):
        return {"type": self.type, "limit": self.limit, "columns": [c.to_dict() for c in self.columns]}


class ComparisonHavingSpec(_Spec):
    _fields = ("type", "aggregation", "value")


class LogicalBinaryOpHavingSpec(_Spec):
    _fields = ("type", "havingSpecs")


class NotOpHavingSpec(_Spec):
    _fields = ("type", "havingSpec")


class NumericTopNMetricSpec(_Spec):
    _fields = ("type", "metric")

    def __init__(self, metric):
        super().__init__("numeric", metric)


class LexiographicTopNMetricSpec(_Spec):
    _fields = ("type", "previousStop")

    def __init__(self, previousStop=""):
        super().__init__("lexicographic", previousStop)


class InvertedTopNMetricSpec(_Spec):
    _fields = ("type", "metric")

    def __init__(self, metric):
        super().__init__("inverted", metric)


class QuerySpecContext(_Spec):
    _fields = ("queryId", "timeout", "priority", "useCache", "populateCache", "bySegment", "chunkPeriod", "minTopNThreshold",
               "maxResults", "maxIntermediateRows", "groupByIsSingleThreaded", "groupByStrategy", "gpuStrategy")
    _defaults = {k: None for k in _fields[1:]}

    def to_dict(self):
        d = super().to_dict()
        d.pop("jsonClass")
        return d


# ---- queries (DruidQuerySpec.scala:573-868) ---------------------------------------------------------
def _gran(g):
    return g if isinstance(g, str) else g


class GroupByQuerySpec(_Spec):
    _fields = ("queryType", "dataSource", "dimensions", "limitSpec", "having", "granularity", "filter", "aggregations",
               "postAggregations", "intervals", "context")

    def __init__(self, dataSource, dimensions, limitSpec, having, granularity, filter, aggregations, postAggregations, intervals,
                 context=None):
        super().__init__("groupBy", dataSource, list(dimensions), limitSpec, having, _gran(granularity), filter, list(aggregations),
                         postAggregations, list(intervals), context)


class TimeSeriesQuerySpec(_Spec):
    _fields = ("queryType", "dataSource", "intervals", "granularity", "filter", "aggregations", "postAggregations", "context")

    def __init__(self, dataSource, intervals, granularity, filter, aggregations, postAggregations=None, context=None):
        super().__init__("timeseries", dataSource, list(intervals), _gran(granularity), filter, list(aggregations), postAggregations,
                         context)


class PagingSpec(_Spec):
    """DruidQuerySpec.scala:976-978"""
    _fields = ("pagingIdentifiers", "threshold", "fromNext")

    def __init__(self, pagingIdentifiers, threshold, fromNext=None):
        super().__init__(dict(pagingIdentifiers), threshold, fromNext)


class SelectSpecWithIntervals(_Spec):
    """DruidQuerySpec.scala:1021-1067"""
    _fields = ("queryType", "dataSource", "dimensions", "metrics", "filter", "pagingSpec", "intervals", "descending", "granularity", "context")

    def __init__(self, dataSource, dimensions, metrics, filter, pagingSpec, intervals, descending=False, granularity="all", context=None):
        super().__init__("select", dataSource, list(dimensions), list(metrics), filter, pagingSpec, list(intervals), descending, granularity, context)

    def withPagingIdentifier(self, ps):
        d = self.to_dict()
        d["pagingSpec"] = dict(d["pagingSpec"], pagingIdentifiers=dict(ps), fromNext=True)
        return d


class InsensitiveContainsSearchQuerySpec(_Spec):
    """DruidQuerySpec.scala:547-552"""
    _fields = ("type", "value")

    def __init__(self, value=""):
        super().__init__("insensitive_contains", value)


class SortSearchQuerySpec(_Spec):
    _fields = ("type",)


class SearchQuerySpec(_Spec):
    """DruidQuerySpec.scala:870-927"""
    _fields = ("queryType", "dataSource", "intervals", "granularity", "filter", "searchDimensions", "query", "limit", "sort", "context")

    def __init__(self, dataSource, intervals, granularity, filter, searchDimensions, query, limit, sort=None, context=None):
        super().__init__("search", dataSource, list(intervals), _gran(granularity), filter, list(searchDimensions), query, limit, sort, context)


class TopNQuerySpec(_Spec):
    _fields = ("queryType", "dataSource", "intervals", "granularity", "filter", "aggregations", "postAggregations", "dimension",
               "threshold", "metric", "context")

    def __init__(self, dataSource, intervals, granularity, filter, aggregations, postAggregations, dimension, threshold, metric,
                 context=None):
        super().__init__("topN", dataSource, list(intervals), _gran(granularity), filter, list(aggregations), postAggregations,
                         dimension, threshold, metric, context)
