"""Select queries (SelectSpec, DruidQuerySpec.scala:980-1098): paging through the rows that pass interval and filter.
The union of all pages must be exactly the oracle's rows; inside a page rows come in segment order then time order, the
threshold is respected, and the pagingIdentifiers of one page resume the next."""
import json
from collections import Counter

import pytest

import helpers as H
import queries as Q
from oracle import druid_oracle as orc

pytestmark = pytest.mark.gpu

SF = 0.01
COLS = H.TEST_COLUMNS


@pytest.fixture(scope="module")
def loaded(gpu_engine):
    from spark_druid_olap_b200 import tpch
    name = "tpch_select"
    try:
        gpu_engine.drop_datasource(name)
    except Exception:
        pass
    tabs = tpch.load_into(gpu_engine, name, SF, COLS, shards=(0, 1), n_shards=2, keep=True)
    yield gpu_engine, H.oracle_table(tabs), name
    gpu_engine.drop_datasource(name)


def _freeze(r, names):
    return tuple(r[n] for n in names)


CASES = [
    ("filtered_small_pages", ["l_returnflag", "s_nation", "o_orderkey"], ["l_extendedprice", "ps_availqty"],
     Q.And(Q.SelectorFilterSpec("c_mktsegment", "BUILDING"), Q.BoundFilterSpec("o_orderdate", "1995-01-01", False, "1995-04-01", True, False)),
     ["1992-01-01T00:00:00.000Z/1999-01-01T00:00:00.000Z"], 17),
    ("interval_only", ["l_shipmode"], ["l_quantity"], None, ["1996-03-10T00:00:00.000Z/1996-04-20T00:00:00.000Z"], 1000),
    ("one_page", ["p_brand", "p_size"], [], Q.SelectorFilterSpec("s_nation", "PERU"), ["1997-01-01T00:00:00.000Z/1997-03-01T00:00:00.000Z"], 100000),
    ("nothing", ["l_returnflag"], ["l_tax"], Q.SelectorFilterSpec("s_nation", "ATLANTIS"), Q.ALL_TIME, 10),
    ("javascript_filter", ["o_orderdate"], ["l_discount"], Q.js_filter_cases()[1][1], Q.ALL_TIME, 5000),
]


@pytest.mark.parametrize("name,dims,mets,flt,intervals,threshold", CASES, ids=[c[0] for c in CASES])
def test_select_pages_cover_the_oracle_rows(loaded, name, dims, mets, flt, intervals, threshold):
    eng, tbl, ds = loaded
    spec = Q.SelectSpecWithIntervals(ds, dims, mets, flt, Q.PagingSpec({}, threshold), intervals)
    want = orc.execute(tbl, json.loads(spec.to_json()))
    names = ["timestamp"] + dims + mets
    got, paging, pages = [], {}, 0
    last = {}  # segment -> (offset, timestamp) of the previous row
    while True:
        q = spec.withPagingIdentifier(paging) if pages else json.loads(spec.to_json())
        res = eng.execute(q)
        pages += 1
        assert res.names == ["timestamp", "segmentId", "offset"] + dims + mets
        rows = H.engine_rows(res)
        assert len(rows) <= threshold
        for r in rows:
            seg = r["segmentId"]
            if seg in last:
                assert r["offset"] == last[seg][0] + 1 and r["timestamp"] >= last[seg][1], (name, r, last[seg])
            else:
                assert r["offset"] == 0
            last[seg] = (r["offset"], r["timestamp"])
        # segments arrive in order: their ids within a page never go back
        segs = [r["segmentId"] for r in rows]
        assert segs == sorted(segs, key=lambda s: (s.split("_")[-3], int(s.split("_")[-1]))) or len(set(segs)) == 1
        got += rows
        paging = res.stats["pagingIdentifiers"]
        for seg, (o, _) in last.items():
            assert paging[seg] == o
        if len(rows) < threshold:
            break
        assert pages < 1000
    assert len(got) == len(want), (name, len(got), len(want))
    g = Counter(_freeze(r, names) for r in got)
    w = Counter(_freeze({k: (float(v) if isinstance(v, float) else v) for k, v in r.items()}, names) for r in want)
    assert g == w, name
    if name == "filtered_small_pages":
        assert pages > 3
    if name == "nothing":
        assert pages == 1 and not got


def test_select_from_next_false_and_errors(loaded):
    import spark_druid_olap_b200 as sdo
    eng, tbl, ds = loaded
    spec = Q.SelectSpecWithIntervals(ds, ["l_returnflag"], ["l_tax"], None, Q.PagingSpec({}, 7), ["1995-06-01T00:00:00.000Z/1995-07-01T00:00:00.000Z"])
    first = eng.execute(spec)
    ids = first.stats["pagingIdentifiers"]
    # fromNext absent / false: the offsets are where the page starts (the 0.9.0 client adds 1 itself)
    q = json.loads(spec.to_json())
    q["pagingSpec"] = {"pagingIdentifiers": ids, "threshold": 7}
    again = H.engine_rows(eng.execute(q))
    assert again[0]["offset"] == max(r["offset"] for r in H.engine_rows(first) if r["segmentId"] == again[0]["segmentId"])
    q["pagingSpec"]["fromNext"] = True
    nxt = H.engine_rows(eng.execute(q))
    assert nxt[0]["offset"] == again[0]["offset"] + 1 or nxt[0]["segmentId"] != again[0]["segmentId"]
    # no dimensions and metrics listed = all of them
    q_all = json.loads(Q.SelectSpecWithIntervals(ds, [], [], None, Q.PagingSpec({}, 3), Q.ALL_TIME).to_json())
    r = eng.execute(q_all)
    assert r.n_rows == 3 and len(r.names) == 3 + len(COLS)  # timestamp, segmentId, offset + every dimension and metric
    with pytest.raises(sdo.UnsupportedQuery):
        eng.execute(dict(q, descending=True))
    with pytest.raises(sdo.InvalidQuery):
        eng.execute(dict(q, pagingSpec={"pagingIdentifiers": {}, "threshold": 0}))
    with pytest.raises(sdo.DruidDataSourceException):
        eng.execute(dict(q, dimensions=["no_such_dim"]))
