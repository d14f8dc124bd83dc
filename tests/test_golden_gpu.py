"""The CUDA engine against the committed answers of tests/golden/results_sf0005.json (produced by the numpy oracle in an
earlier run and checked into the repository), AOT interpreter and run-time specialised kernel."""
import importlib.util
import json
import os

import pytest

import helpers as H

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_result_fixtures", os.path.join(HERE, "golden", "make_result_fixtures.py"))
fixtures = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(fixtures)
GOLDEN = json.load(open(os.path.join(HERE, "golden", "results_sf0005.json")))


@pytest.fixture(scope="module")
def loaded(gpu_engine):
    from spark_druid_olap_b200 import tpch
    name = "tpch_golden"
    try:
        gpu_engine.drop_datasource(name)
    except Exception:
        pass
    tpch.load_into(gpu_engine, name, fixtures.SF, fixtures.COLS, shards=(0, 1), n_shards=2)
    yield gpu_engine, name
    gpu_engine.drop_datasource(name)


@pytest.mark.parametrize("jit", [False, True])
@pytest.mark.parametrize("name,spec,ordered", fixtures.cases(), ids=[c[0] for c in fixtures.cases()])
def test_engine_gives_the_committed_answers(loaded, name, spec, ordered, jit):
    eng, ds = loaded
    q = json.loads(spec.to_json())
    q["dataSource"] = ds
    q["context"] = dict(q.get("context") or {}, gpuJit=jit)
    assert eng.metadata(ds)["numRows"] == GOLDEN["rows"]
    H.assert_rows_equal(H.engine_rows(eng.execute(q)), GOLDEN["results"][name]["rows"], ordered=ordered, what=f"engine vs golden {name} jit={jit}")
