"""sdo_groupby_sum_noalloc, the entry point for callers that may not allocate GPU memory (a plugin
inside another runtime's fixed device reservation, a host that forbids cudaMalloc after start-up).

It is called here through the C ABI on caller-owned device buffers and checked against the regular
entry point (sdo_query_execute) on the same rows and, independently of the engine, against numpy on
the raw generated rows (np.bincount, math.fsum): counts bit-exact, double sums within the 1e-9
relative tolerance the other parity tests use.  Two calls run back to back on different metrics, so
partials left in the shared scratch array by the first call would show up in the second; around the
second call the free device memory (cudaMemGetInfo) must not move, which is the promise the entry
point exists for.
Collected last on purpose (file name), so an -x run has finished every other test before this one."""
import json
import math

import numpy as np
import pytest

import helpers as H
import queries as Q

pytestmark = pytest.mark.gpu

SF = 0.02
DS = "tpch_noalloc"


def _engine_sums(eng, metric):
    spec = Q.GroupByQuerySpec(DS, [Q.DefaultDimensionSpec("l_returnflag")], None, None, "all", None,
                              [Q.agg("count", "n", "count"), Q.agg("doubleSum", "s", metric)], None, Q.ALL_TIME)
    res = eng.execute(spec)
    return {res["l_returnflag"][i]: (int(res["n"][i]), float(res["s"][i])) for i in range(res.n_rows)}


def test_noalloc_matches_query_execute_twice_in_a_row(gpu_engine):
    import torch
    import spark_druid_olap_b200 as sdo
    from spark_druid_olap_b200 import tpch
    eng = gpu_engine
    try:
        eng.drop_datasource(DS)
    except Exception:
        pass
    tab = tpch.load_into(eng, DS, SF, ["l_returnflag", "l_extendedprice", "l_quantity", "count"], keep=True)[0]
    try:
        names = tab.dicts["l_returnflag"]  # generator code -> string; the no-alloc call sees raw codes
        n, nk = tab.n_rows, len(names)
        dev = torch.device("cuda", eng.device)
        keys = torch.from_numpy(tab.data["l_returnflag"].astype(np.uint8)).to(dev)
        counts = torch.full((nk,), -1, dtype=torch.int64, device=dev)
        sums = torch.full((nk,), float("nan"), dtype=torch.float64, device=dev)
        raw_keys = tab.data["l_returnflag"]
        for call, metric in enumerate(("l_extendedprice", "l_quantity")):  # two calls in a row on the shared scratch
            vals = torch.from_numpy(tab.data[metric]).to(dev)
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(dev)[0]  # cudaMemGetInfo
            sdo.groupby_sum_noalloc(eng.device, keys.data_ptr(), vals.data_ptr(), n, nk, counts.data_ptr(), sums.data_ptr(), None)
            torch.cuda.synchronize()
            if call == 1:  # the first call may load the kernel's module; from then on the call takes no device memory at all
                assert torch.cuda.mem_get_info(dev)[0] == free_before
            assert torch.cuda.current_device() == eng.device
            want = _engine_sums(eng, metric)
            got_c, got_s = counts.cpu().numpy(), sums.cpu().numpy()
            assert sum(got_c) == n
            np_counts = np.bincount(raw_keys, minlength=nk)
            for code, name in enumerate(names):
                wc, ws = want.get(name, (0, 0.0))
                assert int(got_c[code]) == wc, (metric, name)
                assert abs(float(got_s[code]) - ws) <= H.F64_RTOL * max(abs(ws), 1e-300), (metric, name, got_s[code], ws)
                # and against the raw rows, without the engine in between
                exact = math.fsum(tab.data[metric][raw_keys == code].tolist())
                assert int(got_c[code]) == int(np_counts[code]), (metric, name)
                assert abs(float(got_s[code]) - exact) <= H.F64_RTOL * max(abs(exact), 1e-300), (metric, name, got_s[code], exact)
        # a row count that is not a multiple of 4 and ids outside n_keys (ignored)
        m = 1001
        sdo.groupby_sum_noalloc(eng.device, keys.data_ptr(), vals.data_ptr(), m, 2, counts.data_ptr(), sums.data_ptr(), None)
        torch.cuda.synchronize()
        k = tab.data["l_returnflag"][:m]
        v = tab.data["l_quantity"][:m]
        for code in range(2):
            assert int(counts[code]) == int((k == code).sum())
            assert abs(float(sums[code]) - float(v[k == code].sum())) <= 1e-9 * max(1.0, abs(float(v[k == code].sum())))
        # sizes beyond the fixed scratch are refused with a clear error, nothing is launched
        with pytest.raises(sdo.InvalidQuery, match="n_keys"):
            sdo.groupby_sum_noalloc(eng.device, keys.data_ptr(), vals.data_ptr(), n, sdo.NOALLOC_MAX_KEYS + 1, counts.data_ptr(), sums.data_ptr(), None)
        with pytest.raises(sdo.InvalidQuery, match="n_rows"):
            sdo.groupby_sum_noalloc(eng.device, keys.data_ptr(), vals.data_ptr(), sdo.NOALLOC_MAX_ROWS + 1, nk, counts.data_ptr(), sums.data_ptr(), None)
    finally:
        eng.drop_datasource(DS)
