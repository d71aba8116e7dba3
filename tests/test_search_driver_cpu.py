"""The product's search driver (gfalign_b200/csrc/search.cpp, compiled unchanged) with its engine calls
answered by the CPU oracle (tests/native/engine_stub.cpp): the tree path of round 2 — incremental
composition (main / fresh set / delta), effective-key speculation roots, incremental path
materialisation — against the reference's golden rows and against the full-scan driver, on CPU."""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

from gfalign_b200 import capi, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "native", "_build")


@pytest.fixture(scope="module")
def cli_cases():
    import gzip
    with gzip.open(os.path.join(ROOT, "tests", "golden", "cli.json.gz"), "rt") as f:
        return {c["name"]: c for c in json.load(f)["cases"]}


@pytest.fixture(scope="module")
def stub():
    os.makedirs(BUILD, exist_ok=True)
    so = os.path.join(BUILD, "libsearch_oracle.so")
    srcs = [os.path.join(ROOT, "gfalign_b200", "csrc", "search.cpp"),
            os.path.join(ROOT, "tests", "native", "engine_stub.cpp")]
    obj = os.path.join(BUILD, "oracle.o")
    subprocess.run(["gcc", "-O2", "-fPIC", "-pthread", "-c", os.path.join(ROOT, "oracle", "oracle.c"), "-o", obj], check=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-x", "c++", srcs[0], srcs[1],
                    "-x", "none", obj, "-o", so], check=True)
    lib = C.CDLL(so)
    lib.stub_set_alignments.argtypes = [capi._u32p, capi._u32p, C.c_uint64]
    lib.stub_counters.argtypes = [C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.gfe_search.argtypes = [C.c_void_p, C.POINTER(capi._Graph), C.POINTER(capi._SearchParams), capi._ROW_FN,
                               C.c_void_p, C.POINTER(capi._SearchStats)]
    return lib


def _search(lib, t, off, st, full_scan=False, **kw):
    off, st = np.ascontiguousarray(off, np.uint32), np.ascontiguousarray(st, np.uint32)
    lib.stub_set_alignments(off.ctypes.data_as(capi._u32p), st.ctypes.data_as(capi._u32p), off.size - 1)
    nodes, counts, node_count = synth.node_table(t)
    handle = C.c_void_p(1)          # the stub never dereferences the engine

    def call(g, p, cb, stats):
        if full_scan:
            os.environ["GFE_SEARCH_FULLSCAN"] = "1"
        try:
            rc = lib.gfe_search(handle, g, p, cb, None, stats)
        finally:
            os.environ.pop("GFE_SEARCH_FULLSCAN", None)
        assert rc == 0, rc

    rows, stats = capi._run_search(call, t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination,
                                   nodes, counts, node_count, kw.get("max_steps", 100000), kw.get("min_nodes", 0),
                                   kw.get("return_all_paths", False), kw.get("speculate", 0), None)
    d, f = C.c_uint64(), C.c_uint64()
    lib.stub_counters(C.byref(d), C.byref(f))
    return rows, stats, d.value, f.value


def test_tree_driver_reproduces_reference_goldens(stub, cli_cases):
    from test_oracle import _search_inputs, search_rows_text
    seen = 0
    for name, c in cli_cases.items():
        if "_search" not in name:
            continue
        t, off, st, kw = _search_inputs(c)
        for spec in (0, 1, 7):
            rows, stats, _, _ = _search(stub, t, off, st, speculate=spec, **kw)
            assert search_rows_text(rows, stats, t.names, kw) == c["stdout"].splitlines(), (name, spec)
        seen += 1
    assert seen >= 13


def test_tree_driver_deep_paths_match_reference_binary_golden(stub):
    """tests/golden/smoke_search.json: stdout of the reference binary on a cfg3 sample whose paths reach
    ~90 steps, so most candidates are scored as deltas to their parents."""
    with open(os.path.join(ROOT, "tests", "golden", "smoke_search.json")) as f:
        gold = json.load(f)
    w = synth.make_workload("cfg3", A=gold["A"])
    t = w.tangle
    rows, stats, deltas, fulls = _search(stub, t, w.offsets, w.steps, max_steps=gold["max_steps"])
    text = "".join(r.columns(t.names) + "\n" for r in rows)
    if stats["extractions"] >= gold["max_steps"]:
        text += f"Reached maximum number of steps ({stats['extractions']})\n"
    assert text == gold["stdout"]
    assert deltas > 3 * fulls > 0      # most candidates are scored as deltas to their parents


def test_tree_driver_equals_full_scan_driver(stub):
    w = synth.make_workload("cfg3", A=400)
    t = w.tangle
    kw = dict(max_steps=1500, return_all_paths=True)
    want, st_full, _, _ = _search(stub, t, w.offsets, w.steps, full_scan=True, **kw)
    for spec in (0, 3, 64, 5000):
        got, st, deltas, _ = _search(stub, t, w.offsets, w.steps, speculate=spec, **kw)
        assert [r.columns(t.names) for r in got] == [r.columns(t.names) for r in want], spec
        assert st["evaluations"] == st_full["evaluations"] and deltas > 0
    assert len(want) > 2
