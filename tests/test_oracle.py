"""CPU suite: pins the oracle (oracle/oracle.c) against the golden fixtures generated from the
compiled reference (tests/golden/make_golden.py), and — when oracle/_ref is present (build
container) — live against the reference's own alignPaths() on fresh random pairs."""
import gzip
import json
import os

import numpy as np
import pytest

from gfalign_b200 import synth
import textio

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def cli_cases():
    with gzip.open(os.path.join(GOLD, "cli.json.gz"), "rt") as f:
        return {c["name"]: c for c in json.load(f)["cases"]}


def _rc(b):
    return (np.asarray(b, dtype=np.uint32)[::-1] ^ 1).astype(np.uint32)


def test_pair_scores_golden(oracle):
    g = np.load(os.path.join(GOLD, "pairs.npz"))
    n = g["t_fw"].size
    assert n > 2500
    for k in range(n):
        a = g["A_st"][g["A_off"][k]:g["A_off"][k + 1]]
        b = g["B_st"][g["B_off"][k]:g["B_off"][k + 1]]
        assert oracle.pair_score(a, b) == g["t_fw"][k], k
        assert oracle.pair_score(a, _rc(b)) == g["t_rc"][k], k


def test_pair_scores_live_reference(oracle, refshim):
    rng = np.random.default_rng(7)
    for _ in range(1500):
        alpha = int(rng.integers(2, 7))
        n, m = int(rng.integers(1, 25)), int(rng.integers(0, 14))
        a = (rng.integers(0, alpha, n) * 2 + rng.integers(0, 2, n)).astype(np.uint32)
        b = (rng.integers(0, alpha, m) * 2 + rng.integers(0, 2, m)).astype(np.uint32)
        tf, tr = refshim.pair_scores(a, b)
        assert (oracle.pair_score(a, b), oracle.pair_score(a, _rc(b))) == (tf, tr)


def _evalpath_expect(case):
    lines = case["stdout"].splitlines()
    scores = [int(l.rsplit("\t", 1)[1]) for l in lines[1:-1]]
    totals = [int(x) for x in lines[-1].split("\t")]
    return scores, totals


def test_evalpath_golden(oracle, cli_cases):
    seen = 0
    for name, c in cli_cases.items():
        if "_evalpath_" not in name:
            continue
        gfa = next(v for k, v in c["files"].items() if k.endswith(".gfa"))
        gaf = next(v for k, v in c["files"].items() if k.endswith(".gaf"))
        names, _, _ = textio.parse_gfa(gfa)
        _, off, st = textio.parse_gaf(gaf, names)
        path = textio.parse_path(c["args"][c["args"].index("-p") + 1], names)
        bad, good, tf, tr = oracle.evaluate_path(path, off, st, False, scores=True)
        scores, totals = _evalpath_expect(c)
        assert np.maximum(tf, tr).tolist() == scores, name
        uniques = len(set((path >> 1).tolist()))
        assert [bad, good, bad - good - uniques, path.size, uniques] == totals, name
        seen += 1
    assert seen >= 18


def test_evalgfa_golden(oracle, cli_cases):
    seen = 0
    for name, c in cli_cases.items():
        if not name.endswith("_evalgfa"):
            continue
        gfa = next(v for k, v in c["files"].items() if k.endswith(".gfa"))
        gaf = next(v for k, v in c["files"].items() if k.endswith(".gaf"))
        names, ef, et = textio.parse_gfa(gfa)
        _, off, st = textio.parse_gaf(gaf, names)
        w = oracle.edge_support(off, st, len(names), ef, et)
        want = [int(l.split("RC:i:")[1].split("\t")[0]) for l in c["out_L_lines"]]
        assert w.tolist() == want, name
        seen += 1
    assert seen >= 6


def _search_inputs(c):
    a = c["args"]
    gfa = c["files"][a[a.index("-f") + 1]]
    nl = c["files"][a[a.index("-n") + 1]]
    t = textio.tangle_from_text(gfa, nl, a[a.index("-s") + 1], a[a.index("-d") + 1])
    if "-g" in a:
        _, off, st = textio.parse_gaf(c["files"][a[a.index("-g") + 1]], t.names)
    else:
        off, st = np.zeros(1, np.uint32), np.zeros(0, np.uint32)
    kw = dict(max_steps=int(a[a.index("-m") + 1]) if "-m" in a else 100000,
              min_nodes=int(a[a.index("--min-nodes") + 1]) if "--min-nodes" in a else 0,
              return_all_paths="--return-all-paths" in a)
    return t, off, st, kw


def search_rows_text(rows, stats, names, kw):
    out = [r.columns(names) for r in rows]
    if stats["extractions"] >= kw["max_steps"]:
        out.append(f"Reached maximum number of steps ({stats['extractions']})")
    return out


def test_search_driver_golden_with_oracle_scorer(oracle, cli_cases):
    """Host logic of gfe_search (heap order, budgets, alt, hamiltonian flag, row emission) against
    the reference CLI's rows, with evaluatePath answered by the oracle through the scorer seam."""
    import gfalign_b200
    seen = 0
    for name, c in cli_cases.items():
        if "_search" not in name:
            continue
        t, off, st, kw = _search_inputs(c)
        nodes, counts, node_count = synth.node_table(t)

        def scorer(po, ps):
            return oracle.evaluate_paths(po, ps, off, st, True, 1)

        for spec in (1, 64):
            rows, stats = gfalign_b200.search_with_scorer(
                scorer, t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination,
                nodes, counts, node_count, speculate=spec, **kw)
            assert search_rows_text(rows, stats, t.names, kw) == c["stdout"].splitlines(), (name, spec)
        seen += 1
    assert seen >= 13


def test_oracle_threads_agree(oracle):
    w = synth.make_workload("cfg3", A=3000)
    po, ps = synth.candidate_paths(w.tangle, 8, seed=1)
    one = oracle.evaluate_paths(po, ps, w.offsets, w.steps, True, 1)
    many = oracle.evaluate_paths(po, ps, w.offsets, w.steps, True, 5)
    assert np.array_equal(one, many) and one.sum() > 0


def test_golden_evalpath_gapped_strings(oracle, cli_cases):
    """Pins the oracle's traceback (orc_pair_alignment) to the reference: the full evalPath stdout —
    every gapped alignment string, score and the totals — of the 18 golden CLI cases."""
    import textio
    from evalpath_text import evalpath_stdout, ops_from_gapped
    seen = 0
    for c in cli_cases.values():
        if "_evalpath_" not in c["name"]:
            continue
        gfa = next(v for k, v in c["files"].items() if k.endswith(".gfa"))
        gaf = next(v for k, v in c["files"].items() if k.endswith(".gaf"))
        names, _, _ = textio.parse_gfa(gfa)
        qn, off, st = textio.parse_gaf(gaf, names)
        path = textio.parse_path(c["args"][c["args"].index("-p") + 1], names)
        _, _, tf, tr = oracle.evaluate_path(path, off, st, False, scores=True)

        def trace(p, B):
            SA, SB, _ = oracle.pair_alignment(p, B)
            return ops_from_gapped(SA, SB)

        assert evalpath_stdout(names, path, qn, off, st, tf, tr, trace) == c["stdout"], c["name"]
        seen += 1
    assert seen >= 18
