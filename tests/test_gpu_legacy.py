"""The non-bitset scoring path (score_paths_kernel, kernels.cu): forced with GFE_FORCE_LEGACY=1, and
reached naturally by a universe whose pair dictionary + node planes do not fit shared memory."""
import os
import subprocess
import sys

import numpy as np
import pytest

from gfalign_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_force_legacy_env_matches_oracle():
    env = dict(os.environ, GFE_FORCE_LEGACY="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "legacy_worker.py")], env=env,
                       capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-3000:]
    assert "legacy ok" in p.stdout


def test_pair_dictionary_overflow_falls_back_and_matches_oracle(engine, oracle):
    """20 000-node universe: the node planes alone take 40 k of the 52 k shared-memory words, the
    graph's ~28 k distinct pairs do not fit the rest -> no dictionary -> position-table kernel."""
    t = synth.make_tangle(20000, 8000, seed=41, minus_frac=0.05, count2_frac=0.05)
    off, st = synth.random_walk_alignments(t.n_nodes, t.edge_from, t.edge_to, 120000,
                                           *synth.HIFI_LENGTHS, seed=42)
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, 48, seed=43, max_len=400)
    engine.set_alignments(off, st, t.n_nodes)
    engine.set_universe(nodes)
    c0 = engine.counters()
    got = engine.score_paths(po, ps, filter=True)
    c1 = engine.counters()
    # prologue + tables + score_paths + exact worklist: the bitset chain would add a second-look launch
    assert c1["kernel_launches"] - c0["kernel_launches"] == 4
    want = oracle.evaluate_paths(po, ps, off, st, True, os.cpu_count() or 1)
    assert np.array_equal(got, want)
    assert int(got.sum()) > 0
    got = engine.score_paths(po, ps, filter=False)
    assert np.array_equal(got, oracle.evaluate_paths(po, ps, off, st, False, os.cpu_count() or 1))


def test_overlong_alignment_is_dropped_by_the_filter_and_flagged_when_it_passes(engine, oracle):
    """An alignment longer than the reference's MAX_N: in search mode the reference drops it at the id
    filter whenever a node of it is missing from the candidate path (src/eval.cpp:81-91), so the table
    is accepted and scored; only a path on which it passes the filter (undefined in the reference:
    dp[1001][1001]) and the unfiltered entry points raise GFE_ERR_LIMIT."""
    import gfalign_b200
    w = synth.make_workload("cfg3", A=20000)
    t = w.tangle
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, 40, seed=5)
    missing = sorted(set(range(t.n_nodes)) - set((ps >> 1).tolist()))
    a, b = int(ps[po[3]]) >> 1, int(ps[po[3] + 1]) >> 1      # two nodes of candidate 3
    # long row 1: names a node no candidate visits -> filtered for every candidate
    assert missing
    long1 = np.array(([a << 1, b << 1] * 700)[:1399] + [missing[0] << 1], dtype=np.uint32)
    off = np.concatenate([w.offsets, [w.offsets[-1] + long1.size]]).astype(np.uint32)
    st = np.concatenate([w.steps, long1]).astype(np.uint32)
    engine.set_alignments(off, st, t.n_nodes)
    engine.set_universe(nodes)
    got = engine.score_paths(po, ps, filter=True)
    assert np.array_equal(got, oracle.evaluate_paths(po, ps, w.offsets, w.steps, True, 4))   # as if the row were absent
    with pytest.raises(gfalign_b200.GfeError) as ex:
        engine.score_paths(po, ps, filter=False)
    assert ex.value.status == -3
    with pytest.raises(gfalign_b200.GfeError):
        engine.score_pairs(ps[po[0]:po[1]])
    # long row 2: only nodes a, b -> passes the filter of candidate 3: reference behaviour undefined -> LIMIT
    long2 = np.array(([a << 1, b << 1] * 700)[:1300], dtype=np.uint32)
    off = np.concatenate([w.offsets, [w.offsets[-1] + long2.size]]).astype(np.uint32)
    st = np.concatenate([w.steps, long2]).astype(np.uint32)
    engine.set_alignments(off, st, t.n_nodes)
    engine.set_universe(nodes)
    with pytest.raises(gfalign_b200.GfeError) as ex:
        engine.score_paths(po, ps, filter=True)
    assert ex.value.status == -3
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)      # and the engine is usable afterwards
    engine.set_universe(nodes)
    assert np.array_equal(engine.score_paths(po, ps, filter=True), got)


def test_search_without_pair_dictionary_falls_back_to_full_scans(engine):
    """A universe too large for the bitset tables has no inverted-index path either: gfe_search must notice
    (GFE_ERR_STATE from the incremental entry) and finish on full scans with the same rows."""
    t = synth.make_tangle(20000, 8000, seed=41, minus_frac=0.05, count2_frac=0.05)
    off, st = synth.random_walk_alignments(t.n_nodes, t.edge_from, t.edge_to, 60000, *synth.HIFI_LENGTHS, seed=42)
    nodes, counts, node_count = synth.node_table(t)
    engine.set_alignments(off, st, t.n_nodes)

    def run():
        rows, stats = engine.search(t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination,
                                    nodes, counts, node_count, max_steps=400, return_all_paths=True)
        return [r.columns(t.names) for r in rows], stats["evaluations"]

    got = run()
    os.environ["GFE_SEARCH_FULLSCAN"] = "1"
    try:
        want = run()
    finally:
        del os.environ["GFE_SEARCH_FULLSCAN"]
    assert got == want and got[1] > 400
