"""GPU parity at BASELINE.json's full sizes through size-independent properties (the oracle checks a
few candidate paths exactly with all host threads; the rest is linearity over alignment shards,
reverse-complement invariance, row-permutation invariance and agreement of the entry points)."""
import os

import numpy as np
import pytest

from gfalign_b200 import sharded, synth

pytestmark = pytest.mark.gpu


def _rc_rows(off, st):
    """every alignment replaced by its reverse complement"""
    m = np.diff(off.astype(np.int64))
    row = np.repeat(np.arange(m.size), m)
    pos = np.arange(st.size) - off[:-1].astype(np.int64)[row]
    src = off[:-1].astype(np.int64)[row] + (m[row] - 1 - pos)
    return (st[src] ^ 1).astype(np.uint32)


def _permute_rows(off, st, rng):
    m = np.diff(off.astype(np.int64))
    perm = rng.permutation(m.size)
    noff = np.zeros(m.size + 1, dtype=np.int64)
    noff[1:] = np.cumsum(m[perm])
    row = np.repeat(np.arange(m.size), m[perm])
    pos = np.arange(st.size) - noff[:-1][row]
    src = off[:-1].astype(np.int64)[perm][row] + pos
    return noff.astype(np.uint32), st[src].astype(np.uint32)


@pytest.mark.parametrize("name,A,P", [("cfg3", 1_000_000, 1024), ("cfg5", 3_000_000, 256)])
def test_search_scoring_properties_full_size(engine, oracle, name, A, P):
    w = synth.make_workload(name, A=A)
    t = w.tangle
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, P, seed=21)
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)
    engine.set_universe(nodes)
    full = engine.score_paths(po, ps, filter=True)
    assert int(full.sum()) > 0 and full.shape == (P, 2)
    # exact check of a few paths against the CPU oracle (all host threads)
    pick = np.linspace(0, P - 1, 6).astype(int)
    so = np.zeros(pick.size + 1, dtype=np.uint32)
    ss = []
    for i, p in enumerate(pick):
        ss.append(ps[po[p]:po[p + 1]])
        so[i + 1] = so[i] + ss[-1].size
    want = oracle.evaluate_paths(so, np.concatenate(ss), w.offsets, w.steps, True, os.cpu_count() or 1)
    assert np.array_equal(full[pick], want)
    # linearity over alignment shards (the multi-GPU decomposition)
    cuts = sharded.shard_bounds(w.offsets, 3)
    acc = np.zeros_like(full, dtype=np.int64)
    for r in range(3):
        o, s = sharded.take_shard(w.offsets, w.steps, int(cuts[r]), int(cuts[r + 1]))
        engine.set_alignments(o, s, t.n_nodes)
        engine.set_universe(nodes)
        acc += engine.score_paths(po, ps, filter=True)
    assert np.array_equal(acc, full.astype(np.int64))
    # reverse-complementing every alignment changes no verdict (max(T_fw, T_rc) is symmetric)
    engine.set_alignments(w.offsets, _rc_rows(w.offsets, w.steps), t.n_nodes)
    engine.set_universe(nodes)
    assert np.array_equal(engine.score_paths(po, ps, filter=True), full)
    # neither does the order of the rows
    o2, s2 = _permute_rows(w.offsets, w.steps, np.random.default_rng(5))
    engine.set_alignments(o2, s2, t.n_nodes)
    engine.set_universe(nodes)
    assert np.array_equal(engine.score_paths(po, ps, filter=True), full)


def test_evalgfa_evalpath_filter_properties_cfg4_shape(engine, oracle):
    """cfg4 shape (500 k nodes) at 2 M alignments: edge support against the oracle, invariance under
    reverse complement and permutation, checksum; evalPath totals == filtered search scoring of the
    same path; filter == numpy."""
    w = synth.make_workload("cfg4", A=2_000_000)
    t = w.tangle
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)
    ew = engine.edge_support(t.edge_from, t.edge_to)
    assert np.array_equal(ew, oracle.edge_support(w.offsets, w.steps, t.n_nodes, t.edge_from, t.edge_to))
    # every consecutive step pair of a walk on the graph is a link: sum of weights over distinct
    # undirected links == number of step pairs (hairpin links, rule 2k-1, do not occur here)
    key = np.minimum(t.edge_from.astype(np.int64) << 32 | t.edge_to,
                     (t.edge_to.astype(np.int64) ^ 1) << 32 | (t.edge_from ^ 1))
    _, first = np.unique(key, return_index=True)
    m = np.diff(w.offsets.astype(np.int64))
    assert int(ew[first].sum(dtype=np.int64)) == int(np.maximum(m - 1, 0).sum())
    path = (np.arange(1000, dtype=np.uint32) << 1)
    tf, tr, bad, good = engine.score_pairs(path, filter=False)
    assert bad + good == m.size and int((np.maximum(tf, tr) < 0).sum()) == bad
    sel = np.linspace(0, m.size - 1, 300).astype(int)
    for a in sel[:60]:
        B = w.steps[w.offsets[a]:w.offsets[a + 1]]
        assert int(tf[a]) == oracle.pair_score(path, B)
    inset = np.zeros(t.n_nodes, dtype=bool)
    inset[::3] = True
    keep = engine.filter_alignments(np.nonzero(inset)[0], 2)
    ok = inset[w.steps >> 1].astype(np.int64)
    cs = np.concatenate([[0], np.cumsum(ok)])
    want = (cs[w.offsets[1:]] - cs[w.offsets[:-1]] == m) & (m >= 2)
    assert np.array_equal(keep, want)
    # invariances of the edge support
    engine.set_alignments(w.offsets, _rc_rows(w.offsets, w.steps), t.n_nodes)
    assert np.array_equal(engine.edge_support(t.edge_from, t.edge_to), ew)
    o2, s2 = _permute_rows(w.offsets, w.steps, np.random.default_rng(6))
    engine.set_alignments(o2, s2, t.n_nodes)
    assert np.array_equal(engine.edge_support(t.edge_from, t.edge_to), ew)
