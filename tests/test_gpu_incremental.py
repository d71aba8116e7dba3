"""Incremental search scoring (gfe_score_paths_incremental): a child path's counts from
its parent's, through the inverted node -> alignment index, must equal the full scan and the oracle."""
import os

import numpy as np
import pytest

from gfalign_b200 import synth

pytestmark = pytest.mark.gpu


def _fresh_rows(off, st, root, long_m=16):
    """row mask of the fresh set F: rows containing the root node, rows longer than 16 steps, empty rows"""
    m = np.diff(off.astype(np.int64))
    hit = (st >> 1) == root
    cs = np.concatenate([[0], np.cumsum(hit.astype(np.int64))])
    has_root = (cs[off[1:].astype(np.int64)] - cs[off[:-1].astype(np.int64)]) > 0
    return has_root | (m > long_m) | (m == 0)


def _subtable(off, st, keep):
    m = np.diff(off.astype(np.int64))
    noff = np.zeros(int(keep.sum()) + 1, dtype=np.int64)
    noff[1:] = np.cumsum(m[keep])
    stepkeep = np.repeat(keep, m)
    return noff.astype(np.uint32), st[stepkeep].astype(np.uint32)


def _check_batch(engine, oracle, off, st, n_nodes, nodes, paths, rng, threads=8):
    """every path as (parent: mode 0, child: mode 1 when long enough else mode 0), random base/suffix cut"""
    root = int(paths[0][0]) >> 1
    engine.set_alignments(off, st, n_nodes)
    engine.set_universe(nodes)
    bases, base_off, item_base, suf, suf_off, mode, kind = [], [0], [], [], [0], [], []
    for p in paths:
        n = len(p)
        if n < 2:
            continue
        cut = int(rng.integers(1, n))            # base = p[:cut]; shared by parent and child
        bases.extend(p[:cut])
        base_off.append(len(bases))
        kb = len(base_off) - 2
        # parent: p[:n-1], full
        item_base.append(kb)
        suf.extend(p[cut:n - 1])
        suf_off.append(len(suf))
        mode.append(0)
        kind.append(("parent", p[:n - 1]))
        # child: p, delta when the parent has >= 16 steps
        item_base.append(kb)
        suf.extend(p[cut:])
        suf_off.append(len(suf))
        mode.append(1 if n - 1 >= 16 else 0)
        kind.append(("child", p))
    got = engine.score_paths_incremental(base_off, bases, item_base, suf_off, suf, mode)
    # oracle: full counts and counts over F for every item path
    po = np.zeros(len(kind) + 1, dtype=np.uint32)
    po[1:] = np.cumsum([len(k[1]) for k in kind])
    ps = np.concatenate([np.asarray(k[1], dtype=np.uint32) for k in kind])
    full = oracle.evaluate_paths(po, ps, off, st, True, threads).astype(np.int64)
    foff, fst = _subtable(off, st, _fresh_rows(off, st, root))
    fresh = oracle.evaluate_paths(po, ps, foff, fst, True, threads).astype(np.int64)
    assert np.array_equal(engine.score_paths(po, ps, filter=True).astype(np.int64), full)
    n_delta = 0
    for i in range(0, len(kind), 2):
        par, ch = got[i].astype(np.int64), got[i + 1].astype(np.int64)
        assert np.array_equal(par[:2], full[i]), ("parent total", i)
        assert np.array_equal(par[2:], fresh[i]), ("parent fresh", i)
        assert np.array_equal(ch[2:], fresh[i + 1]), ("child fresh", i)
        if mode[i + 1] == 1:
            total = (par[:2] - par[2:]) + ch[:2] + ch[2:]
            assert np.array_equal(total, full[i + 1]), ("child delta", i, kind[i + 1][1][-3:])
            n_delta += 1
        else:
            assert np.array_equal(ch[:2], full[i + 1]), ("child total", i)
    return n_delta


def test_incremental_equals_full_scan_cfg3(engine, oracle):
    w = synth.make_workload("cfg3", A=60000)
    t = w.tangle
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, 260, seed=31)
    paths = [ps[po[i]:po[i + 1]].tolist() for i in range(260)]
    rng = np.random.default_rng(3)
    n_delta = _check_batch(engine, oracle, w.offsets, w.steps, t.n_nodes, nodes, paths, rng)
    assert n_delta > 150


def test_incremental_revisits_tiny_alphabet(engine, oracle):
    """arbitrary paths over 6 nodes (every node revisited many times, the start-overhang class and the
    exact DP busy), rows of every length incl. > 16 steps and empty rows"""
    rng = np.random.default_rng(9)
    alpha = 6
    lens = rng.choice([0, 1, 2, 3, 4, 5, 7, 9, 16, 17, 33], size=6000)
    off = np.zeros(6001, dtype=np.uint32)
    off[1:] = np.cumsum(lens)
    st = (rng.integers(0, alpha, size=int(off[-1])) * 2 + rng.integers(0, 2, size=int(off[-1]))).astype(np.uint32)
    paths = []
    for _ in range(120):
        n = int(rng.integers(2, 60))
        p = (rng.integers(0, alpha, size=n) * 2 + rng.integers(0, 2, size=n)).astype(np.uint32)
        p[0] = 2 * 3 + (p[0] & 1)                 # every path starts at node 3 (the root)
        paths.append(p.tolist())
    n_delta = _check_batch(engine, oracle, off, st, alpha, np.arange(alpha, dtype=np.uint32), paths, rng, 4)
    assert n_delta > 60


def test_incremental_cfg5_shape_and_wide_batch(engine, oracle):
    """cfg5 shape (400 nodes, rows up to 16 steps, 10 % of the nodes with two visits), 1500 items in one
    batch: many chunks per node, chunks of mixed fill"""
    w = synth.make_workload("cfg5", A=40000)
    t = w.tangle
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, 750, seed=77)
    paths = [ps[po[i]:po[i + 1]].tolist() for i in range(750)]
    rng = np.random.default_rng(4)
    n_delta = _check_batch(engine, oracle, w.offsets, w.steps, t.n_nodes, nodes, paths, rng, os.cpu_count() or 1)
    assert n_delta > 400


def test_incremental_cfg5_real_shape_3m_rows(engine, oracle):
    """cfg5 at a size where the per-node index lists are long (3 M rows of 4-16 steps over 400 nodes: ~75 k
    rows per node, thousands of claim units per chunk) — parents by full scan, children by delta, both against
    the oracle.  The bench runs the whole search on the 50 M-row table against the full-scan driver; this is
    the oracle-side check of the same kernel regime."""
    w = synth.make_workload("cfg5", A=3_000_000)
    t = w.tangle
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, 64, seed=78)
    paths = [ps[po[i]:po[i + 1]].tolist() for i in range(64)]
    rng = np.random.default_rng(5)
    n_delta = _check_batch(engine, oracle, w.offsets, w.steps, t.n_nodes, nodes, paths, rng, os.cpu_count() or 1)
    assert n_delta > 30


def test_search_rows_incremental_equal_full_scan(engine):
    """gfe_search with incremental scoring (default) emits the rows of the same search scored by full
    scans (GFE_SEARCH_FULLSCAN=1), for several speculation widths; paths here reach ~100 steps."""
    w = synth.make_workload("cfg3", A=40000)
    t = w.tangle
    nodes, counts, node_count = synth.node_table(t)
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)

    def run(**kw):
        rows, stats = engine.search(t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination,
                                    nodes, counts, node_count, max_steps=3000, return_all_paths=True, **kw)
        return [r.columns(t.names) for r in rows], stats

    os.environ["GFE_SEARCH_FULLSCAN"] = "1"
    try:
        want, st_full = run()
    finally:
        del os.environ["GFE_SEARCH_FULLSCAN"]
    assert len(want) > 3
    for spec in (0, 1, 64, 4096):
        got, st = run(speculate=spec)
        assert got == want, spec
        assert st["evaluations"] == st_full["evaluations"]
