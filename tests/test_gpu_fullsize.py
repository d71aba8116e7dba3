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


# ---- BASELINE's configured sizes (cfg5: 50 M alignments, cfg4: 20 M), tables drawn on the device -------

def _need_memory(gib):
    import torch
    free, total = torch.cuda.mem_get_info(0)
    if total < gib * (1 << 30):
        pytest.skip(f"needs a {gib} GiB device")


def _rc_rows_device(off, st):
    """every alignment replaced by its reverse complement, on the device (int32 bit patterns)"""
    import torch
    o = off.to(torch.int64)
    m = o[1:] - o[:-1]
    row = torch.repeat_interleave(torch.arange(m.numel(), device=st.device), m)
    pos = torch.arange(st.numel(), device=st.device) - o[:-1][row]
    src = o[:-1][row] + (m[row] - 1 - pos)
    return (st[src] ^ 1).contiguous()


def test_cfg5_at_the_configured_50M(engine, oracle):
    """cfg5 as BASELINE names it: 400-node dense tangle, 50 M alignments (4.9e8 steps, 32-bit step
    offsets, the P*A-sized worklists).  Oracle on 3 paths over all 50 M rows with every host core;
    linearity over two device-side shards; reverse-complement invariance; host API == device API."""
    import torch
    _need_memory(60)
    t = synth.make_tangle(400, 800, seed=3, minus_frac=0.05, count2_frac=0.10)
    A, P = 50_000_000, 256
    dev = torch.device("cuda", 0)
    d_off, d_st = synth.random_walk_alignments_device(t.n_nodes, t.edge_from, t.edge_to, A, *synth.ONT_LENGTHS,
                                                      seed=1300, device=dev)
    S = int(d_off[-1].item())
    assert S > 4.5e8
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, P, seed=21)
    engine.set_alignments_device(d_off.data_ptr(), d_st.data_ptr(), A, S, t.n_nodes)
    engine.set_universe(nodes)
    full = engine.score_paths(po, ps, filter=True)
    assert full.shape == (P, 2) and int(full.sum()) > 0
    # device entry point agrees and raises no overflow flag at this size
    d_po = torch.from_numpy(po.astype(np.int64)).to(dev).to(torch.int32)
    d_ps = torch.from_numpy(ps.astype(np.int64)).to(dev).to(torch.int32)
    d_counts = torch.zeros(2 * P, dtype=torch.int32, device=dev)
    engine.device_status()
    engine.score_paths_device(d_po.data_ptr(), d_ps.data_ptr(), P, int(ps.size), d_counts.data_ptr(), filter=True)
    assert engine.device_status() == 0
    assert np.array_equal(d_counts.cpu().numpy().view(np.uint32).reshape(P, 2), full)
    # the oracle over the whole table, a few paths, all host cores
    off_h = d_off.cpu().numpy().view(np.uint32)
    st_h = d_st.cpu().numpy().view(np.uint32)
    pick = np.array([0, P // 2, P - 1])
    so = np.zeros(pick.size + 1, dtype=np.uint32)
    ss = []
    for i, p in enumerate(pick):
        ss.append(ps[po[p]:po[p + 1]])
        so[i + 1] = so[i] + ss[-1].size
    want = oracle.evaluate_paths(so, np.concatenate(ss), off_h, st_h, True, os.cpu_count() or 1)
    assert np.array_equal(full[pick], want)
    del off_h, st_h
    # linearity over two shards cut on the device (the multi-GPU decomposition)
    cut = A // 2 + 12345
    acc = np.zeros_like(full, dtype=np.int64)
    for lo, hi in ((0, cut), (cut, A)):
        o = (d_off[lo:hi + 1] - d_off[lo]).contiguous()
        s = d_st[int(d_off[lo].item()):int(d_off[hi].item())].contiguous()
        engine.set_alignments_device(o.data_ptr(), s.data_ptr(), hi - lo, int(s.numel()), t.n_nodes)
        engine.set_universe(nodes)
        acc += engine.score_paths(po, ps, filter=True)
    assert np.array_equal(acc, full.astype(np.int64))
    # reverse-complementing every alignment changes no verdict
    rc = _rc_rows_device(d_off, d_st)
    engine.set_alignments_device(d_off.data_ptr(), rc.data_ptr(), A, S, t.n_nodes)
    engine.set_universe(nodes)
    assert np.array_equal(engine.score_paths(po, ps, filter=True), full)
    engine.set_alignments(np.zeros(1, np.uint32), np.zeros(0, np.uint32), t.n_nodes)   # drop the adopted pointers


def test_cfg4_at_the_configured_20M(engine, oracle):
    """cfg4 as BASELINE names it: 500 k nodes, 20 M alignments.  evalGFA: oracle on a 3 M-row shard,
    checksum over the full table (every step pair of a walk is a link), linearity of the raw traversal
    counters over shards; evalPath (1000-step path): totals, 300 rows against the oracle, scores of
    the traceback entry point; filter == numpy."""
    import torch
    _need_memory(40)
    t = synth.make_tangle(500_000, 200_000, seed=2, minus_frac=0.05, count2_frac=0.0)
    A = 20_000_000
    dev = torch.device("cuda", 0)
    d_off, d_st = synth.random_walk_alignments_device(t.n_nodes, t.edge_from, t.edge_to, A, *synth.HIFI_LENGTHS,
                                                      seed=12, device=dev)
    S = int(d_off[-1].item())
    off_h = d_off.cpu().numpy().view(np.uint32)
    st_h = d_st.cpu().numpy().view(np.uint32)
    m = np.diff(off_h.astype(np.int64))
    engine.set_alignments_device(d_off.data_ptr(), d_st.data_ptr(), A, S, t.n_nodes)
    ew = engine.edge_support(t.edge_from, t.edge_to)
    key = np.minimum(t.edge_from.astype(np.int64) << 32 | t.edge_to,
                     (t.edge_to.astype(np.int64) ^ 1) << 32 | (t.edge_from ^ 1))
    _, first = np.unique(key, return_index=True)
    assert int(ew[first].sum(dtype=np.int64)) == int(np.maximum(m - 1, 0).sum())
    # evalPath on the 1000-step chain path
    path = (np.arange(1000, dtype=np.uint32) << 1)
    tf, tr, bad, good = engine.score_pairs(path, filter=False)
    assert bad + good == A and int((np.maximum(tf, tr) < 0).sum()) == bad
    touching = np.nonzero((tf != -m) | (tr != -m))[0]
    sel = np.concatenate([np.linspace(0, A - 1, 100).astype(np.int64), touching[:200]])
    for a in sel:
        B = st_h[off_h[a]:off_h[a + 1]]
        assert int(tf[a]) == oracle.pair_score(path, B), a
        assert int(tr[a]) == oracle.pair_score(path, (B[::-1] ^ 1).astype(np.uint32)), a
    cols, score = engine.trace_pairs(path, int(touching[0]), 64, np.zeros(64, np.uint8), off_h)
    assert np.array_equal(score, tf[int(touching[0]):int(touching[0]) + 64])
    # filter mode
    inset = np.zeros(t.n_nodes, dtype=bool)
    inset[::3] = True
    keep = engine.filter_alignments(np.nonzero(inset)[0], 2)
    ok = inset[st_h >> 1].astype(np.int64)
    cs = np.concatenate([[0], np.cumsum(ok)])
    want = (cs[off_h[1:].astype(np.int64)] - cs[off_h[:-1].astype(np.int64)] == m) & (m >= 2)
    assert np.array_equal(keep, want)
    # shards: oracle on the first 3 M rows; raw traversal counters add up to the full table's weights
    cut = 3_000_000
    total = torch.zeros(2 * t.edge_from.size, dtype=torch.int64)
    for lo, hi in ((0, cut), (cut, A)):
        o = (d_off[lo:hi + 1] - d_off[lo]).contiguous()
        s = d_st[int(d_off[lo].item()):int(d_off[hi].item())].contiguous()
        engine.set_alignments_device(o.data_ptr(), s.data_ptr(), hi - lo, int(s.numel()), t.n_nodes)
        if lo == 0:
            assert np.array_equal(engine.edge_support(t.edge_from, t.edge_to),
                                  oracle.edge_support(off_h[:cut + 1], st_h[:int(off_h[cut])], t.n_nodes,
                                                      t.edge_from, t.edge_to))
        d = torch.zeros(2 * t.edge_from.size, dtype=torch.int32, device=dev)
        engine.edge_traversals_device(t.edge_from, t.edge_to, d.data_ptr())
        torch.cuda.synchronize()
        total += d.cpu().to(torch.int64)
    assert np.array_equal(engine.edge_weights_from_counts(t.edge_from, t.edge_to, total.numpy().astype(np.uint32)), ew)
    engine.set_alignments(np.zeros(1, np.uint32), np.zeros(0, np.uint32), t.n_nodes)
