"""GPU parity: the CUDA path through the C ABI vs the CPU oracle on the same seeded inputs."""
import numpy as np
import pytest

from gfalign_b200 import synth

pytestmark = pytest.mark.gpu


def _rand_csr(rng, A, n_nodes, lens):
    m = rng.choice(lens, size=A)
    off = np.zeros(A + 1, dtype=np.uint32)
    off[1:] = np.cumsum(m)
    steps = (rng.integers(0, n_nodes, size=int(off[-1])) * 2 + rng.integers(0, 2, size=int(off[-1]))).astype(np.uint32)
    return off, steps


def test_score_paths_cfg3_small(engine, oracle):
    w = synth.make_workload("cfg3", A=30000)
    t = w.tangle
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, 96, seed=3)
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)
    engine.set_universe(nodes)
    for filt in (True, False):
        got = engine.score_paths(po, ps, filter=filt)
        want = oracle.evaluate_paths(po, ps, w.offsets, w.steps, filt, 4)
        assert np.array_equal(got, want), filt


@pytest.mark.parametrize("alpha,seed", [(2, 0), (3, 1), (5, 2), (12, 3)])
def test_score_paths_tiny_alphabet_exact_fallback(engine, oracle, alpha, seed):
    """Tiny alphabets make the ambiguous class (suffix of B == prefix of A) and repeated nodes
    common, so the exact-DP worklist kernels carry most verdicts."""
    rng = np.random.default_rng(seed)
    off, steps = _rand_csr(rng, 4000, alpha, [0, 1, 2, 3, 4, 5, 7, 9, 17, 33, 40])
    P = 40
    plen = rng.integers(1, 30, size=P)
    po = np.zeros(P + 1, dtype=np.uint32)
    po[1:] = np.cumsum(plen)
    ps = (rng.integers(0, alpha, size=int(po[-1])) * 2 + rng.integers(0, 2, size=int(po[-1]))).astype(np.uint32)
    engine.set_alignments(off, steps, alpha)
    engine.set_universe(np.arange(alpha, dtype=np.uint32))
    for filt in (True, False):
        got = engine.score_paths(po, ps, filter=filt)
        want = oracle.evaluate_paths(po, ps, off, steps, filt, 4)
        assert np.array_equal(got, want), (alpha, filt)
    assert engine.counters()["exact_pairs"] > 0


def test_score_pairs_scores(engine, oracle):
    w = synth.make_workload("cfg3", A=20000)
    t = w.tangle
    po, ps = synth.candidate_paths(t, 6, seed=9)
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)
    for p in range(6):
        path = ps[po[p]:po[p + 1]]
        tf, tr, bad, good = engine.score_pairs(path, filter=False)
        ob, og, otf, otr = oracle.evaluate_path(path, w.offsets, w.steps, False, scores=True)
        assert np.array_equal(tf, otf) and np.array_equal(tr, otr)
        assert (bad, good) == (ob, og)
        _, _, bad2, good2 = engine.score_pairs(path, filter=False, want_scores=False)
        assert (bad2, good2) == (ob, og)


def test_score_pairs_long_alignments(engine, oracle):
    rng = np.random.default_rng(5)
    off, steps = _rand_csr(rng, 300, 6, [1, 5, 33, 34, 64, 100, 257])
    engine.set_alignments(off, steps, 6)
    path = (rng.integers(0, 6, size=120) * 2 + rng.integers(0, 2, size=120)).astype(np.uint32)
    tf, tr, bad, good = engine.score_pairs(path)
    ob, og, otf, otr = oracle.evaluate_path(path, off, steps, False, scores=True)
    assert np.array_equal(tf, otf) and np.array_equal(tr, otr) and (bad, good) == (ob, og)


def test_edge_support(engine, oracle):
    w = synth.make_workload("cfg3", A=50000)
    t = w.tangle
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)
    # add duplicates, a twin line and hairpins to the L lines
    ef = np.concatenate([t.edge_from, t.edge_from[:5], t.edge_to[5:9] ^ 1, [6, 7]]).astype(np.uint32)
    et = np.concatenate([t.edge_to, t.edge_to[:5], t.edge_from[5:9] ^ 1, [7, 6]]).astype(np.uint32)
    got = engine.edge_support(ef, et)
    want = oracle.edge_support(w.offsets, w.steps, t.n_nodes, ef, et)
    assert np.array_equal(got, want)
    assert got.sum() > 0


def test_edge_support_hairpins_random(engine, oracle):
    rng = np.random.default_rng(11)
    off, steps = _rand_csr(rng, 5000, 4, [1, 2, 3, 6])
    engine.set_alignments(off, steps, 4)
    ef = rng.integers(0, 8, size=40).astype(np.uint32)
    et = rng.integers(0, 8, size=40).astype(np.uint32)
    got = engine.edge_support(ef, et)
    want = oracle.edge_support(off, steps, 4, ef, et)
    assert np.array_equal(got, want)
    # the sharded flow: raw device counters of two alignment shards, summed, then the host rule
    import torch
    from gfalign_b200 import sharded
    total = torch.zeros(2 * ef.size, dtype=torch.int64)
    for lo, hi in ((0, 2000), (2000, 5000)):
        so, ss = sharded.take_shard(off, steps, lo, hi)
        engine.set_alignments(so, ss, 4)
        d = torch.zeros(2 * ef.size, dtype=torch.int32, device="cuda")
        engine.edge_traversals_device(ef, et, d.data_ptr())
        torch.cuda.synchronize()
        total += d.cpu().to(torch.int64)
    w = engine.edge_weights_from_counts(ef, et, total.numpy().astype(np.uint32))
    assert np.array_equal(w, want)


# ---- against the golden fixtures generated from the compiled reference --------------------------

def _cases():
    import gzip, json, os
    p = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cli.json.gz")
    with gzip.open(p, "rt") as f:
        return json.load(f)["cases"]


def test_golden_search_rows(engine):
    """gfe_search (CUDA-scored) reproduces the reference CLI's emitted rows, in order."""
    from test_oracle import _search_inputs, search_rows_text
    seen = 0
    for c in _cases():
        if "_search" not in c["name"]:
            continue
        t, off, st, kw = _search_inputs(c)
        nodes, counts, node_count = synth.node_table(t)
        engine.set_alignments(off, st, t.n_nodes)
        for spec in (0, 1, 1024):
            rows, stats = engine.search(t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source,
                                        t.destination, nodes, counts, node_count, speculate=spec, **kw)
            assert search_rows_text(rows, stats, t.names, kw) == c["stdout"].splitlines(), (c["name"], spec)
        seen += 1
    assert seen >= 13


def test_golden_evalpath_scores(engine):
    import textio
    from test_oracle import _evalpath_expect
    seen = 0
    for c in _cases():
        if "_evalpath_" not in c["name"]:
            continue
        gfa = next(v for k, v in c["files"].items() if k.endswith(".gfa"))
        gaf = next(v for k, v in c["files"].items() if k.endswith(".gaf"))
        names, _, _ = textio.parse_gfa(gfa)
        _, off, st = textio.parse_gaf(gaf, names)
        path = textio.parse_path(c["args"][c["args"].index("-p") + 1], names)
        engine.set_alignments(off, st, len(names))
        tf, tr, bad, good = engine.score_pairs(path, filter=False)
        scores, totals = _evalpath_expect(c)
        assert np.maximum(tf, tr).tolist() == scores, c["name"]
        uniques = len(set((path >> 1).tolist()))
        assert [bad, good, bad - good - uniques, path.size, uniques] == totals, c["name"]
        seen += 1
    assert seen >= 18


def test_golden_evalgfa_weights(engine):
    import textio
    seen = 0
    for c in _cases():
        if not c["name"].endswith("_evalgfa"):
            continue
        gfa = next(v for k, v in c["files"].items() if k.endswith(".gfa"))
        gaf = next(v for k, v in c["files"].items() if k.endswith(".gaf"))
        names, ef, et = textio.parse_gfa(gfa)
        _, off, st = textio.parse_gaf(gaf, names)
        engine.set_alignments(off, st, len(names))
        want = [int(l.split("RC:i:")[1].split("\t")[0]) for l in c["out_L_lines"]]
        assert engine.edge_support(ef, et).tolist() == want, c["name"]
        seen += 1
    assert seen >= 6


def test_empty_and_ragged_inputs(engine, oracle):
    # no alignments at all; alignments without steps; a single-step path
    engine.set_alignments(np.zeros(1, np.uint32), np.zeros(0, np.uint32), 4)
    engine.set_universe(np.arange(4, dtype=np.uint32))
    po, ps = np.array([0, 1, 3], np.uint32), np.array([0, 2, 5], np.uint32)
    assert engine.score_paths(po, ps).tolist() == [[0, 0], [0, 0]]
    off = np.array([0, 0, 2, 2, 3], np.uint32)
    st = np.array([2, 5, 0], np.uint32)
    engine.set_alignments(off, st, 4)
    engine.set_universe(np.arange(4, dtype=np.uint32))
    for filt in (True, False):
        assert np.array_equal(engine.score_paths(po, ps, filter=filt),
                              oracle.evaluate_paths(po, ps, off, st, filt, 1))
    assert engine.edge_support(np.zeros(0, np.uint32), np.zeros(0, np.uint32)).size == 0


def test_limits_rejected(engine):
    import gfalign_b200
    engine.set_alignments(np.array([0, 1], np.uint32), np.array([0], np.uint32), 2)
    engine.set_universe(np.arange(2, dtype=np.uint32))
    long_path = np.zeros(1001, np.uint32)
    with pytest.raises(gfalign_b200.GfeError):
        engine.score_paths(np.array([0, 1001], np.uint32), long_path)
    # an over-long alignment is accepted into the table (the id filter may drop it, test_gpu_legacy.py) but
    # every unfiltered entry point refuses it: the reference's dp[MAX_N][MAX_N] cannot hold it
    engine.set_alignments(np.array([0, 1001], np.uint32), np.zeros(1001, np.uint32), 2)
    with pytest.raises(gfalign_b200.GfeError):
        engine.score_pairs(np.array([0, 2], np.uint32))
    engine.set_universe(np.arange(2, dtype=np.uint32))
    with pytest.raises(gfalign_b200.GfeError):
        engine.score_paths(np.array([0, 2], np.uint32), np.array([0, 2], np.uint32), filter=False)
    with pytest.raises(gfalign_b200.GfeError):   # node outside the universe
        engine.set_alignments(np.array([0, 1], np.uint32), np.array([0], np.uint32), 4)
        engine.set_universe(np.array([0, 1], np.uint32))
        engine.score_paths(np.array([0, 1], np.uint32), np.array([6], np.uint32))


def test_max_size_path_evalpath(engine, oracle):
    """n = 1000 (the reference's MAX_N - 1) x alignments up to 40 steps, scores bit-exact."""
    rng = np.random.default_rng(3)
    n_nodes = 600
    path = (rng.permutation(n_nodes)[:500].repeat(2)[:1000] * 2 + rng.integers(0, 2, 1000)).astype(np.uint32)
    off, steps = _rand_csr(rng, 3000, n_nodes, [1, 2, 3, 5, 8, 40])
    # plant sub-walks of the path so occurrences and near-occurrences exist
    for a in range(0, 3000, 3):
        m = off[a + 1] - off[a]
        s = int(rng.integers(0, 1000 - m))
        steps[off[a]:off[a + 1]] = path[s:s + m] if a % 2 else (path[s:s + m][::-1] ^ 1)
    engine.set_alignments(off, steps, n_nodes)
    tf, tr, bad, good = engine.score_pairs(path)
    ob, og, otf, otr = oracle.evaluate_path(path, off, steps, False, scores=True)
    assert np.array_equal(tf, otf) and np.array_equal(tr, otr) and (bad, good) == (ob, og)


@pytest.mark.parametrize("seed", range(8))
def test_repeated_nodes_stress(engine, oracle, seed):
    """Paths that revisit nodes many times (and take self-reverse-complement pairs): the bitset
    kernel's broken-triple certification and start-overhang test against the oracle."""
    rng = np.random.default_rng(100 + seed)
    alpha = int(rng.integers(2, 7))
    off, steps = _rand_csr(rng, 6000, alpha, [1, 2, 2, 3, 3, 4, 4, 5, 6, 8, 12, 16])
    P = int(rng.integers(33, 200))
    plen = rng.integers(1, 60, size=P)
    po = np.zeros(P + 1, dtype=np.uint32)
    po[1:] = np.cumsum(plen)
    ps = (rng.integers(0, alpha, size=int(po[-1])) * 2 + rng.integers(0, 2, size=int(po[-1]))).astype(np.uint32)
    # plant sub-walks of the paths (both orientations) so that occurrences are common
    for a in range(0, 6000, 2):
        m = int(off[a + 1] - off[a])
        p = int(rng.integers(0, P))
        n = int(po[p + 1] - po[p])
        if m <= n:
            s = int(rng.integers(0, n - m + 1))
            w = ps[po[p] + s: po[p] + s + m]
            steps[off[a]:off[a + 1]] = w if a % 4 else (w[::-1] ^ 1)
    engine.set_alignments(off, steps, alpha)
    engine.set_universe(np.arange(alpha, dtype=np.uint32))
    for filt in (True, False):
        got = engine.score_paths(po, ps, filter=filt)
        want = oracle.evaluate_paths(po, ps, off, steps, filt, 8)
        assert np.array_equal(got, want), (seed, alpha, filt)


# ---- evalPath detail (traceback columns) and filter mode ------------------------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("alpha,n_hi,m_hi,seed", [(3, 12, 8, 1), (6, 40, 17, 2), (4, 30, 60, 3),
                                                  (12, 200, 10, 4)])
def test_trace_pairs_matches_oracle(engine, oracle, alpha, n_hi, m_hi, seed):
    from evalpath_text import ops_from_gapped
    rng = np.random.default_rng(seed)
    lens = rng.integers(0, m_hi + 1, 400)
    off, st = _rand_csr(rng, 400, alpha, lens)
    engine.set_alignments(off, st, alpha)
    n = int(rng.integers(1, n_hi + 1))
    path = (rng.integers(0, alpha, n) * 2 + rng.integers(0, 2, n)).astype(np.uint32)
    use_rc = rng.integers(0, 2, 400).astype(np.uint8)
    for first, count in ((0, 400), (37, 101), (399, 1), (5, 0)):
        cols, score = engine.trace_pairs(path, first, count, use_rc[first:first + count], off)
        for k in range(count):
            B = st[off[first + k]:off[first + k + 1]]
            if use_rc[first + k]:
                B = (B[::-1] ^ 1).astype(np.uint32)
            SA, SB, sc = oracle.pair_alignment(path, B)
            assert cols[k].tolist() == ops_from_gapped(SA, SB), (first, k)
            assert int(score[k]) == sc == oracle.pair_score(path, B)


@pytest.mark.gpu
def test_golden_evalpath_stdout_through_engine(engine):
    """Full evalPath stdout (gapped strings, scores, totals) of the golden cases from
    gfe_score_pairs + gfe_trace_pairs."""
    import textio
    from evalpath_text import evalpath_stdout
    seen = 0
    for c in _cases():
        if "_evalpath_" not in c["name"]:
            continue
        gfa = next(v for k, v in c["files"].items() if k.endswith(".gfa"))
        gaf = next(v for k, v in c["files"].items() if k.endswith(".gaf"))
        names, _, _ = textio.parse_gfa(gfa)
        qn, off, st = textio.parse_gaf(gaf, names)
        path = textio.parse_path(c["args"][c["args"].index("-p") + 1], names)
        engine.set_alignments(off, st, len(names))
        tf, tr, _, _ = engine.score_pairs(path, filter=False)
        use_rc = (~(tf > tr)).astype(np.uint8)
        cols, score = engine.trace_pairs(path, 0, len(qn), use_rc, off)
        assert np.array_equal(score, np.where(use_rc == 1, tr, tf))
        it = iter(cols)
        text = evalpath_stdout(names, path, qn, off, st, tf, tr, lambda p, B: next(it).tolist())
        assert text == c["stdout"], c["name"]
        seen += 1
    assert seen >= 18


@pytest.mark.gpu
def test_trace_pairs_max_sizes(engine, oracle):
    """n = 1000 path against alignments up to 1000 steps (long-row scratch, several decision words)."""
    from evalpath_text import ops_from_gapped
    rng = np.random.default_rng(9)
    lens = np.array([0, 1, 16, 17, 32, 33, 200, 1000, 5, 999])
    off = np.zeros(lens.size + 1, dtype=np.uint32)
    off[1:] = np.cumsum(lens)
    st = (rng.integers(0, 5, int(off[-1])) * 2 + rng.integers(0, 2, int(off[-1]))).astype(np.uint32)
    path = (rng.integers(0, 5, 1000) * 2 + rng.integers(0, 2, 1000)).astype(np.uint32)
    st2 = st.copy()
    st2[off[7]:off[7] + 400] = path[300:700]          # a long exact sub-walk
    engine.set_alignments(off, st2, 5)
    cols, score = engine.trace_pairs(path, 0, lens.size, np.zeros(lens.size, np.uint8), off)
    for k in range(lens.size):
        SA, SB, sc = oracle.pair_alignment(path, st2[off[k]:off[k + 1]])
        assert cols[k].tolist() == ops_from_gapped(SA, SB), k
        assert int(score[k]) == sc


@pytest.mark.gpu
def test_filter_alignments(engine):
    rng = np.random.default_rng(11)
    lens = rng.choice([0, 1, 2, 3, 5, 9, 300], 5000, p=[.02, .1, .2, .3, .2, .17, .01])
    off, st = _rand_csr(rng, 5000, 50, lens)
    engine.set_alignments(off, st, 50)
    for nodes, mn in ((np.arange(0, 50, 2), 0), (np.arange(50), 3), (np.array([], dtype=np.uint32), 0),
                      (rng.permutation(50)[:45], 2), (np.arange(50), -1)):
        keep = engine.filter_alignments(nodes, mn)
        inset = np.zeros(50, dtype=bool)
        inset[np.asarray(nodes, dtype=np.int64)] = True
        want = np.array([bool(inset[st[off[a]:off[a + 1]] >> 1].all()) and (off[a + 1] - off[a]) >= mn
                         for a in range(5000)])
        assert np.array_equal(keep, want), (nodes.size, mn)


# ---- sharded-path plumbing that also runs on one GPU ----------------------------------------------

def test_peer_sum_single_rank_is_identity_and_epochs_alternate_halves():
    """gfe_peer_sum_u32 with world = 1: the rank publishes to and reads from its own buffer; both
    halves and growing epochs; n < n_cap leaves the rest of `out` untouched."""
    import torch
    from gfalign_b200 import capi
    dev = torch.device("cuda", 0)
    n_cap = 3000
    buf = torch.zeros(2 * n_cap + capi.PEER_MAX, dtype=torch.int32, device=dev)
    out = torch.full((n_cap,), -1, dtype=torch.int32, device=dev)
    g = torch.Generator().manual_seed(5)
    for epoch in range(1, 8):
        n = [1, 2, 2047, 2048, 2049, 3000, 17][epoch - 1]
        rec = torch.randint(0, 1 << 30, (n,), generator=g, dtype=torch.int32).to(dev)
        half = n_cap * (epoch & 1)
        buf[half:half + n].copy_(rec)
        out.fill_(-1)
        capi.peer_sum_u32([buf.data_ptr()], 0, n_cap, n, epoch, out.data_ptr(),
                          torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert torch.equal(out[:n], rec), epoch
        assert bool((out[n:] == -1).all()), epoch
        assert int(buf[2 * n_cap].item()) == epoch
    with pytest.raises(capi.GfeError):
        capi.peer_sum_u32([buf.data_ptr()], 0, n_cap, n_cap + 1, 9, out.data_ptr())


def test_device_status_word_matches_host_status(engine):
    import torch
    w = synth.make_workload("cfg3", A=5000)
    t = w.tangle
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, 8, seed=1)
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)
    engine.set_universe(nodes[:max(2, len(nodes) // 2)])       # half the universe: some steps fall outside
    dev = torch.device("cuda", 0)
    d_po = torch.from_numpy(po.astype(np.int64)).to(dev).to(torch.int32)
    d_ps = torch.from_numpy(ps.astype(np.int64)).to(dev).to(torch.int32)
    d_counts = torch.zeros(2 * 8, dtype=torch.int32, device=dev)
    flag = torch.full((1,), 77, dtype=torch.int32, device=dev)
    engine.score_paths_device(d_po.data_ptr(), d_ps.data_ptr(), 8, int(ps.size), d_counts.data_ptr(), filter=True)
    engine.device_status_to(flag.data_ptr())
    torch.cuda.synchronize()
    assert int(flag.item()) & 1          # a step outside the universe was seen
    assert engine.device_status() == 0   # reported once: the device-side report cleared it
    engine.score_paths_device(d_po.data_ptr(), d_ps.data_ptr(), 8, int(ps.size), d_counts.data_ptr(), filter=True)
    assert engine.device_status() & 1    # and the host-side report sees a new one


def _world1_exchange(n_cap):
    """A one-rank exchange buffer: plain device memory, the rank publishes to and reads from itself."""
    import torch
    from gfalign_b200 import capi
    buf = torch.zeros(2 * n_cap + capi.PEER_MAX, dtype=torch.int32, device="cuda")
    return buf, capi.Exchange([buf.data_ptr()], 0, n_cap)


def test_sharded_entry_world1_scores_and_search_rows(engine, oracle):
    """gfe_score_paths_sharded / gfe_search_sharded with world = 1 (the code path every rank of a
    multi-GPU run takes: accumulate into the exchange buffer, status slot, peer sum, one copy back):
    counts == oracle, search rows == the reference CLI's on every golden search case."""
    from test_oracle import _search_inputs, search_rows_text
    buf, x = _world1_exchange(1 << 18)
    w = synth.make_workload("cfg3", A=30000)
    t = w.tangle
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, 200, seed=8)
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)
    engine.set_universe(nodes)
    for _ in range(3):     # both halves of the buffer, growing epochs
        got = engine.score_paths_sharded(x, po, ps, filter=True)
        assert np.array_equal(got, oracle.evaluate_paths(po, ps, w.offsets, w.steps, True, 4))
    assert np.array_equal(got, engine.score_paths(po, ps, filter=True))
    seen = 0
    for c in _cases():
        if "_search" not in c["name"]:
            continue
        tt, off, st, kw = _search_inputs(c)
        nd, cn, node_count = synth.node_table(tt)
        engine.set_alignments(off, st, tt.n_nodes)
        rows, stats = engine.search_sharded(x, tt.adj_offsets, tt.adj_target, tt.adj_or0, tt.adj_or1,
                                            tt.source, tt.destination, nd, cn, node_count, **kw)
        assert search_rows_text(rows, stats, tt.names, kw) == c["stdout"].splitlines(), c["name"]
        seen += 1
    assert seen >= 13
    # deep paths (deltas through the sharded incremental entry): same rows as the single-engine search
    nd, cn, node_count = synth.node_table(t)
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)
    kw = dict(max_steps=2500, return_all_paths=True)
    rows_s, st_s = engine.search_sharded(x, t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source,
                                         t.destination, nd, cn, node_count, **kw)
    rows_1, st_1 = engine.search(t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination,
                                 nd, cn, node_count, **kw)
    assert [r.columns(t.names) for r in rows_s] == [r.columns(t.names) for r in rows_1] and len(rows_1) > 2
    assert st_s["evaluations"] == st_1["evaluations"]
    # candidate-batch sharding with one rank (it owns every subtree): device-side composition per item
    x.set_batch_sharding(True)
    rows_b, st_b = engine.search_sharded(x, t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source,
                                         t.destination, nd, cn, node_count, **kw)
    assert [r.columns(t.names) for r in rows_b] == [r.columns(t.names) for r in rows_1]
    assert np.array_equal(engine.score_paths_sharded(x, po, ps, filter=True), engine.score_paths(po, ps, filter=True))
    x.close()


def test_device_status_is_sticky_across_batches_and_cleared_by_the_report(engine):
    """A bad step in batch 1 must still be reported after a clean batch 2 (flags of all batches since
    the last report), and the report clears it."""
    import torch
    w = synth.make_workload("cfg3", A=5000)
    t = w.tangle
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, 8, seed=1)
    engine.set_alignments(w.offsets, w.steps, t.n_nodes)
    dev = torch.device("cuda", 0)
    d_po = torch.from_numpy(po.astype(np.int64)).to(dev).to(torch.int32)
    d_ps = torch.from_numpy(ps.astype(np.int64)).to(dev).to(torch.int32)
    d_counts = torch.zeros(2 * 8, dtype=torch.int32, device=dev)
    engine.set_universe(nodes)
    engine.device_status()                                        # clear whatever earlier tests left
    bad = d_ps.clone()
    bad[1] = int((t.n_nodes + 5) << 1)                           # a node outside the universe
    engine.score_paths_device(d_po.data_ptr(), bad.data_ptr(), 8, int(ps.size), d_counts.data_ptr(), filter=True)
    engine.score_paths_device(d_po.data_ptr(), d_ps.data_ptr(), 8, int(ps.size), d_counts.data_ptr(), filter=True)
    assert engine.device_status() & 1
    assert engine.device_status() == 0
    engine.score_paths_device(d_po.data_ptr(), d_ps.data_ptr(), 8, int(ps.size), d_counts.data_ptr(), filter=True)
    assert engine.device_status() == 0
    assert np.array_equal(d_counts.cpu().numpy().view(np.uint32).reshape(8, 2), engine.score_paths(po, ps, True))


def test_multiplicity_compressed_view_equals_plain_view(engine, oracle):
    """Full scans read the deduplicated view (distinct walks up to reverse complement, weights 2^k) by
    default; the counts equal those of the plain view and of the oracle, with and without the filter,
    incl. a table where one walk repeats 70 000 times (classes up to k = 16) and the second-look /
    exact lists (tiny alphabet: heavy multiplicities meet the exact DP)."""
    w = synth.make_workload("cfg3", A=200000)
    t = w.tangle
    nodes, _, _ = synth.node_table(t)
    po, ps = synth.candidate_paths(t, 300, seed=12)
    rep = np.tile(w.steps[w.offsets[5]:w.offsets[6]], 70000)
    m5 = int(w.offsets[6] - w.offsets[5])
    off = np.concatenate([w.offsets.astype(np.int64), w.offsets[-1] + m5 * np.arange(1, 70001)]).astype(np.uint32)
    st = np.concatenate([w.steps, rep]).astype(np.uint32)
    engine.set_alignments(off, st, t.n_nodes)
    res = {}
    for on in (True, False):
        engine.set_dedup(on)
        engine.set_universe(nodes)
        s = engine.dedup_stats()
        assert s["in_use"] == on
        if on:
            assert s["rows"] == off.size - 1 and s["distinct_walks"] < s["rows"] // 3 and s["view_rows"] < s["rows"] // 2
        res[on] = [engine.score_paths(po, ps, filter=f) for f in (True, False)]
    engine.set_dedup(True)
    for k in range(2):
        assert np.array_equal(res[True][k], res[False][k])
    pick = np.linspace(0, 299, 24).astype(int)
    so = np.zeros(pick.size + 1, dtype=np.uint32)
    so[1:] = np.cumsum([po[p + 1] - po[p] for p in pick])
    ss = np.concatenate([ps[po[p]:po[p + 1]] for p in pick])
    assert np.array_equal(res[True][0][pick], oracle.evaluate_paths(so, ss, off, st, True, 8))
    assert np.array_equal(res[True][1][pick], oracle.evaluate_paths(so, ss, off, st, False, 8))
    # tiny alphabet
    rng = np.random.default_rng(2)
    off2, st2 = _rand_csr(rng, 30000, 4, [1, 2, 3, 4, 5, 7, 16, 17, 20])
    P = 48
    plen = rng.integers(1, 30, size=P)
    po2 = np.zeros(P + 1, dtype=np.uint32)
    po2[1:] = np.cumsum(plen)
    ps2 = (rng.integers(0, 4, size=int(po2[-1])) * 2 + rng.integers(0, 2, size=int(po2[-1]))).astype(np.uint32)
    engine.set_alignments(off2, st2, 4)
    engine.set_universe(np.arange(4, dtype=np.uint32))
    assert engine.dedup_stats()["in_use"]
    for filt in (True, False):
        assert np.array_equal(engine.score_paths(po2, ps2, filter=filt),
                              oracle.evaluate_paths(po2, ps2, off2, st2, filt, 8)), filt
