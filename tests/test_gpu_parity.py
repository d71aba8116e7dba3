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
