"""GPU parity, randomised shapes: batch sizes on both sides of every words-per-chunk step of the
bitset kernel (P = 1 .. 2100), universes that are strict subsets of the graph (alignments stepping
outside it), alphabets from 2 to 300 nodes, planted sub-walks in both orientations, rows of every
bucket length plus the long-row kernel (> 16 steps).  Bit-exact against the oracle, filter on/off."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

P_SIZES = [1, 2, 31, 32, 33, 64, 65, 128, 129, 256, 257, 511, 513, 1024, 1025, 2100]


def _case(seed):
    rng = np.random.default_rng(7000 + seed)
    n_nodes = int(rng.choice([2, 3, 5, 9, 20, 60, 150, 300]))
    L = n_nodes if seed % 3 else max(1, int(n_nodes * 0.7))          # universe: prefix of the nodes
    A = int(rng.integers(500, 3500))
    lens = rng.choice([0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 6, 7, 8, 10, 13, 16, 17, 24, 40], size=A)
    off = np.zeros(A + 1, dtype=np.uint32)
    off[1:] = np.cumsum(lens)
    S = int(off[-1])
    hi = n_nodes if seed % 2 else L                                   # odd seeds step outside the universe
    steps = (rng.integers(0, hi, size=S) * 2 + rng.integers(0, 2, size=S)).astype(np.uint32)
    P = P_SIZES[seed % len(P_SIZES)]
    plen = rng.integers(1, int(rng.choice([4, 12, 40, 120])) + 1, size=P)
    po = np.zeros(P + 1, dtype=np.uint32)
    po[1:] = np.cumsum(plen)
    T = int(po[-1])
    ps = (rng.integers(0, L, size=T) * 2 + rng.integers(0, 2, size=T)).astype(np.uint32)
    for a in range(0, A, 2):                                          # plant sub-walks of the paths
        m = int(off[a + 1] - off[a])
        p = int(rng.integers(0, P))
        n = int(po[p + 1] - po[p])
        if 0 < m <= n:
            s = int(rng.integers(0, n - m + 1))
            w = ps[po[p] + s: po[p] + s + m]
            steps[off[a]:off[a + 1]] = w if a % 4 else (w[::-1] ^ 1)
    return n_nodes, L, off, steps, po, ps


@pytest.mark.parametrize("seed", range(32))
def test_score_paths_random_shapes(engine, oracle, seed):
    n_nodes, L, off, steps, po, ps = _case(seed)
    engine.set_alignments(off, steps, n_nodes)
    engine.set_universe(np.arange(L, dtype=np.uint32))
    for filt in (True, False):
        got = engine.score_paths(po, ps, filter=filt)
        want = oracle.evaluate_paths(po, ps, off, steps, filt, 8)
        assert np.array_equal(got, want), (seed, n_nodes, L, len(po) - 1, filt)
