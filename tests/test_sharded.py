"""CPU suite, world_size 2 over gloo: the multi-GPU host logic (alignment shards, one all-reduce of
the count records, lock-step canonical search, sharded edge support).  The per-shard scorer is the
oracle here; on the GPU box it is Engine.score_paths (tests/test_gpu_parity.py, bench.py)."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from gfalign_b200 import sharded, synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _traversals_numpy(off, st, ef, et):
    """2*E raw counters: events equal to L line i, and to its reverse complement."""
    c = np.zeros(2 * ef.size, dtype=np.uint32)
    inner = np.ones(st.size, dtype=bool)
    inner[off[1:-1][off[1:-1] < st.size]] = False        # no pair across a row boundary
    x, y = st[:-1][inner[1:]], st[1:][inner[1:]]
    key = x.astype(np.int64) << 32 | y
    for i in range(ef.size):
        c[2 * i] = np.count_nonzero(key == (int(ef[i]) << 32 | int(et[i])))
        c[2 * i + 1] = np.count_nonzero(key == (int(et[i] ^ 1) << 32 | int(ef[i] ^ 1)))
    return c


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from oracle.binding import Oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = Oracle()
    w = synth.make_workload("cfg3", A=4000)
    t = w.tangle
    cuts = sharded.shard_bounds(w.offsets, world)
    off, st = sharded.take_shard(w.offsets, w.steps, int(cuts[rank]), int(cuts[rank + 1]))
    scorer = sharded.ShardedScorer(lambda po, ps: orc.evaluate_paths(po, ps, off, st, True, 1))
    po, ps = synth.candidate_paths(t, 12, seed=2)
    counts = scorer(po, ps)
    nodes, cnt, node_count = synth.node_table(t)
    rows, stats = sharded.sharded_search(scorer, t, nodes, cnt, node_count, max_steps=40, speculate=32)
    ew = sharded.sharded_edge_support(lambda: _traversals_numpy(off, st, t.edge_from, t.edge_to),
                                      t.edge_from, t.edge_to)
    q.put((rank, counts.tolist(), [r.columns(t.names) for r in rows], ew.tolist(), int(cuts[rank + 1] - cuts[rank])))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_bounds_cover_and_balance():
    rng = np.random.default_rng(0)
    m = rng.integers(0, 30, size=1000)
    off = np.concatenate([[0], np.cumsum(m)]).astype(np.uint32)
    for world in (1, 2, 3, 8):
        cuts = sharded.shard_bounds(off, world)
        assert cuts[0] == 0 and cuts[-1] == 1000 and np.all(np.diff(cuts) >= 0)
        load = [int(off[cuts[r + 1]] - off[cuts[r]]) + int(cuts[r + 1] - cuts[r]) for r in range(world)]
        assert max(load) - min(load) <= 62
    assert sharded.shard_bounds(np.zeros(1, np.uint32), 4).tolist() == [0, 0, 0, 0, 0]


@pytest.mark.timeout(300)
def test_two_rank_gloo_matches_single(oracle):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=240) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    w = synth.make_workload("cfg3", A=4000)
    t = w.tangle
    po, ps = synth.candidate_paths(t, 12, seed=2)
    want = oracle.evaluate_paths(po, ps, w.offsets, w.steps, True, 1)
    import gfalign_b200
    nodes, cnt, node_count = synth.node_table(t)
    rows, _ = gfalign_b200.search_with_scorer(
        lambda a, b: oracle.evaluate_paths(a, b, w.offsets, w.steps, True, 1), t.adj_offsets,
        t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination, nodes, cnt, node_count,
        max_steps=40, speculate=32)
    ew = oracle.edge_support(w.offsets, w.steps, t.n_nodes, t.edge_from, t.edge_to)
    assert sum(g[4] for g in got) == 4000 and all(g[4] > 0 for g in got)
    for g in got:
        assert g[1] == want.tolist()
        assert g[2] == [r.columns(t.names) for r in rows]
        assert g[3] == ew.tolist()
