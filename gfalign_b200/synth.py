"""Seeded synthetic workloads of the shapes BASELINE.json names, as packed arrays
for the engine and as GFA / GAF / nodelist text for the CLIs.

Encoding: oriented step = node << 1 | (orientation == '-').
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np


@dataclass
class Tangle:
    names: List[str]
    edge_from: np.ndarray          # L lines, file order, oriented steps
    edge_to: np.ndarray
    source: int
    destination: int
    nodelist: List[Tuple[int, int]]  # (node, count) as written to the nodelist file
    # reference adjacency (InSequences::buildEdgeGraph, gfalibs/src/gfa.cpp:1076-1097)
    adj_offsets: np.ndarray = field(default=None)
    adj_target: np.ndarray = field(default=None)
    adj_or0: np.ndarray = field(default=None)
    adj_or1: np.ndarray = field(default=None)

    @property
    def n_nodes(self) -> int:
        return len(self.names)


def build_adjacency(n_nodes: int, edge_from: Sequence[int], edge_to: Sequence[int]):
    """Bidirected adjacency exactly as the reference builds it: for each L line, the forward
    entry {or1, id2, or2} at id1 unless already present (then the line is skipped entirely), and
    the reverse entry {!or2, id1, !or1} at id2 WITHOUT a presence check
    (gfalibs/src/gfa.cpp:1082-1093).  Returns CSR (offsets, target, or0, or1) in insertion order."""
    rows: List[List[Tuple[int, int, int]]] = [[] for _ in range(n_nodes)]
    for x, y in zip(edge_from, edge_to):
        x, y = int(x), int(y)
        fw = (x & 1, y >> 1, y & 1)
        if fw in rows[x >> 1]:
            continue
        rows[x >> 1].append(fw)
        rows[y >> 1].append(((y & 1) ^ 1, x >> 1, (x & 1) ^ 1))
    offsets = np.zeros(n_nodes + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(r) for r in rows])
    flat = [e for r in rows for e in r]
    target = np.array([e[1] for e in flat], dtype=np.uint32)
    or0 = np.array([e[0] for e in flat], dtype=np.uint8)
    or1 = np.array([e[2] for e in flat], dtype=np.uint8)
    return offsets, target, or0, or1


def node_table(tangle: Tangle):
    """NodeTable of dijkstra() (reference src/eval.cpp:125-128, include/nodetable.h:16-54):
    nodelist records with count >= 1, then add(source), add(destination); insert() keeps an
    existing record but nodeCount grows regardless.  -> (nodes, counts, node_count)"""
    rec: Dict[int, int] = {}
    node_count = 0
    for node, count in tangle.nodelist:
        if count < 1:
            continue
        node_count += count
        rec.setdefault(node, count)
    for node in (tangle.source, tangle.destination):
        rec.setdefault(node, 1)
        node_count += 1
    nodes = np.array(list(rec.keys()), dtype=np.uint32)
    counts = np.array(list(rec.values()), dtype=np.uint32)
    return nodes, counts, node_count


def make_tangle(n_nodes: int, extra_edges: int, seed: int, minus_frac: float = 0.05,
                count2_frac: float = 0.05, prefix: str = "utig4-") -> Tangle:
    """Backbone i+ -> (i+1)+ plus `extra_edges` random a -> b links (a != b), `minus_frac` of all
    links with a '-' endpoint; source = node 0, destination = node n-1; nodelist = interior nodes,
    count 1 (`count2_frac` of them count 2)."""
    rng = np.random.default_rng(seed)
    ef = [(i << 1) for i in range(n_nodes - 1)]
    et = [((i + 1) << 1) for i in range(n_nodes - 1)]
    a = rng.integers(0, n_nodes, size=extra_edges)
    b = rng.integers(0, n_nodes - 1, size=extra_edges)
    b = np.where(b >= a, b + 1, b)
    ef += [int(x) << 1 for x in a]
    et += [int(x) << 1 for x in b]
    ef = np.array(ef, dtype=np.uint32)
    et = np.array(et, dtype=np.uint32)
    flip = rng.random(ef.size) < minus_frac
    side = rng.random(ef.size) < 0.5
    ef = np.where(flip & side, ef | 1, ef).astype(np.uint32)
    et = np.where(flip & ~side, et | 1, et).astype(np.uint32)
    names = [f"{prefix}{i}" for i in range(n_nodes)]
    two = rng.random(n_nodes) < count2_frac
    nodelist = [(i, 2 if two[i] else 1) for i in range(1, n_nodes - 1)]
    t = Tangle(names, ef, et, 0, n_nodes - 1, nodelist)
    t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1 = build_adjacency(n_nodes, ef, et)
    return t


def oriented_successors(n_nodes: int, edge_from: np.ndarray, edge_to: np.ndarray):
    """Successor CSR over oriented steps (2*n_nodes rows): x -> y for every link and its twin."""
    fr = np.concatenate([edge_from, edge_to ^ 1]).astype(np.int64)
    to = np.concatenate([edge_to, edge_from ^ 1]).astype(np.int64)
    key = np.unique(fr * (2 * n_nodes) + to)
    fr, to = key // (2 * n_nodes), key % (2 * n_nodes)
    off = np.zeros(2 * n_nodes + 1, dtype=np.int64)
    np.add.at(off, fr + 1, 1)
    off = np.cumsum(off)
    return off, to.astype(np.uint32)


def random_walk_alignments(n_nodes: int, edge_from: np.ndarray, edge_to: np.ndarray, A: int,
                           lengths: Sequence[int], probs: Sequence[float], seed: int,
                           rc_frac: float = 0.5):
    """A random walks on the bidirected graph -> CSR (offsets u32[A+1], steps u32).  Walk length
    drawn from (lengths, probs); a walk that reaches a dead end is cut there; `rc_frac` of the
    walks are emitted reverse-complemented."""
    rng = np.random.default_rng(seed)
    soff, sto = oriented_successors(n_nodes, edge_from, edge_to)
    deg = (soff[1:] - soff[:-1])
    want = rng.choice(np.asarray(lengths), size=A, p=np.asarray(probs)).astype(np.int64)
    maxlen = int(want.max()) if A else 0
    walk = np.zeros((A, maxlen), dtype=np.uint32)
    cur = rng.integers(0, 2 * n_nodes, size=A).astype(np.int64)
    got = np.ones(A, dtype=np.int64)
    if maxlen:
        walk[:, 0] = cur
    alive = want > 1
    for j in range(1, maxlen):
        alive &= (deg[cur] > 0) & (want > j)
        idx = np.nonzero(alive)[0]
        if idx.size == 0:
            break
        c = cur[idx]
        k = (rng.random(idx.size) * deg[c]).astype(np.int64)
        nxt = sto[soff[c] + k].astype(np.int64)
        cur[idx] = nxt
        walk[idx, j] = nxt
        got[idx] = j + 1
    offsets = np.zeros(A + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum(got)
    assert offsets[-1] < 2 ** 32
    rc = rng.random(A) < rc_frac
    col = np.arange(maxlen, dtype=np.int64)[None, :]
    valid = col < got[:, None]
    # reverse-complement rows: step j <- walk[got-1-j] ^ 1
    src_col = np.where(rc[:, None], got[:, None] - 1 - col, col)
    src_col = np.clip(src_col, 0, max(maxlen - 1, 0))
    out = np.take_along_axis(walk, src_col, axis=1) ^ rc[:, None].astype(np.uint32)
    steps = out[valid].astype(np.uint32)
    return offsets.astype(np.uint32), steps


HIFI_LENGTHS = ([3, 4, 5, 6, 8], [0.35, 0.30, 0.20, 0.10, 0.05])       # cfg3 / cfg4
ONT_LENGTHS = (list(range(4, 17)), [1.0 / 13] * 13)                    # cfg5


@dataclass
class Workload:
    name: str
    tangle: Tangle
    offsets: np.ndarray
    steps: np.ndarray


def make_workload(name: str, scale: float = 1.0, A: Optional[int] = None) -> Workload:
    """cfg3: 200-node HiFi tangle, 1M alignments; cfg5: 400-node ONT tangle, 50M alignments;
    cfg4: 500k-node graph, 20M alignments.  `scale` shrinks A (tests)."""
    if name == "cfg3":
        t = make_tangle(200, 200, seed=1, minus_frac=0.05, count2_frac=0.05)
        n_aln, (ls, ps), seed = 1_000_000, HIFI_LENGTHS, 11
    elif name == "cfg5":
        t = make_tangle(400, 800, seed=3, minus_frac=0.05, count2_frac=0.10)
        n_aln, (ls, ps), seed = 50_000_000, ONT_LENGTHS, 13
    elif name == "cfg4":
        t = make_tangle(500_000, 200_000, seed=2, minus_frac=0.05, count2_frac=0.0)
        n_aln, (ls, ps), seed = 20_000_000, HIFI_LENGTHS, 12
    else:
        raise ValueError(name)
    n_aln = int(A if A is not None else max(1, round(n_aln * scale)))
    off, st = random_walk_alignments(t.n_nodes, t.edge_from, t.edge_to, n_aln, ls, ps, seed)
    return Workload(name, t, off, st)


def candidate_paths(t: Tangle, P: int, seed: int, max_len: Optional[int] = None):
    """P candidate paths of the kind the search expands: walks from the source that follow the
    adjacency with orientation continuity and the nodelist visit budgets, cut at a uniformly
    drawn length.  -> CSR (offsets, steps)."""
    rng = np.random.default_rng(seed)
    nodes, counts, _ = node_table(t)
    budget0 = dict(zip(nodes.tolist(), counts.tolist()))
    if max_len is None:
        max_len = min(t.n_nodes, 1000)
    offs, steps = [0], []
    while len(offs) - 1 < P:
        want = int(rng.integers(2, max_len + 1))
        budget = dict(budget0)
        path = [t.source << 1]
        first = True
        while len(path) < want:
            node = path[-1] >> 1
            lo, hi = int(t.adj_offsets[node]), int(t.adj_offsets[node + 1])
            cand = [k for k in range(lo, hi)
                    if (first or int(t.adj_or0[k]) == (path[-1] & 1))
                    and budget.get(int(t.adj_target[k]), 0) > 0]
            if not cand:
                break
            k = cand[int(rng.integers(0, len(cand)))]
            if first:
                path[0] = (path[0] & ~1) | int(t.adj_or0[k])
                first = False
            v = int(t.adj_target[k])
            path.append((v << 1) | int(t.adj_or1[k]))
            if v == t.destination:
                break
            budget[v] -= 1
        if len(path) >= 2:
            steps += path
            offs.append(len(steps))
    return np.array(offs, dtype=np.uint32), np.array(steps, dtype=np.uint32)


# ---- text formats (small scale: CLIs, reference binary) ----------------------------------------

def step_str(names: Sequence[str], s: int, gaf: bool = False) -> str:
    if gaf:
        return (">" if not (s & 1) else "<") + names[s >> 1]
    return names[s >> 1] + ("-" if s & 1 else "+")


def write_gfa(path: str, t: Tangle, seq: str = "ACGT"):
    with open(path, "w") as f:
        f.write("H\tVN:Z:1.2\n")
        for nm in t.names:
            f.write(f"S\t{nm}\t{seq}\tLN:i:{len(seq)}\n")
        for x, y in zip(t.edge_from.tolist(), t.edge_to.tolist()):
            f.write(f"L\t{t.names[x >> 1]}\t{'-' if x & 1 else '+'}\t{t.names[y >> 1]}\t"
                    f"{'-' if y & 1 else '+'}\t0M\n")


def write_gaf(path: str, names: Sequence[str], offsets: np.ndarray, steps: np.ndarray,
              read_prefix: str = "read"):
    off = offsets.tolist()
    st = steps.tolist()
    with open(path, "w") as f:
        for a in range(len(off) - 1):
            walk = "".join(step_str(names, s, gaf=True) for s in st[off[a]:off[a + 1]])
            m = off[a + 1] - off[a]
            ln = 4 * max(m, 1)
            f.write(f"{read_prefix}{a}\t{ln}\t0\t{ln}\t+\t{walk}\t{ln}\t0\t{ln}\t{ln}\t{ln}\t60\n")


def write_nodelist(path: str, t: Tangle):
    with open(path, "w") as f:
        for node, count in t.nodelist:
            f.write(f"{t.names[node]}\t{count}\n")


# ---- full-size tables generated on the device (bench.py: cfg4 / cfg5) ---------------------------

def random_walk_alignments_device(n_nodes: int, edge_from: np.ndarray, edge_to: np.ndarray, A: int,
                                  lengths: Sequence[int], probs: Sequence[float], seed: int, device,
                                  rc_frac: float = 0.5, chunk: int = 4_000_000):
    """The same walks as random_walk_alignments (same rules: uniform oriented start, uniform
    successor, cut at dead ends, `rc_frac` emitted reverse-complemented), drawn with torch on
    `device` in chunks of rows -> (offsets int32[A+1], steps int32[S]) device tensors (bit patterns
    of the uint32 CSR).  Deterministic for a given (seed, A, chunk) on one GPU model; NOT the same
    stream of random numbers as the numpy generator."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    soff_np, sto_np = oriented_successors(n_nodes, edge_from, edge_to)
    soff = torch.from_numpy(soff_np.astype(np.int64)).to(device)
    sto = torch.from_numpy(sto_np.astype(np.int64)).to(device)
    if sto.numel() == 0:
        sto = torch.zeros(1, dtype=torch.int64, device=device)
    deg = soff[1:] - soff[:-1]
    lens = torch.tensor(list(lengths), dtype=torch.int64, device=device)
    pr = torch.tensor(list(probs), dtype=torch.float32, device=device)
    maxlen = int(max(lengths)) if A else 0
    col = torch.arange(maxlen, device=device)[None, :]
    got_all, step_chunks = [], []
    for lo in range(0, A, chunk):
        n = min(chunk, A - lo)
        want = lens[torch.multinomial(pr, n, replacement=True, generator=g)]
        cur = torch.randint(0, 2 * n_nodes, (n,), generator=g, device=device)
        walk = torch.zeros((n, maxlen), dtype=torch.int64, device=device)
        walk[:, 0] = cur
        got = torch.ones(n, dtype=torch.int64, device=device)
        alive = want > 1
        for j in range(1, maxlen):
            alive = alive & (deg[cur] > 0) & (want > j)
            k = (torch.rand(n, generator=g, device=device) * deg[cur]).to(torch.int64)
            k = torch.minimum(k, torch.clamp(deg[cur] - 1, min=0))
            nxt = sto[torch.clamp(soff[cur] + k, max=sto.numel() - 1)]
            cur = torch.where(alive, nxt, cur)
            walk[:, j] = torch.where(alive, nxt, torch.zeros_like(nxt))
            got += alive.to(torch.int64)
        rc = torch.rand(n, generator=g, device=device) < rc_frac
        src_col = torch.where(rc[:, None], got[:, None] - 1 - col, col).clamp_(0, max(maxlen - 1, 0))
        out = torch.gather(walk, 1, src_col) ^ rc[:, None].to(torch.int64)
        step_chunks.append(out[col < got[:, None]].to(torch.int32))
        got_all.append(got)
        del walk, out, src_col
    got = torch.cat(got_all) if got_all else torch.zeros(0, dtype=torch.int64, device=device)
    offsets = torch.zeros(A + 1, dtype=torch.int64, device=device)
    offsets[1:] = torch.cumsum(got, 0)
    assert int(offsets[-1]) < 2 ** 31
    steps = torch.cat(step_chunks) if step_chunks else torch.zeros(0, dtype=torch.int32, device=device)
    return offsets.to(torch.int32), steps


def table_shard_device(n_nodes: int, edge_from: np.ndarray, edge_to: np.ndarray, A_total: int,
                       lengths: Sequence[int], probs: Sequence[float], seed: int, device, rank: int = 0,
                       world: int = 1, pieces: int = 8):
    """Rank `rank`'s shard of ONE table that is the same for every world size dividing `pieces`: the table is
    drawn as `pieces` pieces of A_total / pieces rows (seed + piece) and the rank holds the pieces
    [rank * pieces / world, (rank + 1) * pieces / world), concatenated.  Summed counts and search rows over
    the ranks are therefore identical at N = 1, 2, 4, 8.  -> (offsets int32[A+1], steps int32[S]) on device."""
    import torch
    assert pieces % world == 0 and A_total % pieces == 0
    per = pieces // world
    offs, sts, base = [], [], 0
    for pc in range(rank * per, (rank + 1) * per):
        o, s_ = random_walk_alignments_device(n_nodes, edge_from, edge_to, A_total // pieces, lengths, probs,
                                              seed=seed + pc, device=device)
        offs.append(o[:-1] + base)
        sts.append(s_)
        base += int(o[-1].item())
    assert base < 2 ** 31
    offs.append(torch.tensor([base], dtype=torch.int32, device=device))
    return torch.cat(offs), torch.cat(sts)
