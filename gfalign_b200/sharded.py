"""Alignment shards over the GPUs of one box (one process per GPU, torch.distributed).

The hot path shards by alignment: every rank keeps a contiguous range of the CSR in its own HBM,
scores the same batch of candidate paths against it, and the small P x {bad, good} record is
summed with ONE all-reduce (NCCL over NVLink on GPUs; gloo in the CPU tests).  Integer addition
commutes, so the result is bit-identical to the single-GPU one.  The search itself is not
sharded: every rank replays the same canonical heap (reference src/eval.cpp:134-189) on the same
summed counts, so all ranks stay in lock step and emit the same rows.

evalGFA shards the same way: per-rank traversal counters (2 per L line), one all-reduce, then the
reference's 2k-1 hairpin rule on the summed counters.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import numpy as np

from . import capi


def shard_bounds(offsets: np.ndarray, world: int) -> np.ndarray:
    """Row boundaries (world+1) of contiguous shards balanced on streamed bytes (m + 1 words per
    row).  Deterministic; empty shards are legal."""
    offsets = np.asarray(offsets, dtype=np.int64)
    A = offsets.size - 1
    cost = offsets[1:] + np.arange(1, A + 1, dtype=np.int64)      # prefix of (m + 1)
    total = int(cost[-1]) if A else 0
    cuts = [0]
    for r in range(1, world):
        cuts.append(int(np.searchsorted(cost, total * r / world, side="left")) if A else 0)
    cuts.append(A)
    return np.maximum.accumulate(np.array(cuts, dtype=np.int64))


def take_shard(offsets: np.ndarray, steps: np.ndarray, lo: int, hi: int) -> Tuple[np.ndarray, np.ndarray]:
    o = np.asarray(offsets[lo:hi + 1], dtype=np.int64)
    return (o - o[0]).astype(np.uint32), np.ascontiguousarray(steps[int(o[0]):int(o[-1])], dtype=np.uint32)


class PeerExchange:
    """The exchange step over NVLink peer memory: a symmetric buffer (torch symmetric memory, one per
    rank, mapped into every process of the box) holds two halves of n_cap count slots plus the
    epoch flags; `half_ptr()` is where this rank's kernels accumulate the next batch's records and
    `sum_into(out, n, stream)` launches gfe_peer_sum_u32 (csrc/peer.cu) behind them.  Every rank
    must call both in the same order (they do: the search is replayed in lock step)."""

    def __init__(self, n_cap: int, device, group=None):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        self.torch = torch
        self.n_cap = int(n_cap)
        group = group if group is not None else dist.group.WORLD
        words = 2 * self.n_cap + capi.PEER_MAX
        self.buf = symm_mem.empty(words, dtype=torch.int32, device=device)
        self.buf.zero_()
        self.hdl = symm_mem.rendezvous(self.buf, group)
        self.rank, self.world = int(self.hdl.rank), int(self.hdl.world_size)
        if self.world > capi.PEER_MAX:
            raise ValueError(f"PeerExchange: at most {capi.PEER_MAX} ranks")
        self.bases = [int(p) for p in self.hdl.buffer_ptrs]
        torch.cuda.synchronize(device)
        dist.barrier(group=group)          # every rank's flags are zero before the first epoch
        self.epoch = 0

    def c_exchange(self) -> "capi.Exchange":
        """The same buffer as a gfe_exchange for gfe_score_paths_sharded / gfe_search_sharded.  Use either
        this handle or half_ptr()/sum_into() on one PeerExchange, not both (each keeps its own epoch)."""
        if getattr(self, "_cx", None) is None:
            if self.epoch:
                raise RuntimeError("PeerExchange already used through sum_into()")
            self._cx = capi.Exchange(self.bases, self.rank, self.n_cap)
        return self._cx

    def half_ptr(self) -> int:
        """Device address of the half the NEXT sum_into() reads on this rank."""
        return self.bases[self.rank] + 4 * self.n_cap * ((self.epoch + 1) & 1)

    def half_tensor(self, n: int):
        o = self.n_cap * ((self.epoch + 1) & 1)
        return self.buf[o:o + n]

    def sum_into(self, out_ptr: int, n: int, stream: int = 0) -> None:
        self.epoch += 1
        capi.peer_sum_u32(self.bases, self.rank, self.n_cap, n, self.epoch, out_ptr, stream)


class ShardedScorer:
    """scorer(path_offsets, path_steps) -> summed (P, 2) counts over all ranks of `group`.

    `local` scores the batch against this rank's shard: Engine.score_paths on a GPU rank.  With
    `device` set (a torch cuda device) the counts travel as a device tensor (NCCL)."""

    def __init__(self, local: Callable[[np.ndarray, np.ndarray], np.ndarray], group=None, device=None):
        import torch.distributed as dist
        self.local, self.group, self.device, self.dist = local, group, device, dist
        self.calls = 0

    def __call__(self, path_offsets: np.ndarray, path_steps: np.ndarray) -> np.ndarray:
        import torch
        counts = np.ascontiguousarray(self.local(path_offsets, path_steps), dtype=np.uint32)
        t = torch.from_numpy(counts.astype(np.int64))
        if self.device is not None:
            t = t.to(self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        self.calls += 1
        return t.cpu().numpy().astype(np.uint32)


class PeerShardedScorer:
    """The same contract as ShardedScorer for GPU ranks of one box, through ONE C call per batch
    (gfe_score_paths_sharded): the batch goes host -> device once, the kernels accumulate this rank's
    records into the exchange buffer, the peer-memory kernel sums all shards, and only the summed
    P x 2 record comes back (no NCCL call, no extra device<->host hop).  A rank's worklist overflow
    travels as one more slot of the record, so all ranks repeat that batch together.  Batches above
    `(n_cap - 1) / 2` paths go to `fallback` (a ShardedScorer); P is the same on every rank, so the
    decision is too."""

    def __init__(self, engine, exchange: PeerExchange, device=None, filter: bool = True, fallback=None):
        self.engine, self.px, self.cx = engine, exchange, exchange.c_exchange()
        self.filter, self.fallback = filter, fallback
        self.calls = 0

    def __call__(self, path_offsets: np.ndarray, path_steps: np.ndarray) -> np.ndarray:
        P = int(path_offsets.size) - 1
        if 2 * P + 1 > self.px.n_cap:
            if self.fallback is None:
                raise ValueError(f"batch of {P} paths exceeds the exchange buffer")
            return self.fallback(path_offsets, path_steps)
        self.calls += 1
        return self.engine.score_paths_sharded(self.cx, path_offsets, path_steps, self.filter)


def sharded_search(scorer: ShardedScorer, tangle, nodes, counts, node_count, **kw):
    """`gfalign search` over sharded alignments: identical rows on every rank."""
    return capi.search_with_scorer(scorer, tangle.adj_offsets, tangle.adj_target, tangle.adj_or0,
                                   tangle.adj_or1, tangle.source, tangle.destination, nodes, counts,
                                   node_count, **kw)


def sharded_edge_support(local_traversals: Callable[[], np.ndarray], edge_from, edge_to,
                         group=None, device=None) -> np.ndarray:
    """`local_traversals()` -> this rank's 2*E raw traversal counters (slot 2i = events equal to L
    line i, 2i+1 = its reverse complement; gfe_edge_traversals_device).  Returns RC:i weights."""
    import torch
    import torch.distributed as dist
    c = np.ascontiguousarray(local_traversals(), dtype=np.uint32)
    t = torch.from_numpy(c.astype(np.int64))
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return capi.Engine.edge_weights_from_counts(edge_from, edge_to, t.cpu().numpy().astype(np.uint32))
