"""Shared by the CPU and GPU suites: render `gfalign evalPath` stdout (reference src/eval.cpp:71-72,
100-105, 241 and PairwisePathAlignment::getAlignment, include/alignments.h:111-120) from per-pair
gapped alignments, whoever computed them."""
from typing import Callable, List, Sequence

import numpy as np

GAP = 0xFFFFFFFF


def ops_from_gapped(SA, SB) -> List[int]:
    """column op codes of include/gfalign_b200.h (gfe_trace_pairs) from two gapped sequences"""
    out = []
    for a, b in zip(SA.tolist(), SB.tolist()):
        out.append(2 if b == GAP else 3 if a == GAP else 0 if a == b else 1)
    return out


def render_line(names: Sequence[str], path, B, ops) -> str:
    i = j = 0
    s = []
    for op in ops:
        if op == 2:
            s.append("-" * (len(names[int(path[i]) >> 1]) + 1))
        else:
            b = int(B[j]); j += 1
            s.append("." * (len(names[b >> 1]) + 1) if op == 0
                     else names[b >> 1] + ("-" if b & 1 else "+"))
        if op != 3:
            i += 1
    return "".join(x + "," for x in s)


def evalpath_stdout(names, path, qnames, offsets, steps, tfw, trc,
                    trace: Callable[[np.ndarray, np.ndarray], List[int]]) -> str:
    """`trace(path, B)` returns the op codes of alignPaths(path, B)."""
    path = np.asarray(path, dtype=np.uint32)
    lines = [",".join(names[int(s) >> 1] + ("-" if s & 1 else "+") for s in path.tolist())]
    bad = good = 0
    for a, qn in enumerate(qnames):
        B = np.asarray(steps[offsets[a]:offsets[a + 1]], dtype=np.uint32)
        if not tfw[a] > trc[a]:
            B = (B[::-1] ^ 1).astype(np.uint32)
        best = max(int(tfw[a]), int(trc[a]))
        bad += best < 0
        good += best >= 0
        lines.append(f"{render_line(names, path, B, trace(path, B))}\t{qn}\t{best}")
    uniq = len(set((path >> 1).tolist()))
    lines.append(f"{bad}\t{good}\t{bad - good - uniq}\t{path.size}\t{uniq}")
    return "\n".join(lines) + "\n"
