"""Minimal GFA1 / GAF / nodelist text -> packed arrays, for tests and small tools (the drop-in CLI
has its own C++ packer in csrc/host/).  Follows the reference's reading of these files:

  * segments get dense ids in order of first appearance (S lines, then names first seen on L
    lines: gfalibs/src/input-gfa.cpp:497-641); only S and L lines matter to the hot path
  * a GAF walk `>a<b...` becomes oriented steps ('>' -> '+', anything else -> '-',
    reference src/alignments.cpp:75-94); a name the graph does not know aliases to id 0, as the
    reference's headersToIds[...] default-inserts 0 (src/alignments.cpp:86)
  * nodelist: `name[\\tcount]`, count < 1 drops the line (include/nodetable.h:16-42)
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np

from gfalign_b200 import synth


def parse_gfa(text: str):
    """-> (names, edge_from, edge_to) with L lines in file order as oriented steps."""
    ids: Dict[str, int] = {}
    names: List[str] = []

    def nid(nm: str) -> int:
        if nm not in ids:
            ids[nm] = len(names)
            names.append(nm)
        return ids[nm]

    lines = text.splitlines()
    for ln in lines:
        if ln.startswith("S\t"):
            nid(ln.split("\t")[1])
    ef, et = [], []
    for ln in lines:
        if ln.startswith("L\t"):
            c = ln.split("\t")
            ef.append(nid(c[1]) << 1 | (c[2][0] == "-"))
            et.append(nid(c[3]) << 1 | (c[4][0] == "-"))
    return names, np.array(ef, dtype=np.uint32), np.array(et, dtype=np.uint32)


def parse_walk(walk: str, ids: Dict[str, int]) -> List[int]:
    steps, i = [], 0
    while i < len(walk):
        j = i + 1
        while j < len(walk) and walk[j] not in "<>":
            j += 1
        steps.append(ids.get(walk[i + 1:j], 0) << 1 | (walk[i] != ">"))
        i = j
    return steps


def parse_gaf(text: str, names: List[str]):
    """-> (qnames, offsets u32[A+1], steps u32) in file order."""
    ids = {nm: i for i, nm in enumerate(names)}
    qn, off, st = [], [0], []
    for ln in text.splitlines():
        if not ln:
            continue
        c = ln.split("\t")
        qn.append(c[0])
        st.extend(parse_walk(c[5], ids))
        off.append(len(st))
    return qn, np.array(off, dtype=np.uint32), np.array(st, dtype=np.uint32)


def parse_nodelist(text: str, names: List[str]) -> List[Tuple[int, int]]:
    ids = {nm: i for i, nm in enumerate(names)}
    out = []
    for ln in text.splitlines():
        c = ln.split("\t")
        count = int(c[1]) if len(c) > 1 else 1
        if count < 1:
            continue
        out.append((ids[c[0]], count))
    return out


def parse_path(path: str, names: List[str]) -> np.ndarray:
    """`a+,b-;c+` (reference src/eval.cpp:203-226) -> oriented steps."""
    ids = {nm: i for i, nm in enumerate(names)}
    comps = path.replace(";", ",").split(",")
    return np.array([ids[c[:-1]] << 1 | (c[-1] == "-") for c in comps], dtype=np.uint32)


def tangle_from_text(gfa: str, nodelist: str, source: str, destination: str) -> synth.Tangle:
    names, ef, et = parse_gfa(gfa)
    ids = {nm: i for i, nm in enumerate(names)}
    t = synth.Tangle(names, ef, et, ids[source], ids[destination], parse_nodelist(nodelist, names))
    t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1 = synth.build_adjacency(len(names), ef, et)
    return t
