"""ctypes binding of the CPU oracle (oracle/liboracle.so) and of the compiled reference
(oracle/_ref/).  TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs, never from gfalign_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_u32p = C.POINTER(C.c_uint32)
_i32p = C.POINTER(C.c_int32)
FILTERED = -(2 ** 31)


def build(ref: bool = True) -> None:
    """Compile liboracle.so (always) and, when /root/reference is present, oracle/_ref/."""
    subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)
    if ref and os.path.exists("/root/reference/src/eval.cpp"):
        subprocess.run(["make", "-s", "-j8", "-C", HERE, "ref"], check=True)


def ref_binary() -> Optional[str]:
    p = os.path.join(HERE, "_ref", "gfalign")
    return p if os.path.exists(p) else None


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


def _p(a, t=_u32p):
    return a.ctypes.data_as(t)


class Oracle:
    def __init__(self):
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build(ref=False)
        self.lib = C.CDLL(path)
        L = self.lib
        L.orc_pair_score.restype = C.c_int32
        L.orc_pair_score.argtypes = [_u32p, C.c_uint32, _u32p, C.c_uint32]
        L.orc_pair_alignment.restype = C.c_uint32
        L.orc_pair_alignment.argtypes = [_u32p, C.c_uint32, _u32p, C.c_uint32, _u32p, _u32p, _i32p]
        L.orc_evaluate_path.restype = None
        L.orc_evaluate_path.argtypes = [_u32p, C.c_uint32, _u32p, _u32p, C.c_uint64, C.c_int,
                                        _u32p, _u32p, _u32p, _i32p, _i32p]
        L.orc_evaluate_paths.restype = None
        L.orc_evaluate_paths.argtypes = [_u32p, _u32p, C.c_uint32, _u32p, _u32p, C.c_uint64,
                                         C.c_int, C.c_int, _u32p]
        L.orc_edge_support.restype = None
        L.orc_edge_support.argtypes = [_u32p, _u32p, C.c_uint64, C.c_uint32, _u32p, _u32p,
                                       C.c_uint32, _u32p]

    def pair_score(self, A, B) -> int:
        A, B = _u32(A), _u32(B)
        return int(self.lib.orc_pair_score(_p(A), A.size, _p(B), B.size))

    def pair_alignment(self, A, B):
        """-> (SA, SB, score): gapped sequences in path order, 0xFFFFFFFF = gap."""
        A, B = _u32(A), _u32(B)
        SA = np.zeros(A.size + B.size + 1, dtype=np.uint32)
        SB = np.zeros(A.size + B.size + 1, dtype=np.uint32)
        sc = C.c_int32()
        n = self.lib.orc_pair_alignment(_p(A), A.size, _p(B), B.size, _p(SA), _p(SB), C.byref(sc))
        return SA[:n], SB[:n], sc.value

    def evaluate_path(self, path, offsets, steps, filter: bool, scores: bool = False):
        path, offsets, steps = _u32(path), _u32(offsets), _u32(steps)
        A = offsets.size - 1
        bad, good, un = C.c_uint32(), C.c_uint32(), C.c_uint32()
        tf = np.zeros(A, dtype=np.int32) if scores else None
        tr = np.zeros(A, dtype=np.int32) if scores else None
        self.lib.orc_evaluate_path(_p(path), path.size, _p(offsets), _p(steps), A, int(filter),
                                   C.byref(bad), C.byref(good), C.byref(un),
                                   _p(tf, _i32p) if scores else None,
                                   _p(tr, _i32p) if scores else None)
        return bad.value, good.value, tf, tr

    def evaluate_paths(self, path_offsets, path_steps, offsets, steps, filter: bool = True,
                       threads: int = 1) -> np.ndarray:
        po, ps, off, st = _u32(path_offsets), _u32(path_steps), _u32(offsets), _u32(steps)
        P = po.size - 1
        counts = np.zeros((P, 2), dtype=np.uint32)
        self.lib.orc_evaluate_paths(_p(po), _p(ps), P, _p(off), _p(st), off.size - 1, int(filter),
                                    threads, _p(counts))
        return counts

    def edge_support(self, offsets, steps, n_nodes: int, edge_from, edge_to) -> np.ndarray:
        off, st, ef, et = _u32(offsets), _u32(steps), _u32(edge_from), _u32(edge_to)
        w = np.zeros(ef.size, dtype=np.uint32)
        self.lib.orc_edge_support(_p(off), _p(st), off.size - 1, n_nodes, _p(ef), _p(et), ef.size,
                                  _p(w))
        return w


class RefShim:
    """The reference's own alignPaths() arithmetic (oracle/_ref/libgfalign_ref.so)."""

    def __init__(self):
        path = os.path.join(HERE, "_ref", "libgfalign_ref.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.lib.ref_pair_scores.argtypes = [_u32p, C.c_uint32, _u32p, C.c_uint32, _i32p, _i32p]
        self.lib.ref_path_scores.argtypes = [_u32p, C.c_uint32, _u32p, _u32p, C.c_uint64, _i32p,
                                             _i32p]

    def pair_scores(self, A, B):
        A, B = _u32(A), _u32(B)
        tf, tr = C.c_int32(), C.c_int32()
        rc = self.lib.ref_pair_scores(_p(A), A.size, _p(B), B.size, C.byref(tf), C.byref(tr))
        assert rc == 0
        return tf.value, tr.value

    def path_scores(self, path, offsets, steps):
        path, off, st = _u32(path), _u32(offsets), _u32(steps)
        A = off.size - 1
        tf = np.zeros(A, dtype=np.int32)
        tr = np.zeros(A, dtype=np.int32)
        rc = self.lib.ref_path_scores(_p(path), path.size, _p(off), _p(st), A, _p(tf, _i32p),
                                      _p(tr, _i32p))
        assert rc == 0
        return tf, tr
