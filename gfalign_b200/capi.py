"""ctypes binding of include/gfalign_b200.h (the drop-in C ABI).

Mirrors the three reference entry points the ABI replaces — evaluatePath / evalPath scoring /
InAlignments::buildEdgeGraph and the dijkstra() driver (reference src/eval.cpp:34-242) — with the
same argument meaning; errors raise GfeError carrying the library's message.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import numpy as np

_u32p = C.POINTER(C.c_uint32)
_i32p = C.POINTER(C.c_int32)
_u8p = C.POINTER(C.c_uint8)

GFE_SCORE_FILTERED = -(2 ** 31)
GFE_MAX_STEPS = 1000


class GfeError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"gfalign_b200 error {status}: {message}")
        self.status = status


def lib_path() -> str:
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "libgfalign_b200.so")


class _SearchRow(C.Structure):
    _fields_ = [("path_counter", C.c_uint64), ("bad", C.c_uint32), ("good", C.c_uint32),
                ("alt", C.c_int32), ("size", C.c_uint32), ("uniques", C.c_uint32),
                ("hamiltonian", C.c_int), ("steps", _u32p)]


class _Graph(C.Structure):
    _fields_ = [("n_nodes", C.c_uint32), ("adj_offsets", _u32p), ("adj_target", _u32p),
                ("adj_or0", _u8p), ("adj_or1", _u8p)]


class _TextView(C.Structure):
    _fields_ = [("n_nodes", C.c_uint32), ("n_links", C.c_uint32), ("edge_from", _u32p),
                ("edge_to", _u32p), ("graph", _Graph), ("A", C.c_uint64), ("offsets", _u32p),
                ("steps", _u32p)]


class _SearchParams(C.Structure):
    _fields_ = [("source", C.c_uint32), ("destination", C.c_uint32), ("budget_nodes", _u32p),
                ("budget_counts", _u32p), ("n_budget", C.c_uint32), ("node_count", C.c_uint32),
                ("max_steps", C.c_uint32), ("min_nodes", C.c_uint32),
                ("return_all_paths", C.c_int), ("speculate", C.c_uint32)]


class _SearchStats(C.Structure):
    _fields_ = [("extractions", C.c_uint64), ("evaluations", C.c_uint64),
                ("kernel_batches", C.c_uint64), ("paths_scored", C.c_uint64), ("rows", C.c_uint64),
                ("reached_max_steps", C.c_int), ("seconds_scoring", C.c_double),
                ("seconds_total", C.c_double)]


class _PathBatch(C.Structure):
    _fields_ = [("n_bases", C.c_uint32), ("base_offsets", _u32p), ("base_steps", _u32p), ("P", C.c_uint32),
                ("item_base", _u32p), ("suffix_offsets", _u32p), ("suffix_steps", _u32p), ("mode", _u8p)]


class _Counters(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("exact_pairs", C.c_uint64),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("slow_pairs", C.c_uint64)]


_ROW_FN = C.CFUNCTYPE(None, C.POINTER(_SearchRow), C.c_void_p)
_SCORE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, _u32p, _u32p, C.c_uint32, _u32p)

# every symbol include/gfalign_b200.h declares (checked by tests/test_abi.py)
EXPORTS = [
    "gfe_create", "gfe_destroy", "gfe_last_error", "gfe_version", "gfe_set_alignments",
    "gfe_set_alignments_device", "gfe_set_universe", "gfe_score_paths", "gfe_score_paths_device",
    "gfe_score_pairs", "gfe_edge_support", "gfe_edge_traversals_device",
    "gfe_edge_weights_from_counts", "gfe_search", "gfe_search_with_scorer", "gfe_get_counters", "gfe_set_timing",
    "gfe_last_kernel_ms", "gfe_kernel_time_stats", "gfe_device_status",
    "gfe_trace_pairs", "gfe_filter_alignments",
    "gfe_text_open", "gfe_text_close", "gfe_text_get", "gfe_text_node_name", "gfe_text_node_id",
    "gfe_text_qname", "gfe_peer_sum_u32", "gfe_device_status_to",
    "gfe_exchange_create", "gfe_exchange_destroy", "gfe_score_paths_sharded", "gfe_search_sharded",
    "gfe_score_paths_incremental", "gfe_score_paths_incremental_sharded", "gfe_exchange_set_batch_sharding",
    "gfe_last_aux_kernel_ms", "gfe_set_dedup", "gfe_get_dedup_stats",
    "gfe_filter_alignments_static",
]

PEER_MAX = 16      # GFE_PEER_MAX

_lib = None


def peer_sum_u32(peer_bases, rank: int, n_cap: int, n: int, epoch: int, d_out: int, stream: int = 0) -> None:
    """gfe_peer_sum_u32: sum the n count slots of all ranks out of their peer-mapped exchange
    buffers (peer_bases[q] = rank q's buffer as mapped here) into the device array at d_out."""
    lib = load_library()
    arr = (C.c_uint64 * len(peer_bases))(*[int(b) for b in peer_bases])
    rc = lib.gfe_peer_sum_u32(arr, len(peer_bases), rank, n_cap, n, epoch, C.c_void_p(d_out), C.c_void_p(stream))
    if rc != 0:
        raise GfeError(rc, "gfe_peer_sum_u32: bad arguments or launch failure")


class Exchange:
    """gfe_exchange: this rank's view of the box's peer-mapped exchange buffer (see sharded.PeerExchange
    for the allocation through torch symmetric memory)."""

    def __init__(self, peer_bases, rank: int, n_cap: int):
        lib = load_library()
        arr = (C.c_uint64 * len(peer_bases))(*[int(b) for b in peer_bases])
        h = C.c_void_p()
        rc = lib.gfe_exchange_create(arr, len(peer_bases), rank, n_cap, C.byref(h))
        if rc != 0:
            raise GfeError(rc, "gfe_exchange_create: bad arguments")
        self.handle, self._lib, self.n_cap = h, lib, n_cap

    def set_batch_sharding(self, on: bool = True):
        """Replicated tables: deal the candidates of a batch over the ranks instead of the alignments."""
        self._lib.gfe_exchange_set_batch_sharding(self.handle, int(on))

    def close(self):
        if getattr(self, "handle", None):
            self._lib.gfe_exchange_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def load_library() -> C.CDLL:
    """Load libgfalign_b200.so (built in-tree by __graft_entry__.build()). Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise ImportError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; "
                          f"g.build()'` (nvcc, sm_100a). gfalign_b200 has no CPU fallback.")
    lib = C.CDLL(path)
    lib.gfe_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    lib.gfe_destroy.argtypes = [C.c_void_p]
    lib.gfe_destroy.restype = None
    lib.gfe_last_error.argtypes = [C.c_void_p]
    lib.gfe_last_error.restype = C.c_char_p
    lib.gfe_version.restype = C.c_char_p
    lib.gfe_set_alignments.argtypes = [C.c_void_p, _u32p, _u32p, C.c_uint64, C.c_uint32]
    lib.gfe_set_alignments_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64,
                                              C.c_uint64, C.c_uint32]
    lib.gfe_set_universe.argtypes = [C.c_void_p, _u32p, C.c_uint32]
    lib.gfe_score_paths.argtypes = [C.c_void_p, _u32p, _u32p, C.c_uint32, C.c_int, _u32p]
    lib.gfe_score_paths_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                                           C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
    lib.gfe_score_pairs.argtypes = [C.c_void_p, _u32p, C.c_uint32, C.c_int, _i32p, _i32p, _u32p,
                                    _u32p]
    lib.gfe_edge_support.argtypes = [C.c_void_p, _u32p, _u32p, C.c_uint32, _u32p]
    lib.gfe_edge_traversals_device.argtypes = [C.c_void_p, _u32p, _u32p, C.c_uint32, C.c_void_p,
                                               C.c_void_p]
    lib.gfe_edge_weights_from_counts.argtypes = [_u32p, _u32p, C.c_uint32, _u32p, _u32p]
    lib.gfe_device_status_to.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.gfe_peer_sum_u32.argtypes = [C.POINTER(C.c_uint64), C.c_int, C.c_int, C.c_uint32, C.c_uint32,
                                     C.c_uint32, C.c_void_p, C.c_void_p]
    lib.gfe_score_paths_incremental.argtypes = [C.c_void_p, C.POINTER(_PathBatch), _i32p]
    lib.gfe_score_paths_incremental_sharded.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(_PathBatch), _i32p]
    lib.gfe_exchange_create.argtypes = [C.POINTER(C.c_uint64), C.c_int, C.c_int, C.c_uint32,
                                        C.POINTER(C.c_void_p)]
    lib.gfe_exchange_set_batch_sharding.argtypes = [C.c_void_p, C.c_int]
    lib.gfe_exchange_destroy.argtypes = [C.c_void_p]
    lib.gfe_exchange_destroy.restype = None
    lib.gfe_score_paths_sharded.argtypes = [C.c_void_p, C.c_void_p, _u32p, _u32p, C.c_uint32, C.c_int, _u32p]
    lib.gfe_search_sharded.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(_Graph), C.POINTER(_SearchParams),
                                       _ROW_FN, C.c_void_p, C.POINTER(_SearchStats)]
    lib.gfe_search.argtypes = [C.c_void_p, C.POINTER(_Graph), C.POINTER(_SearchParams), _ROW_FN,
                               C.c_void_p, C.POINTER(_SearchStats)]
    lib.gfe_search_with_scorer.argtypes = [_SCORE_FN, C.c_void_p, C.POINTER(_Graph),
                                           C.POINTER(_SearchParams), _ROW_FN, C.c_void_p,
                                           C.POINTER(_SearchStats)]
    lib.gfe_get_counters.argtypes = [C.c_void_p, C.POINTER(_Counters)]
    lib.gfe_set_timing.argtypes = [C.c_void_p, C.c_int]
    lib.gfe_set_dedup.argtypes = [C.c_void_p, C.c_int]
    lib.gfe_get_dedup_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                        C.POINTER(C.c_uint64), C.POINTER(C.c_int)]
    lib.gfe_last_kernel_ms.argtypes = [C.c_void_p]
    lib.gfe_last_kernel_ms.restype = C.c_float
    lib.gfe_last_aux_kernel_ms.argtypes = [C.c_void_p]
    lib.gfe_last_aux_kernel_ms.restype = C.c_float
    lib.gfe_device_status.argtypes = [C.c_void_p, _u32p]
    lib.gfe_kernel_time_stats.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_uint32)]
    lib.gfe_trace_pairs.argtypes = [C.c_void_p, _u32p, C.c_uint32, C.c_uint64, C.c_uint32, _u8p, _u8p,
                                    C.c_uint64, C.POINTER(C.c_uint64), _u32p, _i32p]
    lib.gfe_filter_alignments.argtypes = [C.c_void_p, _u32p, C.c_uint32, C.c_int32, _u8p]
    lib.gfe_filter_alignments_static.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32,
                                                 C.c_void_p, C.c_uint32, C.c_int32, C.c_void_p, C.c_void_p,
                                                 C.c_char_p, C.c_size_t]
    lib.gfe_text_open.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_void_p), C.c_char_p,
                                  C.c_size_t]
    lib.gfe_text_close.argtypes = [C.c_void_p]
    lib.gfe_text_close.restype = None
    lib.gfe_text_get.argtypes = [C.c_void_p, C.POINTER(_TextView)]
    lib.gfe_text_node_name.argtypes = [C.c_void_p, C.c_uint32]
    lib.gfe_text_node_name.restype = C.c_char_p
    lib.gfe_text_node_id.argtypes = [C.c_void_p, C.c_char_p]
    lib.gfe_text_node_id.restype = C.c_int64
    lib.gfe_text_qname.argtypes = [C.c_void_p, C.c_uint64, _u32p]
    lib.gfe_text_qname.restype = C.c_void_p
    _lib = lib
    return lib


def _u32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.uint32)


def _ptr(a: np.ndarray, typ=_u32p):
    return a.ctypes.data_as(typ)


@dataclass
class SearchRow:
    """One emitted row of `gfalign search` (reference src/eval.cpp:183)."""
    path_counter: int
    bad: int
    good: int
    alt: int
    size: int
    uniques: int
    hamiltonian: bool
    steps: List[int]

    def columns(self, names: Sequence[str]) -> str:
        path = ",".join(f"{names[s >> 1]}{'-' if s & 1 else '+'}" for s in self.steps)
        return "\t".join([str(self.path_counter), str(self.bad), str(self.good), str(self.alt),
                          str(self.size), str(self.uniques), "T" if self.hamiltonian else "F", path])


class Engine:
    """One engine = one CUDA device.  Thin, typed wrapper over the gfe_* C ABI."""

    def __init__(self, device: int = 0):
        self._lib = load_library()
        h = C.c_void_p()
        rc = self._lib.gfe_create(device, C.byref(h))
        if rc != 0:
            raise GfeError(rc, self._lib.gfe_last_error(None).decode())
        self._h = h
        self._keep = []

    def close(self):
        if getattr(self, "_h", None):
            self._lib.gfe_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc: int):
        if rc != 0:
            raise GfeError(rc, self._lib.gfe_last_error(self._h).decode())

    # -- alignment table ---------------------------------------------------------------------
    def set_alignments(self, offsets, steps, n_nodes: int):
        offsets, steps = _u32(offsets), _u32(steps)
        if offsets.size < 1:
            raise ValueError("offsets must have A+1 entries")
        self._check(self._lib.gfe_set_alignments(self._h, _ptr(offsets), _ptr(steps),
                                                 offsets.size - 1, n_nodes))

    def set_alignments_device(self, d_offsets_ptr: int, d_steps_ptr: int, A: int, total_steps: int,
                              n_nodes: int):
        self._check(self._lib.gfe_set_alignments_device(self._h, d_offsets_ptr, d_steps_ptr, A,
                                                        total_steps, n_nodes))
        self._A_hint = int(A)

    def set_universe(self, nodes):
        nodes = _u32(nodes)
        self._check(self._lib.gfe_set_universe(self._h, _ptr(nodes), nodes.size))

    # -- scoring -----------------------------------------------------------------------------
    def score_paths(self, path_offsets, path_steps, filter: bool = True) -> np.ndarray:
        """evaluatePath for P candidate paths -> uint32 array (P, 2) of (bad, good)."""
        po, ps = _u32(path_offsets), _u32(path_steps)
        P = po.size - 1
        counts = np.zeros((max(P, 0), 2), dtype=np.uint32)
        self._check(self._lib.gfe_score_paths(self._h, _ptr(po), _ptr(ps), P, int(filter),
                                              _ptr(counts)))
        return counts

    def score_paths_incremental(self, base_offsets, base_steps, item_base, suffix_offsets, suffix_steps,
                                mode) -> np.ndarray:
        """gfe_score_paths_incremental: path i = bases[item_base[i]] ++ suffix i -> int32 (P, 4):
        mode 0: bad, good over all alignments, then bad, good over the fresh set F;
        mode 1: bad/good DIFFERENCE to the parent (path minus its last step) outside F, then F."""
        bo, bs, ib = _u32(base_offsets), _u32(base_steps), _u32(item_base)
        so, ss = _u32(suffix_offsets), _u32(suffix_steps)
        md = np.ascontiguousarray(mode, dtype=np.uint8)
        P = ib.size
        out = np.zeros((max(P, 1), 4), dtype=np.int32)
        b = _PathBatch(bo.size - 1, _ptr(bo), _ptr(bs), P, _ptr(ib), _ptr(so), _ptr(ss), _ptr(md, _u8p))
        self._check(self._lib.gfe_score_paths_incremental(self._h, C.byref(b), _ptr(out, _i32p)))
        return out[:P]

    def score_paths_sharded(self, exchange, path_offsets, path_steps, filter: bool = True) -> np.ndarray:
        """gfe_score_paths_sharded: (P, 2) counts summed over every rank's alignment shard."""
        po, ps = _u32(path_offsets), _u32(path_steps)
        P = po.size - 1
        counts = np.zeros((max(P, 0), 2), dtype=np.uint32)
        self._check(self._lib.gfe_score_paths_sharded(self._h, exchange.handle, _ptr(po), _ptr(ps), P,
                                                      int(filter), _ptr(counts)))
        return counts

    def search_sharded(self, exchange, adj_offsets, adj_target, adj_or0, adj_or1, source: int,
                       destination: int, budget_nodes, budget_counts, node_count: int,
                       max_steps: int = 100000, min_nodes: int = 0, return_all_paths: bool = False,
                       speculate: int = 0, on_row=None):
        """gfe_search_sharded: dijkstra() replayed on every rank over the summed shard counts."""
        return _run_search(lambda g, p, cb, st: self._check(
            self._lib.gfe_search_sharded(self._h, exchange.handle, g, p, cb, None, st)),
            adj_offsets, adj_target, adj_or0, adj_or1, source, destination, budget_nodes,
            budget_counts, node_count, max_steps, min_nodes, return_all_paths, speculate, on_row)

    def score_paths_device(self, d_path_offsets: int, d_path_steps: int, P: int, total: int,
                           d_counts: int, filter: bool = True, stream: int = 0):
        self._check(self._lib.gfe_score_paths_device(self._h, d_path_offsets, d_path_steps, P,
                                                     total, int(filter), d_counts, stream))

    def score_pairs(self, path_steps, filter: bool = False, want_scores: bool = True):
        """evalPath scoring: (t_fw, t_rc, bad, good) of one path against every alignment."""
        ps = _u32(path_steps)
        A = self._A_hint
        bad, good = C.c_uint32(), C.c_uint32()
        if want_scores:
            tf = np.empty(A, dtype=np.int32)
            tr = np.empty(A, dtype=np.int32)
            self._check(self._lib.gfe_score_pairs(self._h, _ptr(ps), ps.size, int(filter),
                                                  _ptr(tf, _i32p), _ptr(tr, _i32p),
                                                  C.byref(bad), C.byref(good)))
            return tf, tr, bad.value, good.value
        self._check(self._lib.gfe_score_pairs(self._h, _ptr(ps), ps.size, int(filter), None, None,
                                              C.byref(bad), C.byref(good)))
        return None, None, bad.value, good.value

    def trace_pairs(self, path_steps, first: int, count: int, use_rc, offsets):
        """evalPath detail: per pair the traceback columns (list of uint8 arrays; 0 equal, 1
        different, 2 path step vs gap, 3 alignment step vs gap in the path) and the score."""
        ps = _u32(path_steps)
        off = np.asarray(offsets, dtype=np.int64)
        cap = int(off[first + count] - off[first]) + count * ps.size
        ops = np.zeros(max(cap, 1), dtype=np.uint8)
        rc = np.ascontiguousarray(use_rc, dtype=np.uint8)
        start = np.zeros(max(count, 1), dtype=np.uint64)
        ln = np.zeros(max(count, 1), dtype=np.uint32)
        score = np.zeros(max(count, 1), dtype=np.int32)
        self._check(self._lib.gfe_trace_pairs(self._h, _ptr(ps), ps.size, first, count, _ptr(rc, _u8p),
                                              _ptr(ops, _u8p), cap,
                                              start.ctypes.data_as(C.POINTER(C.c_uint64)), _ptr(ln),
                                              _ptr(score, _i32p)))
        cols = [ops[int(start[k]):int(start[k]) + int(ln[k])] for k in range(count)]
        return cols, score[:count]

    def filter_alignments(self, nodes, min_nodes: int = 0) -> np.ndarray:
        """filter mode: keep[a] = all nodes of alignment a in `nodes` and m >= min_nodes."""
        nd = _u32(nodes)
        keep = np.zeros(max(self._A_hint, 1), dtype=np.uint8)
        self._check(self._lib.gfe_filter_alignments(self._h, _ptr(nd), nd.size, int(min_nodes),
                                                    _ptr(keep, _u8p)))
        return keep[:self._A_hint].astype(bool)

    # -- edge support ------------------------------------------------------------------------
    def edge_support(self, edge_from, edge_to) -> np.ndarray:
        ef, et = _u32(edge_from), _u32(edge_to)
        w = np.zeros(ef.size, dtype=np.uint32)
        self._check(self._lib.gfe_edge_support(self._h, _ptr(ef), _ptr(et), ef.size, _ptr(w)))
        return w

    def edge_traversals_device(self, edge_from, edge_to, d_counts: int, stream: int = 0):
        ef, et = _u32(edge_from), _u32(edge_to)
        self._check(self._lib.gfe_edge_traversals_device(self._h, _ptr(ef), _ptr(et), ef.size,
                                                         d_counts, stream))

    @staticmethod
    def edge_weights_from_counts(edge_from, edge_to, counts) -> np.ndarray:
        lib = load_library()
        ef, et, cn = _u32(edge_from), _u32(edge_to), _u32(counts)
        w = np.zeros(ef.size, dtype=np.uint32)
        rc = lib.gfe_edge_weights_from_counts(_ptr(ef), _ptr(et), ef.size, _ptr(cn), _ptr(w))
        if rc != 0:
            raise GfeError(rc, "gfe_edge_weights_from_counts: bad arguments")
        return w

    # -- search ------------------------------------------------------------------------------
    def search(self, adj_offsets, adj_target, adj_or0, adj_or1, source: int, destination: int,
               budget_nodes, budget_counts, node_count: int, max_steps: int = 100000,
               min_nodes: int = 0, return_all_paths: bool = False, speculate: int = 0,
               on_row: Optional[Callable[[SearchRow], None]] = None):
        """dijkstra() of the reference on packed inputs -> (rows, stats dict)."""
        return _run_search(lambda g, p, cb, st: self._check(
            self._lib.gfe_search(self._h, g, p, cb, None, st)),
            adj_offsets, adj_target, adj_or0, adj_or1, source, destination, budget_nodes,
            budget_counts, node_count, max_steps, min_nodes, return_all_paths, speculate, on_row)

    # -- introspection -----------------------------------------------------------------------
    def counters(self) -> dict:
        c = _Counters()
        self._check(self._lib.gfe_get_counters(self._h, C.byref(c)))
        return {f: getattr(c, f) for f, _ in _Counters._fields_}

    def set_dedup(self, on: bool):
        """Multiplicity compression of the scoring view for full scans; applies at the next set_universe."""
        self._check(self._lib.gfe_set_dedup(self._h, int(on)))

    def dedup_stats(self) -> dict:
        r, d, v, u = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_int()
        self._check(self._lib.gfe_get_dedup_stats(self._h, C.byref(r), C.byref(d), C.byref(v), C.byref(u)))
        return {"rows": r.value, "distinct_walks": d.value, "view_rows": v.value, "in_use": bool(u.value)}

    def set_timing(self, on: bool):
        self._check(self._lib.gfe_set_timing(self._h, int(on)))

    def last_kernel_ms(self) -> float:
        return float(self._lib.gfe_last_kernel_ms(self._h))

    def last_aux_kernel_ms(self) -> float:
        return float(self._lib.gfe_last_aux_kernel_ms(self._h))

    def device_status(self) -> int:
        """Sticky flags of the *_device calls (bit0 path outside universe, bit1 worklist overflow)."""
        f = C.c_uint32()
        self._check(self._lib.gfe_device_status(self._h, C.byref(f)))
        return f.value

    def device_status_to(self, d_flags: int, stream: int = 0):
        """device_status() written to the device word at d_flags, stream-ordered, no synchronisation."""
        self._check(self._lib.gfe_device_status_to(self._h, C.c_void_p(d_flags), C.c_void_p(stream)))

    def kernel_time_stats(self):
        """(total ms, launches) of the dominant kernel since the previous call; synchronises."""
        ms, n = C.c_double(), C.c_uint32()
        self._check(self._lib.gfe_kernel_time_stats(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    # set_alignments records A so score_pairs can size its outputs
    _A_hint = 0


def _run_search(call, adj_offsets, adj_target, adj_or0, adj_or1, source, destination,
                budget_nodes, budget_counts, node_count, max_steps, min_nodes, return_all_paths,
                speculate, on_row):
    ao, at = _u32(adj_offsets), _u32(adj_target)
    o0 = np.ascontiguousarray(adj_or0, dtype=np.uint8)
    o1 = np.ascontiguousarray(adj_or1, dtype=np.uint8)
    bn, bc = _u32(budget_nodes), _u32(budget_counts)
    g = _Graph(ao.size - 1, _ptr(ao), _ptr(at), _ptr(o0, _u8p), _ptr(o1, _u8p))
    p = _SearchParams(source, destination, _ptr(bn), _ptr(bc), bn.size, node_count, max_steps,
                      min_nodes, int(return_all_paths), speculate)
    rows: List[SearchRow] = []

    def _cb(rp, _user):
        r = rp.contents
        row = SearchRow(r.path_counter, r.bad, r.good, r.alt, r.size, r.uniques,
                        bool(r.hamiltonian), [r.steps[i] for i in range(r.size)])
        rows.append(row)
        if on_row:
            on_row(row)

    cb = _ROW_FN(_cb)
    st = _SearchStats()
    call(C.byref(g), C.byref(p), cb, C.byref(st))
    stats = {f: getattr(st, f) for f, _ in _SearchStats._fields_}
    return rows, stats


def search_with_scorer(scorer: Callable[[np.ndarray, np.ndarray], np.ndarray], adj_offsets,
                       adj_target, adj_or0, adj_or1, source: int, destination: int, budget_nodes,
                       budget_counts, node_count: int, max_steps: int = 100000, min_nodes: int = 0,
                       return_all_paths: bool = False, speculate: int = 0,
                       on_row: Optional[Callable[[SearchRow], None]] = None):
    """gfe_search_with_scorer: the reference's dijkstra() replay with evaluatePath answered by
    `scorer(path_offsets, path_steps) -> (P, 2) uint32 (bad, good)` — e.g. alignment shards on
    several GPUs summed with one all-reduce.  Needs no CUDA device by itself."""
    lib = load_library()
    failure = []

    def _score(_user, po, ps, P, counts):
        try:
            off = np.ctypeslib.as_array(po, shape=(P + 1,)).copy()
            st = np.ctypeslib.as_array(ps, shape=(int(off[-1]),)).copy()
            out = np.ascontiguousarray(scorer(off, st), dtype=np.uint32).reshape(-1)
            if out.size != 2 * P:
                raise ValueError("scorer must return P x 2 counts")
            C.memmove(counts, out.ctypes.data, out.nbytes)
            return 0
        except Exception as exc:  # surfaced after the C call returns
            failure.append(exc)
            return -2

    fn = _SCORE_FN(_score)

    def call(g, p, cb, st):
        rc = lib.gfe_search_with_scorer(fn, None, g, p, cb, None, st)
        if failure:
            raise failure[0]
        if rc != 0:
            raise GfeError(rc, "gfe_search_with_scorer failed")

    return _run_search(call, adj_offsets, adj_target, adj_or0, adj_or1, source, destination,
                       budget_nodes, budget_counts, node_count, max_steps, min_nodes,
                       return_all_paths, speculate, on_row)


_orig_set = Engine.set_alignments


def _set_alignments(self, offsets, steps, n_nodes):
    _orig_set(self, offsets, steps, n_nodes)
    self._A_hint = int(np.asarray(offsets).size - 1)


Engine.set_alignments = _set_alignments


class TextPack:
    """GFA1 (+ GAF) text packed by the C++ host packers (gfe_text_*); arrays are copies."""

    def __init__(self, gfa_path: str, gaf_path: Optional[str] = None, threads: int = 0):
        lib = load_library()
        h = C.c_void_p()
        err = C.create_string_buffer(512)
        rc = lib.gfe_text_open(gfa_path.encode(), gaf_path.encode() if gaf_path else None, threads,
                               C.byref(h), err, 512)
        if rc != 0:
            raise GfeError(rc, err.value.decode(errors="replace"))
        try:
            v = _TextView()
            lib.gfe_text_get(h, C.byref(v))
            take = lambda p, n, dt: (np.ctypeslib.as_array(p, shape=(n,)).astype(dt).copy()  # noqa: E731
                                     if n else np.zeros(0, dtype=dt))
            self.n_nodes = v.n_nodes
            self.names = [lib.gfe_text_node_name(h, i).decode() for i in range(v.n_nodes)]
            self.edge_from = take(v.edge_from, v.n_links, np.uint32)
            self.edge_to = take(v.edge_to, v.n_links, np.uint32)
            self.adj_offsets = take(v.graph.adj_offsets, v.n_nodes + 1, np.uint32)
            n_adj = int(self.adj_offsets[-1]) if v.n_nodes else 0
            self.adj_target = take(v.graph.adj_target, n_adj, np.uint32)
            self.adj_or0 = take(v.graph.adj_or0, n_adj, np.uint8)
            self.adj_or1 = take(v.graph.adj_or1, n_adj, np.uint8)
            self.A = int(v.A)
            self.offsets = take(v.offsets, self.A + 1, np.uint32) if gaf_path else np.zeros(1, np.uint32)
            self.steps = take(v.steps, int(self.offsets[-1]), np.uint32)
            self.qnames = []
            ln = C.c_uint32()
            for a in range(self.A if self.A <= 200000 else 0):
                p = lib.gfe_text_qname(h, a, C.byref(ln))
                self.qnames.append(C.string_at(p, ln.value).decode(errors="replace"))
            assert all(lib.gfe_text_node_id(h, nm.encode()) == i
                       for i, nm in enumerate(self.names[:1000]))
            self._ids = {nm: i for i, nm in enumerate(self.names)}
            self.node_id = lambda name: self._ids.get(name, -1)
        finally:
            lib.gfe_text_close(h)
