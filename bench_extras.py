#!/usr/bin/env python
"""bench_extras.py — the other kernels of the path on BASELINE's
cfg4 shape (500 k nodes, 20 M alignments), each with its roofline, the host-to-host call (`e2e`) and
the CPU baselines timed on the same box.  bench.py imports `run_cfg4()` and puts the result into its
JSON line as `extras`; run by hand it prints one JSON object.

  python bench_extras.py [--alignments A] [--reps R] [--no-cpu]

  evalgfa   K3 edge_stream_kernel      edge-steps/s; algorithmic 4(m+1) B per alignment vs HBM
  evalpath  K2 pair_classify + exact   pairs/s of one 1000-step path vs every alignment; DP cells/s
  trace     pair_trace_kernel          DP cells/s
  filter    filter_mark/rows           alignments/s, GB/s
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def write_gfa_fast(path, t):
    with open(path, "w") as f:
        f.write("H\tVN:Z:1.2\n")
        f.write("".join(f"S\t{nm}\t*\n" for nm in t.names))
        ef, et = t.edge_from.tolist(), t.edge_to.tolist()
        f.write("".join(f"L\t{t.names[x >> 1]}\t{'-' if x & 1 else '+'}\t{t.names[y >> 1]}\t"
                        f"{'-' if y & 1 else '+'}\t0M\n" for x, y in zip(ef, et)))


def write_gaf_fast(path, names, offsets, steps):
    off = offsets.tolist()
    st = steps[:off[-1]].tolist()
    tok = [(">" if not (s & 1) else "<") + names[s >> 1] for s in range(2 * len(names))] \
        if len(names) <= 100000 else None
    with open(path, "w") as f:
        buf = []
        for a in range(len(off) - 1):
            seg = st[off[a]:off[a + 1]]
            walk = "".join(tok[s] for s in seg) if tok else \
                "".join((">" if not (s & 1) else "<") + names[s >> 1] for s in seg)
            ln = 4 * max(len(seg), 1)
            buf.append(f"read{a}\t{ln}\t0\t{ln}\t+\t{walk}\t{ln}\t0\t{ln}\t{ln}\t{ln}\t60\n")
            if len(buf) >= 100000:
                f.write("".join(buf))
                buf = []
        f.write("".join(buf))


# backwards-compatible names (profiles/*.py of round 1)
_write_gfa_fast, _write_gaf_fast = write_gfa_fast, write_gaf_fast


def reference_cfg4_scaled(timeout_s: float = 120.0):
    """The UNMODIFIED reference CLI on a cfg4-SHAPED sample it can finish: 5 000 nodes / 7 000 links
    (1 % of cfg4's graph: the reference's evalGFA does O(segments) scans per record — it needs > 6 min
    on the 500 k-node graph even with 1 % of the alignments, measured in the build container) and
    200 000 alignments for evalGFA, 20 000 for evalPath on a 1000-step path.  Hot-path seconds =
    wall(tool) - wall(same inputs, hot path off).  Returns {} when oracle/_ref is absent."""
    from gfalign_b200 import synth
    from oracle import binding
    binary = binding.ref_binary()
    if not binary:
        return {}
    out = {}
    t = synth.make_tangle(5000, 2000, seed=2, minus_frac=0.05, count2_frac=0.0)
    off, st = synth.random_walk_alignments(t.n_nodes, t.edge_from, t.edge_to, 200_000,
                                           *synth.HIFI_LENGTHS, seed=12)

    def wall(cmd, **kw):
        t0 = time.perf_counter()
        subprocess.run(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=True,
                       timeout=timeout_s, **kw)
        return time.perf_counter() - t0

    with tempfile.TemporaryDirectory() as wd:
        gfa, gaf, gaf2 = (os.path.join(wd, x) for x in ("g.gfa", "a.gaf", "b.gaf"))
        write_gfa_fast(gfa, t)
        write_gaf_fast(gaf, t.names, off, st)
        n2 = 20_000
        write_gaf_fast(gaf2, t.names, off[:n2 + 1], st)
        try:
            base = wall([binary, "evalGFA", "-j", "1", "-f", gfa, "-g", gaf])
            full = wall([binary, "evalGFA", "-j", "1", "-f", gfa, "-g", gaf, "-o", os.path.join(wd, "o.gfa")])
            ev = int(off[-1]) - 200_000
            out["evalgfa"] = {"value": ev / max(full - base, 1e-9), "unit": "edge-steps/s", "cores": 1,
                              "kind": "reference",
                              "sample": f"cfg4 shape at 1 % (5 000 nodes, 7 000 links, 200 000 alignments, {ev} "
                                        f"edge-steps): `evalGFA -o` {full:.2f} s minus `evalGFA` {base:.2f} s "
                                        "(edge support + GFA write); single-threaded hot path"}
        except Exception as ex:   # noqa: BLE001
            out["evalgfa"] = {"unavailable": repr(ex)[:200]}
        try:
            path = ",".join(f"{t.names[i]}+" for i in range(1000))
            load = wall([binary, "evalGFA", "-j", "1", "-f", gfa, "-g", gaf2])
            full = wall([binary, "evalPath", "-j", "1", "-f", gfa, "-g", gaf2, "-p", path])
            cells = 2.0 * 1000 * float(off[n2])
            out["evalpath"] = {"value": n2 / max(full - load, 1e-9), "unit": "pairs/s", "cores": 1,
                               "kind": "reference", "dp_cells_per_s": cells / max(full - load, 1e-9),
                               "sample": f"cfg4 shape at 1 % of the graph, {n2} alignments, 1000-step path: "
                                         f"`evalPath` {full:.2f} s minus load {load:.2f} s (`evalGFA` without -o "
                                         "on the same files); stdout to /dev/null"}
        except Exception as ex:   # noqa: BLE001
            out["evalpath"] = {"unavailable": repr(ex)[:200]}
    return out


def run_cfg4(device_index: int = 0, A: int = 20_000_000, reps: int = 5, cpu: bool = True,
             trace_pairs: int = 500_000, flush=None):
    """-> dict of sections for cfg4 (see module docstring).  The table is drawn on the device."""
    import torch
    import gfalign_b200
    from gfalign_b200 import capi, synth
    dev = torch.device("cuda", device_index)
    peak, peak_src = peak_gbs()
    if flush is None:
        flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)
    t0 = time.perf_counter()
    t = synth.make_tangle(500_000, 200_000, seed=2, minus_frac=0.05, count2_frac=0.0)
    d_off, d_st = synth.random_walk_alignments_device(t.n_nodes, t.edge_from, t.edge_to, A,
                                                      *synth.HIFI_LENGTHS, seed=12, device=dev)
    torch.cuda.synchronize(dev)
    gen_s = time.perf_counter() - t0
    S = int(d_off[-1].item())
    mbar = S / A
    out = {"workload": "cfg4 synthetic human-scale graph: 500k nodes, 700k links, 20M alignments "
                       "(random walks, HiFi length pmf); evalGFA + evalPath on a 1000-step path",
           "nodes": t.n_nodes, "links": int(t.edge_from.size), "alignments": A, "mean_steps": mbar,
           "hbm_peak_gbs": peak, "peak_source": peak_src, "synth_seconds": gen_s,
           "l2": "flushed (512 MiB write) before every timed launch"}
    off_h = d_off.cpu().numpy().view(np.uint32)
    st_h = d_st.cpu().numpy().view(np.uint32)
    eng = gfalign_b200.Engine(device_index)
    t0 = time.perf_counter()
    eng.set_alignments(off_h, st_h, t.n_nodes)          # host CSR in: the upload is part of a user's first call
    out["table_upload_seconds"] = time.perf_counter() - t0
    del d_off, d_st
    eng.set_timing(True)
    alg_bytes = 4.0 * (S + A)      # every step + one offset per alignment, read once

    def timed(fn):
        """(mean ms of the dominant kernel, mean wall seconds of the whole host call)"""
        ms, wall = [], []
        for r in range(reps + 2):
            flush.zero_()
            torch.cuda.synchronize(dev)
            w0 = time.perf_counter()
            fn()
            w1 = time.perf_counter()
            if r >= 2:
                ms.append(eng.last_kernel_ms())
                wall.append(w1 - w0)
        return float(np.mean(ms)), float(np.mean(wall))

    def roof(nbytes, ms, note=None):
        r = {"bound": "hbm", "achieved": nbytes / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
             "frac": nbytes / (ms * 1e-3) / 1e9 / peak, "traffic": None,
             "algorithmic_bytes_per_launch": nbytes, "kernel_ms": ms}
        if note:
            r["note"] = note
        return r

    ref = reference_cfg4_scaled() if cpu else {}
    orc = None
    if cpu:
        from oracle.binding import Oracle
        orc = Oracle()

    # ---- evalGFA: K3 ------------------------------------------------------------------------------
    c0 = eng.counters()
    weights = eng.edge_support(t.edge_from, t.edge_to)
    c1 = eng.counters()
    k_ms, call_s = timed(lambda: eng.edge_support(t.edge_from, t.edge_to))
    sec = {"kernel": "edge_stream_kernel", "kernel_ms": k_ms, "edge_steps_per_s": (S - A) / (k_ms * 1e-3),
           "alignments_per_s": A / (k_ms * 1e-3),
           "roofline": roof(alg_bytes, k_ms, "algorithmic = every step + one offset per alignment, once; the kernel "
                                             "issues one table probe + one RED per edge-step: LSU/L2-atomic bound"),
           "e2e": {"value": (S - A) / call_s, "unit": "edge-steps/s", "call": "gfe_edge_support(host link records) -> "
                   "host RC:i weights; alignment table resident", "seconds": call_s,
                   "h2d_bytes_per_step": int(c1["h2d_bytes"] - c0["h2d_bytes"]),
                   "d2h_bytes_per_step": int(c1["d2h_bytes"] - c0["d2h_bytes"])},
           "weight_sum": int(weights.sum(dtype=np.uint64))}
    if cpu:
        a_s = min(A, 4_000_000)
        t0 = time.perf_counter()
        ow = orc.edge_support(off_h[:a_s + 1], st_h[:int(off_h[a_s])], t.n_nodes, t.edge_from, t.edge_to)
        dt = time.perf_counter() - t0
        sec["cpu_baseline"] = {"value": (int(off_h[a_s]) - a_s) / dt, "unit": "edge-steps/s", "cores": 1, "kind": "port",
                               "sample": f"oracle port on the first {a_s} alignments of the same CSR and the full "
                                         f"graph ({dt:.2f} s)", "reference_cli": ref.get("evalgfa")}
        if a_s == A:
            sec["parity_check"] = {"ok": bool(np.array_equal(ow, weights)), "records": int(weights.size)}
    out["evalgfa"] = sec

    # ---- evalPath: K2 -----------------------------------------------------------------------------
    path = (np.arange(1000, dtype=np.uint32) << 1)
    n = int(path.size)
    c0 = eng.counters()
    tf, tr, bad, good = eng.score_pairs(path, filter=False)
    c1 = eng.counters()
    exact = int(c1["exact_pairs"] - c0["exact_pairs"])
    k_ms, call_s = timed(lambda: eng.score_pairs(path, filter=False, want_scores=False))
    exact_ms = eng.last_aux_kernel_ms()
    view_bytes = 2.0 * (S + A) + 8.0 * A
    sec = {"kernel": "pair_classify_kernel (+ pair_exact_kernel for rows touching the path)", "path_steps": n,
           "classify_kernel_ms": k_ms, "exact_dp_kernel_ms": exact_ms, "call_seconds_totals_only": call_s, "pairs_per_s": A / call_s,
           "exact_dp_pairs": exact, "dp_cells_dense_equivalent": 2.0 * n * mbar * exact, "bad": int(bad),
           "good": int(good),
           "roofline": roof(view_bytes, k_ms, "classify pass streams the 16-bit view (2(m+1) B) + 8 B of scores "
                                              "per row; rows touching the path go to the sparse-row exact DP"),
           "e2e": {"value": A / call_s, "unit": "pairs/s", "seconds": call_s,
                   "call": "gfe_score_pairs(host path) -> host totals (bad, good); set_universe included",
                   "h2d_bytes_per_step": 16 + 2 * n, "d2h_bytes_per_step": 16}}
    if cpu:
        a_s = min(A, 100_000)
        t0 = time.perf_counter()
        ob, og, otf, otr = orc.evaluate_path(path, off_h[:a_s + 1], st_h[:int(off_h[a_s])], False, scores=True)
        dt = time.perf_counter() - t0
        sec["cpu_baseline"] = {"value": a_s / dt, "unit": "pairs/s", "cores": 1, "kind": "port",
                               "dp_cells_per_s": 2.0 * n * float(off_h[a_s]) / dt,
                               "sample": f"oracle port, the same path against the first {a_s} alignments ({dt:.2f} s)",
                               "reference_cli": ref.get("evalpath")}
        sec["parity_check"] = {"ok": bool(np.array_equal(tf[:a_s], otf) and np.array_equal(tr[:a_s], otr)),
                               "pairs": a_s}
    out["evalpath"] = sec

    # ---- evalPath detail: traceback columns ----------------------------------------------------------
    cnt = min(A, trace_pairs)
    rc = np.zeros(cnt, dtype=np.uint8)
    off64 = off_h.astype(np.int64)
    cap = int(off64[cnt] - off64[0]) + cnt * n
    ops = np.zeros(cap, dtype=np.uint8)
    start = np.zeros(cnt, dtype=np.uint64)
    ln = np.zeros(cnt, dtype=np.uint32)
    sc = np.zeros(cnt, dtype=np.int32)
    p32 = np.ascontiguousarray(path, dtype=np.uint32)

    def trace_call():
        rcode = eng._lib.gfe_trace_pairs(eng._h, p32.ctypes.data_as(capi._u32p), n, 0, cnt,
                                         rc.ctypes.data_as(capi._u8p), ops.ctypes.data_as(capi._u8p),
                                         cap, start.ctypes.data_as(C.POINTER(C.c_uint64)),
                                         ln.ctypes.data_as(capi._u32p), sc.ctypes.data_as(capi._i32p))
        assert rcode == 0, rcode
    trace_call()
    k_ms, call_s = timed(trace_call)
    cells = float(n) * float(off64[cnt] - off64[0])
    sec = {"kernel": "pair_trace_kernel", "path_steps": n, "pairs": cnt, "kernel_ms_last_subbatch": k_ms,
           "dp_cells": cells, "dp_cells_per_s_call": cells / call_s,
           "e2e": {"value": cnt / call_s, "unit": "pairs/s", "seconds": call_s,
                   "call": "gfe_trace_pairs(host path) -> host op bytes (n + m per pair)",
                   "h2d_bytes_per_step": 4 * n + cnt, "d2h_bytes_per_step": cap + 16 * cnt},
           "scores_agree_with_score_pairs": bool(np.array_equal(sc, tf[:cnt]))}
    out["trace"] = sec

    # ---- filter mode -----------------------------------------------------------------------------------
    nodes = np.arange(0, t.n_nodes, 2, dtype=np.uint32)
    keep = eng.filter_alignments(nodes, 2)
    k_ms, call_s = timed(lambda: eng.filter_alignments(nodes, 2))
    sec = {"kernel": "filter_mark_kernel + filter_rows_kernel", "kernel_ms": k_ms, "kept": int(keep.sum()),
           "alignments_per_s": A / (k_ms * 1e-3), "roofline": roof(alg_bytes + A, k_ms),
           "e2e": {"value": A / call_s, "unit": "alignments/s", "seconds": call_s,
                   "call": "gfe_filter_alignments(host node set) -> host keep[A]",
                   "h2d_bytes_per_step": int((t.n_nodes + 31) // 32 * 4), "d2h_bytes_per_step": A}}
    if cpu:
        inset = np.zeros(t.n_nodes, dtype=bool)
        inset[nodes] = True
        t0 = time.perf_counter()
        ok = inset[st_h >> 1].astype(np.int64)
        cs = np.concatenate([[0], np.cumsum(ok)])
        m = np.diff(off64)
        want = (cs[off64[1:]] - cs[off64[:-1]] == m) & (m >= 2)
        dt = time.perf_counter() - t0
        sec["cpu_baseline"] = {"value": A / dt, "unit": "alignments/s", "cores": 1, "kind": "port",
                               "sample": f"numpy restatement of isContained + pathNodesCount on the full CSR ({dt:.2f} s)"}
        sec["parity_check"] = {"ok": bool(np.array_equal(keep, want)), "alignments": A}
    out["filter"] = sec
    eng.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--alignments", type=int, default=20_000_000)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--trace-pairs", type=int, default=500_000)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench_extras.py needs a CUDA device: gfalign_b200 has no CPU fallback")
    print(json.dumps(run_cfg4(0, args.alignments, args.reps, not args.no_cpu, args.trace_pairs)))


if __name__ == "__main__":
    main()
