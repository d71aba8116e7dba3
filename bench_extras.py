#!/usr/bin/env python
"""bench_extras.py — the other kernels of the path against their
rooflines, one JSON line per measurement.  Not the driver's contract (that is bench.py); run by hand
on the GPU box and committed under profiles/.

  python bench_extras.py [--workload cfg3|cfg4] [--alignments A] [--reps R]

  evalgfa   K3 edge_support_kernel     alignments/s, algorithmic 4(m+1) B per alignment vs HBM
  evalpath  K2 pair_classify (+exact)  pairs/s of one n-step path vs every alignment, GB/s, DP cells/s
  trace     pair_trace_kernel          DP cells/s (2 decision bits per cell written + read back)
  filter    filter_rows_kernel         alignments/s, GB/s
  ingest    gfe_text_open              GAF text -> CSR, MB/s and records/s (host threads)
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except Exception:
        return 6650.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg3", choices=["cfg3", "cfg4"])
    ap.add_argument("--alignments", type=int, default=0)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--path-steps", type=int, default=0, help="evalPath path length (default: 200 / 1000)")
    ap.add_argument("--trace-pairs", type=int, default=500_000)
    ap.add_argument("--ingest-records", type=int, default=2_000_000)
    ap.add_argument("--skip", default="", help="comma list of sections to skip")
    args = ap.parse_args()
    skip = set(args.skip.split(",")) if args.skip else set()

    import torch
    import gfalign_b200
    from gfalign_b200 import capi, synth
    if not torch.cuda.is_available():
        raise SystemExit("bench_extras.py needs a CUDA device: gfalign_b200 has no CPU fallback")
    peak = peak_gbs()
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    t0 = time.perf_counter()
    if args.workload == "cfg3":
        w = synth.make_workload("cfg3", A=args.alignments or 1_000_000)
        n_path = args.path_steps or 150
    else:
        w = synth.make_workload("cfg4", A=args.alignments or 20_000_000)
        n_path = args.path_steps or 1000
    t = w.tangle
    A = w.offsets.size - 1
    S = int(w.offsets[-1])
    mbar = S / A
    gen_s = time.perf_counter() - t0
    base = {"workload": args.workload, "nodes": t.n_nodes, "links": int(t.edge_from.size),
            "alignments": A, "mean_steps": mbar, "hbm_peak_gbs": peak, "synth_seconds": gen_s}

    def emit(section, **kw):
        print(json.dumps({"section": section, **base, **kw}), flush=True)

    eng = gfalign_b200.Engine(0)
    t0 = time.perf_counter()
    eng.set_alignments(w.offsets, w.steps, t.n_nodes)
    upload_s = time.perf_counter() - t0
    eng.set_timing(True)
    alg_bytes = 4.0 * (S + A)      # every step + one offset per alignment, read once

    def timed(fn):
        ms = []
        for r in range(args.reps + 2):
            flush.zero_()
            torch.cuda.synchronize()
            fn()
            if r >= 2:
                ms.append(eng.last_kernel_ms())
        return float(np.mean(ms)), float(np.min(ms))

    if "evalgfa" not in skip:
        wall0 = time.perf_counter()
        weights = eng.edge_support(t.edge_from, t.edge_to)
        wall = time.perf_counter() - wall0
        mean_ms, min_ms = timed(lambda: eng.edge_support(t.edge_from, t.edge_to))
        emit("evalgfa", kernel="edge_support_kernel", kernel_ms=mean_ms, kernel_ms_min=min_ms,
             alignments_per_s=A / (mean_ms * 1e-3), edge_steps_per_s=(S - A) / (mean_ms * 1e-3),
             roofline={"bound": "hbm", "achieved": alg_bytes / (mean_ms * 1e-3) / 1e9, "peak": peak,
                       "unit": "GB/s", "frac": alg_bytes / (mean_ms * 1e-3) / 1e9 / peak,
                       "algorithmic_bytes_per_launch": alg_bytes},
             e2e_seconds_first_call=wall, weight_sum=int(weights.sum(dtype=np.uint64)),
             table_upload_seconds=upload_s)

    # evalPath: a candidate walk from the source (cfg3) / the first n chain nodes (cfg4)
    if args.workload == "cfg3":
        po, ps = synth.candidate_paths(t, 64, seed=3, max_len=n_path)
        k = int(np.argmax(np.diff(po)))
        path = ps[po[k]:po[k + 1]]
    else:
        path = (np.arange(n_path, dtype=np.uint32) << 1)
    n = int(path.size)
    if "evalpath" not in skip:
        c0 = eng.counters()
        wall0 = time.perf_counter()
        tf, tr, bad, good = eng.score_pairs(path, filter=False)
        wall = time.perf_counter() - wall0
        c1 = eng.counters()
        mean_ms, min_ms = timed(lambda: eng.score_pairs(path, filter=False, want_scores=False))
        w0 = time.perf_counter()
        for _ in range(3):
            eng.score_pairs(path, filter=False, want_scores=False)
        call_s = (time.perf_counter() - w0) / 3
        exact = int(c1["exact_pairs"] - c0["exact_pairs"])
        view_bytes = 2.0 * (S + A) + 8.0 * A
        emit("evalpath", kernel="pair_classify_kernel (+ pair_exact)", path_steps=n,
             classify_kernel_ms=mean_ms, classify_kernel_ms_min=min_ms, call_seconds_totals_only=call_s,
             call_seconds_with_scores_first=wall, pairs_per_s_call=A / call_s,
             exact_dp_pairs=exact, dp_cells_exact=2.0 * n * mbar * exact, bad=int(bad), good=int(good),
             roofline={"bound": "hbm", "achieved": view_bytes / (mean_ms * 1e-3) / 1e9, "peak": peak,
                       "unit": "GB/s", "frac": view_bytes / (mean_ms * 1e-3) / 1e9 / peak,
                       "algorithmic_bytes_per_launch": view_bytes,
                       "note": "classify pass streams the 16-bit view (2(m+1) B) + 8 B of scores per row"})

    if "trace" not in skip:
        cnt = min(A, args.trace_pairs)
        rc = np.zeros(cnt, dtype=np.uint8)
        wall0 = time.perf_counter()
        cols, score = eng.trace_pairs(path, 0, cnt, rc, w.offsets) if cnt <= 20000 else (None, None)
        # large batches: call the ABI directly, skip the per-pair python slicing
        import ctypes as C
        off = w.offsets.astype(np.int64)
        cap = int(off[cnt] - off[0]) + cnt * n
        ops = np.zeros(cap, dtype=np.uint8)
        start = np.zeros(cnt, dtype=np.uint64)
        ln = np.zeros(cnt, dtype=np.uint32)
        sc = np.zeros(cnt, dtype=np.int32)
        p32 = np.ascontiguousarray(path, dtype=np.uint32)

        def call():
            rcode = eng._lib.gfe_trace_pairs(eng._h, p32.ctypes.data_as(capi._u32p), n, 0, cnt,
                                             rc.ctypes.data_as(capi._u8p), ops.ctypes.data_as(capi._u8p),
                                             cap, start.ctypes.data_as(C.POINTER(C.c_uint64)),
                                             ln.ctypes.data_as(capi._u32p), sc.ctypes.data_as(capi._i32p))
            assert rcode == 0, rcode
        w0 = time.perf_counter()
        call()
        call_s = time.perf_counter() - w0
        mean_ms, min_ms = timed(call)
        cells = float(n) * float(off[cnt] - off[0])
        emit("trace", kernel="pair_trace_kernel", path_steps=n, pairs=cnt, kernel_ms_last_subbatch=mean_ms,
             call_seconds=call_s, pairs_per_s_call=cnt / call_s, dp_cells=cells,
             dp_cells_per_s_call=cells / call_s, ops_bytes=cap, score_sum=int(sc.sum(dtype=np.int64)))

    if "filter" not in skip:
        nodes = np.arange(0, t.n_nodes, 2, dtype=np.uint32)
        keep = eng.filter_alignments(nodes, 2)
        mean_ms, min_ms = timed(lambda: eng.filter_alignments(nodes, 2))
        emit("filter", kernel="filter_rows_kernel", kernel_ms=mean_ms, kernel_ms_min=min_ms,
             kept=int(keep.sum()), alignments_per_s=A / (mean_ms * 1e-3),
             roofline={"bound": "hbm", "achieved": (alg_bytes + A) / (mean_ms * 1e-3) / 1e9, "peak": peak,
                       "unit": "GB/s", "frac": (alg_bytes + A) / (mean_ms * 1e-3) / 1e9 / peak,
                       "algorithmic_bytes_per_launch": alg_bytes + A})

    if "ingest" not in skip:
        n_rec = min(A, args.ingest_records)
        with tempfile.TemporaryDirectory() as wd:
            gfa, gaf = os.path.join(wd, "g.gfa"), os.path.join(wd, "a.gaf")
            synth.write_gfa(gfa, t) if t.n_nodes <= 5000 else _write_gfa_fast(gfa, t)
            _write_gaf_fast(gaf, t.names, w.offsets[:n_rec + 1], w.steps)
            size = os.path.getsize(gaf)
            res = {}
            for threads in (1, 0):
                w0 = time.perf_counter()
                tp = _open_only(gfa, gaf, threads)
                res["all" if threads == 0 else "1"] = time.perf_counter() - w0
            emit("ingest", call="gfe_text_open (GFA + GAF -> CSR)", records=n_rec, gaf_bytes=size,
                 host_cores=os.cpu_count(), seconds_1_thread=res["1"], seconds_all_threads=res["all"],
                 mb_per_s_all_threads=size / res["all"] / 1e6, records_per_s_all_threads=n_rec / res["all"],
                 checked=bool(tp))
    eng.close()


def _write_gfa_fast(path, t):
    with open(path, "w") as f:
        f.write("H\tVN:Z:1.2\n")
        f.write("".join(f"S\t{nm}\t*\n" for nm in t.names))
        ef, et = t.edge_from.tolist(), t.edge_to.tolist()
        f.write("".join(f"L\t{t.names[x >> 1]}\t{'-' if x & 1 else '+'}\t{t.names[y >> 1]}\t"
                        f"{'-' if y & 1 else '+'}\t0M\n" for x, y in zip(ef, et)))


def _write_gaf_fast(path, names, offsets, steps):
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


def _open_only(gfa, gaf, threads):
    import ctypes as C
    from gfalign_b200 import capi
    lib = capi.load_library()
    h = C.c_void_p()
    err = C.create_string_buffer(512)
    rc = lib.gfe_text_open(gfa.encode(), gaf.encode(), threads, C.byref(h), err, 512)
    assert rc == 0, err.value
    v = capi._TextView()
    lib.gfe_text_get(h, C.byref(v))
    ok = v.A > 0
    lib.gfe_text_close(h)
    return ok


if __name__ == "__main__":
    main()
