#!/usr/bin/env python
"""bench.py — headline benchmark of the gfalign hot path on B200 (contract: see DESIGN.md §6).

One "step" = one pass of the hot path over one batch: P candidate paths (the kind `gfalign search`
expands) scored against every alignment of the resident table with the id filter on
(evaluatePath x P, reference src/eval.cpp:67-108).  metric = path x alignment checks per second,
whole job.  N > 1 (torchrun): every rank holds its own alignment shard of the same shape (weak
scaling, `value`), scores the same batch, and the P x {bad, good} records are summed by the peer-memory
exchange kernel inside the timed step.  Every run — at every N — checks a few of the scored paths
against the CPU oracle (`parity_check`) and fails loudly when they differ.

Blocks of the one JSON line (beside the contract's keys):
  parity_check   oracle spot check of the timed configuration (all shards), peer and NCCL exchange
  roofline       dominant kernel against the SM issue roof (it is integer-ALU/issue bound), with the
                 HBM view (algorithmic bytes, measured DRAM traffic) next to it; counters from a live
                 `ncu --metrics` child run when ncu is on the box, else from the committed capture
  e2e            host buffers in, host counts out, through the C ABI (N > 1: gfe_score_paths_sharded)
  search         `gfalign search -m 100000` driver: search paths/s (BASELINE's second metric)
  strong_cfg5    BASELINE configs[4]: 50 M alignments split over the N GPUs (N = 1: the whole table)
  extras         N = 1: BASELINE configs[3] (cfg4): evalGFA / evalPath / trace / filter kernels
  cpu_baseline   N = 1: the unmodified reference CLI on a bounded sample + the oracle port

  python bench.py [--gpus N] [--steps K] [--warmup W] [--paths P] [--no-extras] [--no-cfg5] ...
  python bench.py --impl reference ...   # the reference's own CPU implementation, bounded sample
"""
from __future__ import annotations

import argparse
import csv
import io
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "path_x_alignment_checks_per_sec"
UNIT = "checks/s"
# the committed capture the roofline falls back to when no live ncu run is possible
FALLBACK_CAPTURE = "profiles/r02_score_bitset_ncu_counters.json"


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.stop_flag = threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True,
                                     text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            for nm, v in zip(names, s[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(self.samples)}


# ---- reference / CPU baselines -----------------------------------------------------------------

REF_SAMPLE_A, REF_SAMPLE_M = 20000, 60      # oracle-port fallback only (no reference binary on the box)
# The reference arm runs the reference on the arm's OWN table (cfg3, 1 M alignments: the table of GPU rank 0,
# same generator and seed) — not on a subsample.  One evaluatePath call there costs ~0.35 s (it copies all A
# paths and zero-fills its 4 MB dp per call, reference src/eval.cpp:67,79), so a step is `search -m 4`: the 8
# evaluatePath calls = 8 M path x alignment checks of the same workload.
REF_FULL_M = 4
CFG3_DESC = "cfg3 synthetic HiFi tangle: 200-node utig4 nodelist, 1M q60 alignments per GPU"


def arm_config(A, mbar, P, world, exchange):
    """`config` of the bench line — the same dict for the GPU arm and the reference arm."""
    return {"workload": CFG3_DESC, "alignments_per_gpu": A, "mean_steps": mbar,
            "candidate_paths_per_step": P, "filter": True,
            "l2": "flushed (512 MiB write) before every timed step",
            "parallelism": f"alignment shards x{world}, counts summed by {exchange}" if world > 1 else "single GPU"}


def reference_sample(workdir: str, A: int = REF_SAMPLE_A):
    """Text files of a bounded cfg3 sample for the reference CLI."""
    from gfalign_b200 import synth
    w = synth.make_workload("cfg3", A=A)
    t = w.tangle
    gfa, gaf, nl = (os.path.join(workdir, x) for x in ("s.gfa", "s.gaf", "s.tsv"))
    synth.write_gfa(gfa, t)
    synth.write_gaf(gaf, t.names, w.offsets, w.steps)
    synth.write_nodelist(nl, t)
    return w, gfa, gaf, nl


def _ref_cmd(binary, gfa, gaf, nl, t):
    return [binary, "search", "-j", "1", "-f", gfa, "-g", gaf, "-n", nl, "-s", t.names[t.source],
            "-d", t.names[t.destination]]


def reference_cli_setup(binary: str, gfa: str, gaf: str, nl: str, t, m_steps: int):
    """Taken ONCE per arm: seconds of start-up + input parsing (`-m 0`, the faster of two runs) and the number of
    evaluatePath calls of `-m m_steps`, read from the --verbose log (src/eval.cpp:145; deterministic)."""
    base = _ref_cmd(binary, gfa, gaf, nl, t)
    load = 1e30
    for _ in range(2):
        t0 = time.perf_counter()
        subprocess.run(base + ["-m", "0"], capture_output=True, check=True, timeout=900)
        load = min(load, time.perf_counter() - t0)
    v = subprocess.run(base + ["-m", str(m_steps), "--verbose"], capture_output=True, text=True, timeout=900)
    return load, v.stderr.count("We can visit this node")


def time_reference_cli(binary: str, gfa: str, gaf: str, nl: str, t, m_steps: int, load=None, calls=None):
    """Hot-path seconds of ONE `gfalign search -m m_steps` run = wall(-m m_steps) - wall(-m 0)."""
    if load is None or calls is None:
        load, calls = reference_cli_setup(binary, gfa, gaf, nl, t, m_steps)
    t0 = time.perf_counter()
    subprocess.run(_ref_cmd(binary, gfa, gaf, nl, t) + ["-m", str(m_steps)], capture_output=True, check=True,
                   timeout=900)
    full = time.perf_counter() - t0
    return max(full - load, 1e-9), calls, load


REF_WORKLOAD = ("cfg3 HiFi tangle (200 nodes, 1M alignments) — EXTRAPOLATED SAMPLE: the reference CLI runs "
                f"on a {REF_SAMPLE_A}-alignment subsample (2% of the table) with `search -m {REF_SAMPLE_M}`; its "
                "cost per check is flat in A (one loop over the alignments per evaluatePath call, "
                "src/eval.cpp:80-99), so checks/s carries over, paths/s scales with 1/A")


def run_reference_arm(args):
    """--impl reference: the UNMODIFIED reference CLI (oracle/_ref/gfalign, built by
    oracle/Makefile from the reference sources) on a bounded cfg3 sample; its hot path is
    single-threaded, so cores = 1.  Falls back to the oracle port when the binary
    is not there."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import binding
    binary = binding.ref_binary()
    A_s, m_steps = (args.alignments or 1_000_000, REF_FULL_M) if binary else (REF_SAMPLE_A, REF_SAMPLE_M)
    rates, prates = [], []
    with tempfile.TemporaryDirectory() as wd:
        w, gfa, gaf, nl = reference_sample(wd, A_s)
        t = w.tangle
        mbar = float(w.offsets[-1]) / A_s
        if binary:
            kind, cores = "reference", 1
            load, calls = reference_cli_setup(binary, gfa, gaf, nl, t, m_steps)
            sample = (f"the arm's own cfg3 table (A={A_s} alignments, the table of GPU rank 0), unmodified reference "
                      f"CLI `gfalign search -j 1 -m {m_steps}` per step: {calls} evaluatePath calls = {calls * A_s} "
                      f"checks; hot-path time = wall - {load:.2f} s (start-up + parsing, `-m 0`, taken once); the "
                      "hot path is single-threaded (src/eval.cpp:134-189)")
            for i in range(args.warmup + args.steps):
                sec, _, _ = time_reference_cli(binary, gfa, gaf, nl, t, m_steps, load, calls)
                if i >= args.warmup:
                    rates.append(calls * A_s / sec)
                    prates.append(calls / sec)
            ms = 1e3 * calls * A_s / np.mean(rates)
        else:
            from gfalign_b200 import synth
            orc = binding.Oracle()
            kind, cores = "port", os.cpu_count() or 1
            po, ps = synth.candidate_paths(t, 64, seed=7)
            sample = f"cfg3 subsample A={A_s}, 64 candidate paths, oracle port, {cores} threads"
            for i in range(args.warmup + args.steps):
                t0 = time.perf_counter()
                orc.evaluate_paths(po, ps, w.offsets, w.steps, True, cores)
                sec = time.perf_counter() - t0
                if i >= args.warmup:
                    rates.append(64 * A_s / sec)
                    prates.append(64 / sec)
            ms = 1e3 * 64 * A_s / np.mean(rates)
    value = float(np.mean(rates))
    line = {"metric": METRIC, "value": value, "unit": UNIT, "impl": "reference",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic",
            "config": (arm_config(A_s, mbar, args.paths, args.gpus, "peer-memory sum kernel (NVLink P2P loads)")
                       if binary else {"workload": REF_WORKLOAD, "sample": sample, "extrapolated": True,
                                       "alignments_sample": A_s, "alignments_named_config": 1_000_000}),
            "sample": sample, "extrapolated": not binary,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": sample},
            "search_paths_per_s": {"sample": float(np.mean(prates)), "alignments_sample": A_s,
                                   "at_1M_alignments": float(np.mean(prates)) * A_s / 1e6},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def cpu_baseline_block(workload=None):
    """cpu_baseline of the default run: reference CLI on a bounded sample (kind=reference), with
    the oracle port's 1-thread and all-core rates on the same sample AND on the GPU arm's own CSR
    and paths (`same_inputs_port`) alongside."""
    from oracle import binding
    from gfalign_b200 import synth
    A_s, m_steps = REF_SAMPLE_A, REF_SAMPLE_M
    out = {}
    cores = os.cpu_count() or 1
    with tempfile.TemporaryDirectory() as wd:
        w, gfa, gaf, nl = reference_sample(wd, A_s)
        t = w.tangle
        binary = binding.ref_binary()
        orc = binding.Oracle()
        po, ps = synth.candidate_paths(t, 32, seed=7)
        t0 = time.perf_counter()
        orc.evaluate_paths(po, ps, w.offsets, w.steps, True, 1)
        port1 = 32 * A_s / (time.perf_counter() - t0)
        t0 = time.perf_counter()
        orc.evaluate_paths(po, ps, w.offsets, w.steps, True, cores)
        portn = 32 * A_s / (time.perf_counter() - t0)
        if binary:
            # the reference itself on the arm's own table (1 M alignments, the table of GPU rank 0): ~25 s of CPU
            with tempfile.TemporaryDirectory() as wd2:
                A_f = 1_000_000
                wf, gfa2, gaf2, nl2 = reference_sample(wd2, A_f)
                sec, calls, load = time_reference_cli(binary, gfa2, gaf2, nl2, wf.tangle, REF_FULL_M)
            out = {"value": calls * A_f / sec, "unit": UNIT, "cores": 1, "kind": "reference",
                   "sample": f"the arm's own cfg3 table (A={A_f}), unmodified reference CLI "
                             f"`search -j 1 -m {REF_FULL_M}`: {calls} evaluatePath calls in {sec:.2f} s "
                             f"(wall minus {load:.2f} s start-up + parsing); hot path is single-threaded",
                   "search_paths_per_s": {"sample": calls / sec, "alignments_sample": A_f,
                                          "at_1M_alignments": calls / sec * A_f / 1e6}}
        else:
            out = {"value": portn, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"cfg3 subsample A={A_s}, 32 candidate paths, oracle port"}
        out["port_1core"] = port1
        out["port_allcores"] = {"value": portn, "cores": cores}
    if workload is not None:
        off, steps, po, pst, k = workload
        sel_o = np.zeros(k + 1, dtype=np.uint32)
        sel_o[1:] = np.cumsum(np.diff(po[:k + 1].astype(np.int64)))
        sel_s = pst[:int(po[k])]
        t0 = time.perf_counter()
        orc.evaluate_paths(sel_o, sel_s, off, steps, True, cores)
        dt = time.perf_counter() - t0
        out["same_inputs_port"] = {"value": k * (off.size - 1) / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                   "sample": f"oracle port, all cores, the GPU arm's own CSR (A={off.size - 1}) and "
                                             f"its first {k} candidate paths ({dt:.2f} s)"}
    return out


# ---- helpers of the GPU arm -----------------------------------------------------------------------

def parity_check(orc, eng, off, steps, po, pst, pick, summed_rows, dist, dev, threads, exchange_name):
    """k candidate paths against the CPU oracle over EVERY rank's shard: each rank runs the oracle on
    its own shard (counts are sums over alignments, src/eval.cpp:80-99, so the oracle over the union is
    the sum of the per-shard oracles), the oracle counts are summed over the ranks and compared on
    every rank with `summed_rows` (what the exchange delivered for those paths)."""
    import torch
    k = len(pick)
    so = np.zeros(k + 1, dtype=np.uint32)
    ss = []
    for i, p in enumerate(pick):
        ss.append(pst[po[p]:po[p + 1]])
        so[i + 1] = so[i] + ss[-1].size
    t0 = time.perf_counter()
    want = orc.evaluate_paths(so, np.concatenate(ss), off, steps, True, threads).astype(np.int64)
    sec = time.perf_counter() - t0
    tw = torch.from_numpy(want).to(dev)
    if dist is not None:
        dist.all_reduce(tw)
    want = tw.cpu().numpy()
    ok = bool(np.array_equal(want, summed_rows.astype(np.int64)))
    t_ok = torch.tensor([1 if ok else 0], device=dev)
    if dist is not None:
        dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
    return {"paths": k, "path_ids": [int(p) for p in pick], "ok": bool(t_ok.item()),
            "checked_against": "CPU oracle (oracle/liboracle.so) over every rank's shard, summed",
            "exchange": exchange_name, "oracle_seconds": sec, "oracle_threads": threads,
            "bad_good_of_first": want[0].tolist()}


NCU_METRICS = ("dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,"
               "smsp__issue_active.avg.pct_of_peak_sustained_elapsed,"
               "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed,"
               "sm__warps_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum")


def ncu_counters(args):
    """Counters of ONE launch of the dominant kernel from a live `ncu --metrics` child run of this very
    script (`--counters-child`: same workload, a few steps, nothing timed).  Durations printed under ncu
    are not used.  -> (dict, source string) or (None, reason)."""
    ncu = shutil.which("ncu") or ("/usr/local/cuda/bin/ncu" if os.path.exists("/usr/local/cuda/bin/ncu") else None)
    if not ncu or os.environ.get("GFE_BENCH_NO_NCU"):
        return None, "ncu not available"
    cmd = [ncu, "--metrics", NCU_METRICS, "--clock-control", "none", "-k", "regex:score_bitset", "-s", "3",
           "-c", "1", "--csv", sys.executable, os.path.abspath(__file__), "--counters-child", "--paths",
           str(args.paths)] + (["--alignments", str(args.alignments)] if args.alignments else [])
    try:
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT,
                           env=dict(os.environ, GFE_BENCH_NO_NCU="1"))
    except Exception as ex:   # noqa: BLE001
        return None, f"ncu child failed: {ex!r}"[:200]
    rows = [r for r in csv.reader(io.StringIO(p.stdout)) if len(r) > 3]
    if not rows or "Metric Name" not in rows[0]:
        hdr = [i for i, r in enumerate(rows) if "Metric Name" in r]
        if not hdr:
            return None, ("ncu produced no metric table: " + (p.stderr or p.stdout)[-200:].replace("\n", " "))
        rows = rows[hdr[0]:]
    h = rows[0]
    ni, ui, vi = h.index("Metric Name"), h.index("Metric Unit"), h.index("Metric Value")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "inst": 1.0, "%": 1.0, "": 1.0,
             "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0, "warp": 1.0}
    out = {}
    for r in rows[1:]:
        try:
            out[r[ni]] = float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
        except ValueError:
            pass
    if "smsp__inst_executed.sum" not in out:
        return None, "ncu output lacks smsp__inst_executed.sum"
    return out, "ncu --metrics child run of this bench (live, this box)"


def counters_child(args):
    """The workload of the timed region, a few steps, under ncu (see ncu_counters)."""
    import torch
    import gfalign_b200
    from gfalign_b200 import synth
    t = synth.make_tangle(200, 200, seed=1, minus_frac=0.05, count2_frac=0.05)
    A = args.alignments or 1_000_000
    off, steps = synth.random_walk_alignments(t.n_nodes, t.edge_from, t.edge_to, A, *synth.HIFI_LENGTHS, seed=11)
    nodes, _, _ = synth.node_table(t)
    po, pst = synth.candidate_paths(t, args.paths, seed=5)
    eng = gfalign_b200.Engine(0)
    eng.set_dedup(False)
    eng.set_alignments(off, steps, t.n_nodes)
    eng.set_universe(nodes)
    dev = torch.device("cuda", 0)
    d_po = torch.from_numpy(po.astype(np.int64)).to(dev).to(torch.int32)
    d_ps = torch.from_numpy(pst.astype(np.int64)).to(dev).to(torch.int32)
    d_counts = torch.zeros(args.paths * 2, dtype=torch.int32, device=dev)
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)
    for _ in range(5):
        flush.zero_()
        eng.score_paths_device(d_po.data_ptr(), d_ps.data_ptr(), args.paths, int(pst.size), d_counts.data_ptr(),
                               filter=True, stream=0)
    torch.cuda.synchronize()
    eng.close()


CFG5_PIECES = 8
# rows of `search -m 20000` on the default cfg5 table (50 M alignments, 8 pieces), as emitted at N = 1 with every
# candidate scored by full scans AND by the incremental path (they agreed); runs at N > 1 must reproduce it
CFG5_SEARCH_SHA1 = {(50_000_000, 20000): "211e21f36920c6cd"}      # {(A_total, max_steps): sha1[:16]}


def _row_key(r):
    return (int(r.path_counter), int(r.bad), int(r.good), int(r.alt), int(r.uniques), int(r.hamiltonian),
            tuple(int(v) for v in r.steps))


def _rows_sha1(rows):
    import hashlib
    return hashlib.sha1(repr([_row_key(r) for r in rows]).encode()).hexdigest()[:16]


def search_cfg5(args, rank, world, dist, dev, eng, t, nodes, counts, node_count, fixed_table):
    """BASELINE configs[4] as the reference would run it: dijkstra() end to end (reference src/eval.cpp:110-193,
    one evaluatePath per expansion :162) on the 400-node tangle with the 50 M alignments split over the ranks
    — gfe_search at N = 1, gfe_search_sharded (incremental records summed by the peer kernel) at N > 1; -m is
    the same at every N, so seconds_search_loop over N is strong scaling."""
    import torch
    from gfalign_b200.sharded import PeerExchange
    m_steps = args.cfg5_search_steps
    cx = px_s = None
    if dist is not None:
        px_s = PeerExchange(1 << 18, dev)     # must outlive the handle: it owns the peer-mapped buffer
        cx = px_s.c_exchange()
        dist.barrier()
    torch.cuda.synchronize(dev)
    kw = dict(max_steps=m_steps)
    s0 = time.perf_counter()
    if cx is not None:
        rows, st = eng.search_sharded(cx, t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination,
                                      nodes, counts, node_count, **kw)
    else:
        rows, st = eng.search(t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination,
                              nodes, counts, node_count, **kw)
    wall = time.perf_counter() - s0
    sha = _rows_sha1(rows)
    tm = torch.tensor([st["seconds_total"], st["seconds_scoring"], wall], dtype=torch.float64, device=dev)
    agree = True
    if dist is not None:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        sig = torch.tensor([int(sha, 16) >> 2], dtype=torch.int64, device=dev)
        lo, hi = sig.clone(), sig.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        agree = bool(torch.equal(lo, hi))
    loop_s = float(tm[0].item())
    out = {"workload": "gfalign search on the cfg5 table (sharded over the ranks), source/destination of the tangle",
           "max_steps": m_steps, "n_gpus": world, "alignments_total": args.cfg5_alignments,
           "evaluatePath_calls": st["evaluations"], "paths_per_s": st["evaluations"] / loop_s,
           "seconds_search_loop": loop_s, "seconds_scoring": float(tm[1].item()),
           "seconds_incl_universe_and_index": float(tm[2].item()),
           "kernel_batches": st["kernel_batches"], "paths_scored": st["paths_scored"],
           "speculation_hit_rate": st["evaluations"] / max(st["paths_scored"], 1), "rows": len(rows),
           "rows_sha1": sha, "rows_identical_on_all_ranks": agree}
    if not agree:
        raise SystemExit("cfg5 search rows differ between ranks")
    if world == 1 and not args.no_cfg5_fullscan:
        # the checker of the incremental path at this size: every candidate through a full scan
        os.environ["GFE_SEARCH_FULLSCAN"] = "1"
        try:
            f0 = time.perf_counter()
            rows_fs, st_fs = eng.search(t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination,
                                        nodes, counts, node_count, **kw)
            out["fullscan_driver_seconds"] = time.perf_counter() - f0
        finally:
            del os.environ["GFE_SEARCH_FULLSCAN"]
        out["rows_equal_fullscan_driver"] = ([_row_key(r) for r in rows] == [_row_key(r) for r in rows_fs]
                                             and st_fs["evaluations"] == st["evaluations"])
        if not out["rows_equal_fullscan_driver"]:
            raise SystemExit("cfg5: incremental search rows differ from the full-scan driver's")
    del cx, px_s
    want = CFG5_SEARCH_SHA1.get((args.cfg5_alignments, m_steps)) if fixed_table else None
    out["rows_sha1_expected_from_n1"] = want
    out["rows_equal_n1"] = None if want is None else (want == sha)
    if want is not None and want != sha:
        raise SystemExit(f"cfg5 search rows differ from the N = 1 rows: {sha} != {want}")
    return out


def strong_cfg5(args, rank, world, local, dist, dev, orc, flush):
    """BASELINE configs[4]: 400-node dense ONT tangle, 50 M alignments split over the N ranks (N = 1: the
    whole table on one GPU), P candidate paths per step, exchange inside the timed step."""
    import torch
    import gfalign_b200
    from gfalign_b200 import synth
    from gfalign_b200.sharded import PeerExchange
    A_total = args.cfg5_alignments
    A = A_total // world
    P = args.cfg5_paths
    t = synth.make_tangle(400, 800, seed=3, minus_frac=0.05, count2_frac=0.10)
    nodes, counts, node_count = synth.node_table(t)
    t0 = time.perf_counter()
    # ONE table at every N: it is drawn as CFG5_PIECES pieces (seed 1300 + piece) and rank r of N holds the
    # pieces [r * PIECES / N, (r + 1) * PIECES / N), so the union over the ranks — and with it every summed
    # count and every search row — is the same for N = 1, 2, 4, 8
    fixed_table = CFG5_PIECES % world == 0 and A_total % CFG5_PIECES == 0
    if fixed_table:
        d_off, d_st = synth.table_shard_device(t.n_nodes, t.edge_from, t.edge_to, A_total, *synth.ONT_LENGTHS,
                                               seed=1300, device=dev, rank=rank, world=world, pieces=CFG5_PIECES)
    else:
        d_off, d_st = synth.random_walk_alignments_device(t.n_nodes, t.edge_from, t.edge_to, A, *synth.ONT_LENGTHS,
                                                          seed=1300 + rank, device=dev)
    torch.cuda.synchronize(dev)
    gen_s = time.perf_counter() - t0
    S = int(d_off[-1].item())
    po, pst = synth.candidate_paths(t, P, seed=5)
    eng = gfalign_b200.Engine(local)
    eng.set_dedup(False)          # brute force, like the headline; the compressed view is timed separately below
    eng.set_alignments_device(d_off.data_ptr(), d_st.data_ptr(), A, S, t.n_nodes)
    eng.set_universe(nodes)
    d_po = torch.from_numpy(po.astype(np.int64)).to(dev).to(torch.int32)
    d_ps = torch.from_numpy(pst.astype(np.int64)).to(dev).to(torch.int32)
    d_counts = torch.zeros(P * 2, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    px = None
    if dist is not None:
        try:
            px = PeerExchange(P * 2, dev)
        except Exception as ex:   # noqa: BLE001
            sys.stderr.write(f"[bench cfg5] peer exchange unavailable ({ex!r}); using NCCL\n")

    def step():
        if px is not None:
            eng.score_paths_device(d_po.data_ptr(), d_ps.data_ptr(), P, int(pst.size), px.half_ptr(),
                                   filter=True, stream=stream)
            px.sum_into(d_counts.data_ptr(), P * 2, stream)
            return
        eng.score_paths_device(d_po.data_ptr(), d_ps.data_ptr(), P, int(pst.size), d_counts.data_ptr(),
                               filter=True, stream=stream)
        if dist is not None:
            dist.all_reduce(d_counts)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    steps_k = max(3, min(args.steps, 5))
    for _ in range(3):
        flush.zero_()
        step()
    barrier()
    assert eng.device_status() == 0, "cfg5: device flags raised (worklist overflow / bad path)"
    eng.set_timing(True)
    eng.kernel_time_stats()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps_k)]
    c0 = eng.counters()
    barrier()
    for k in range(steps_k):
        flush.zero_()
        ev[k][0].record()
        step()
        ev[k][1].record()
    barrier()
    k1_ms, k1_n = eng.kernel_time_stats()
    eng.set_timing(False)
    c1 = eng.counters()
    assert eng.device_status() == 0, "cfg5: device flags raised in the timed region"
    total_ms = float(sum(a.elapsed_time(b) for a, b in ev))
    tm = torch.tensor([total_ms, k1_ms / max(k1_n, 1)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    total_ms, k1_mean = float(tm[0].item()), float(tm[1].item())
    # exchange alone (peer kernel incl. the wait for the slowest rank, nothing to wait for here)
    ex_ms = None
    if px is not None:
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            px.sum_into(d_counts.data_ptr(), P * 2, stream)
        e1.record()
        torch.cuda.synchronize(dev)
        ex_ms = e0.elapsed_time(e1) / 20
        step()          # leave d_counts holding real sums again
        barrier()
    # oracle spot check over all shards
    off_h = d_off.cpu().numpy().view(np.uint32)
    st_h = d_st.cpu().numpy().view(np.uint32)
    pick = np.linspace(0, P - 1, 3).astype(int)
    got = d_counts.cpu().numpy().view(np.uint32).reshape(P, 2)[pick]
    par = parity_check(orc, eng, off_h, st_h, po, pst, pick, got, dist, dev,
                       max(1, (os.cpu_count() or 1) // world),
                       "peer-memory sum kernel" if px is not None else ("nccl all-reduce" if dist is not None else "none"))
    ms_step = total_ms / steps_k
    out = {"workload": "cfg5 synthetic dense ONT tangle: 400 nodes, 50M alignments split over the GPUs "
                       "(table drawn on the device in 8 pieces, seed 1300 + piece: the same union at N = 1/2/4/8)"
                       if fixed_table else "cfg5 synthetic dense ONT tangle (table drawn on the device, seed 1300 + rank)",
           "same_table_at_every_n": fixed_table,
           "A_total": A * world, "alignments_per_gpu": A, "mean_steps": S / A, "candidate_paths_per_step": P,
           "n_gpus": world, "steps": steps_k, "ms_per_step": ms_step,
           "checks_per_s": float(P) * A * world / (ms_step * 1e-3), "k1_ms": k1_mean,
           "worklist_and_launch_ms": ms_step - k1_mean - (ex_ms or 0.0), "exchange_ms": ex_ms,
           "slow_pairs_per_step": None, "exact_pairs_per_step": int(c1["exact_pairs"] - c0["exact_pairs"]) // steps_k,
           "synth_seconds": gen_s, "parity_check": par, "scaling": "strong",
           "l2": "flushed (512 MiB write) before every timed step"}
    # the same step through the multiplicity-compressed view (product default), a separate figure
    brute = d_counts.clone()
    eng.set_dedup(True)
    eng.set_universe(nodes)
    ds = eng.dedup_stats()
    for _ in range(2):
        step()
    barrier()
    evd = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps_k)]
    for k in range(steps_k):
        flush.zero_()
        evd[k][0].record()
        step()
        evd[k][1].record()
    barrier()
    dd = torch.tensor([float(sum(a.elapsed_time(b) for a, b in evd)) / steps_k], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(dd, op=dist.ReduceOp.MAX)
    same = bool(torch.equal(brute, d_counts))
    out["dedup"] = {"note": "separate figure: equal walks scored once (exact); ms_per_step above is brute force",
                    "in_use": ds["in_use"], "rows": ds["rows"], "view_rows": ds["view_rows"],
                    "row_reduction": ds["rows"] / max(ds["view_rows"], 1), "ms_per_step": float(dd.item()),
                    "counts_equal_brute_force": same}
    if not args.no_search:
        out["search"] = search_cfg5(args, rank, world, dist, dev, eng, t, nodes, counts, node_count, fixed_table)
    eng.close()
    del d_off, d_st
    if not par["ok"]:
        raise SystemExit("cfg5 parity check failed: " + json.dumps(par))
    if not same:
        raise SystemExit("cfg5: dedup view counts differ from the brute-force counts")
    return out


# ---- the GPU arm -------------------------------------------------------------------------------

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--paths", type=int, default=1024, help="candidate paths per step")
    ap.add_argument("--alignments", type=int, default=0, help="override A per GPU (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-search", action="store_true")
    ap.add_argument("--search-steps", type=int, default=100000)
    ap.add_argument("--no-extras", action="store_true", help="skip the cfg4 block (N = 1)")
    ap.add_argument("--extras-alignments", type=int, default=20_000_000)
    ap.add_argument("--no-cfg5", action="store_true", help="skip the cfg5 strong-scaling block")
    ap.add_argument("--cfg5-alignments", type=int, default=50_000_000)
    ap.add_argument("--cfg5-paths", type=int, default=1024)
    ap.add_argument("--cfg5-search-steps", type=int, default=20000, help="-m of the cfg5 search block (same at every N)")
    ap.add_argument("--no-cfg5-fullscan", action="store_true", help="skip the N = 1 full-scan check of the cfg5 search")
    ap.add_argument("--no-ncu", action="store_true")
    ap.add_argument("--counters-child", action="store_true", help=argparse.SUPPRESS)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        run_reference_arm(args)
        return
    if args.counters_child:
        counters_child(args)
        return

    import torch
    import gfalign_b200
    from gfalign_b200 import synth
    from oracle.binding import Oracle   # the checker of parity_check / cpu_baseline only

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: gfalign_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        # stdout carries exactly one JSON line: NCCL's printf banner goes to stderr (fd 1 -> fd 2 for
        # the C level, Python keeps a private copy of the real stdout)
        sys.stdout.flush()
        _real = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        sys.stdout = _real
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    orc = Oracle()
    cores = os.cpu_count() or 1

    # workload (headline): every rank builds the same tangle and its own alignment shard (seed + rank)
    t = synth.make_tangle(200, 200, seed=1, minus_frac=0.05, count2_frac=0.05)
    A = args.alignments or 1_000_000
    desc = CFG3_DESC
    off, steps = synth.random_walk_alignments(t.n_nodes, t.edge_from, t.edge_to, A, *synth.HIFI_LENGTHS,
                                              seed=11 + rank)
    nodes, counts, node_count = synth.node_table(t)
    po, pst = synth.candidate_paths(t, args.paths, seed=5)
    P = args.paths
    mbar = float(off[-1]) / A
    bytes_per_check = 4.0 * (mbar + 1.0)

    eng_sms = torch.cuda.get_device_properties(local).multi_processor_count
    eng = gfalign_b200.Engine(local)
    eng.set_dedup(False)          # headline = brute force: every alignment row scored for every path
    eng.set_alignments(off, steps, t.n_nodes)
    eng.set_universe(nodes)

    dev = torch.device("cuda", local)
    d_po = torch.from_numpy(po.astype(np.int64)).to(dev).to(torch.int32)
    d_ps = torch.from_numpy(pst.astype(np.int64)).to(dev).to(torch.int32)
    d_counts = torch.zeros(P * 2, dtype=torch.int32, device=dev)
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)   # 4x L2
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream

    # the exchange step of the sharded path: one peer-memory kernel per rank (csrc/peer.cu) reading
    # every shard's record over NVLink; GFE_EXCHANGE=nccl selects a NCCL all-reduce instead
    px = px_e2e = None
    exchange = "none"
    if dist is not None:
        exchange = "nccl all-reduce"
        if os.environ.get("GFE_EXCHANGE", "peer") == "peer":
            try:
                from gfalign_b200.sharded import PeerExchange
                px = PeerExchange(P * 2, dev)
                px_e2e = PeerExchange(1 << 18, dev)     # also carries the search's incremental batches
                exchange = "peer-memory sum kernel (NVLink P2P loads)"
            except Exception as ex:   # no peer access between the GPUs of this box
                sys.stderr.write(f"[bench] peer exchange unavailable ({ex!r}); using NCCL\n")
                px = px_e2e = None

    def step(use_px=True):
        if px is not None and use_px:
            eng.score_paths_device(d_po.data_ptr(), d_ps.data_ptr(), P, int(pst.size),
                                   px.half_ptr(), filter=True, stream=stream)
            px.sum_into(d_counts.data_ptr(), P * 2, stream)
            return
        eng.score_paths_device(d_po.data_ptr(), d_ps.data_ptr(), P, int(pst.size),
                               d_counts.data_ptr(), filter=True, stream=stream)
        if dist is not None:
            dist.all_reduce(d_counts)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        flush.zero_()
        step()
    barrier()
    ref_counts = d_counts.clone()

    # ---- parity: the configuration about to be timed, against the CPU oracle, on every rank's shard ----
    pick = np.linspace(0, P - 1, 4).astype(int)
    got = ref_counts.cpu().numpy().view(np.uint32).reshape(P, 2)
    parity = parity_check(orc, eng, off, steps, po, pst, pick, got[pick], dist, dev, max(1, cores // world),
                          exchange)
    if dist is not None:
        step(use_px=False)                       # the same batch through NCCL
        barrier()
        same = bool(torch.equal(ref_counts, d_counts))
        parity["nccl_exchange_equals_peer_exchange"] = same
        parity["ok"] = parity["ok"] and same
        step()
        barrier()
    if not parity["ok"]:
        raise SystemExit("parity check failed: " + json.dumps(parity))

    # ---- timed region: K steps, device-timed, L2 flushed between steps -------------------------
    sampler = ClockSampler(local)
    sampler.start()
    c0 = eng.counters()
    eng.set_timing(True)
    eng.kernel_time_stats()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    barrier()
    w0 = time.perf_counter()
    for k in range(args.steps):
        flush.zero_()
        ev[k][0].record()
        step()
        ev[k][1].record()
    barrier()
    wall = time.perf_counter() - w0
    kern_ms, kern_n = eng.kernel_time_stats()
    eng.set_timing(False)
    c1 = eng.counters()
    sampler.stop_flag.set()
    sampler.join()
    clocks = sampler.summary()
    step_ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = float(sum(step_ms))
    t_max = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_max, op=dist.ReduceOp.MAX)
    total_ms = float(t_max.item())
    assert eng.device_status() == 0, "device flags raised (worklist overflow / bad path)"
    assert torch.equal(ref_counts, d_counts), "counts changed between steps"
    checks_per_step = float(P) * A * world
    value = checks_per_step * args.steps / (total_ms * 1e-3)

    # ---- roofline of the dominant kernel -----------------------------------------------------------
    peak, peak_src = measured_peak_gbs()
    k1_ms = kern_ms / max(kern_n, 1)
    alg_bytes = float(P) * A * bytes_per_check
    legacy = os.environ.get("GFE_FORCE_LEGACY") == "1"
    ctr, ctr_src = (None, "skipped")
    if rank == 0 and world == 1 and not args.no_ncu and not legacy:
        ctr, ctr_src = ncu_counters(args)
    if ctr is None:
        try:
            with open(os.path.join(ROOT, FALLBACK_CAPTURE)) as f:
                cap = json.load(f)
            if cap.get("paths") == P and cap.get("alignments") == A and not legacy:
                ctr, ctr_src = cap["counters"], f"{FALLBACK_CAPTURE} (commit {cap.get('commit', '?')}); live ncu: {ctr_src}"
        except Exception:
            pass
    sm_hz = (clocks["sm_mhz"] or 1965.0) * 1e6
    issue_peak = eng_sms * 4 * sm_hz                      # 4 schedulers per SM, one warp instruction per clock each
    roofline = {"kernel": "score_paths_kernel" if legacy else "score_bitset_kernel", "kernel_ms": k1_ms,
                "kernel_launches_timed": kern_n, "counters_source": ctr_src}
    hbm = {"algorithmic_bytes_per_launch": alg_bytes, "algorithmic_gbs": alg_bytes / (k1_ms * 1e-3) / 1e9,
           "algorithmic_frac": alg_bytes / (k1_ms * 1e-3) / 1e9 / peak, "peak": peak, "peak_source": peak_src,
           "note": "algorithmic bytes = P x A x 4(m+1) (SURVEY.md 8d: one row read per check); one pass over the "
                   "rows serves all P paths from shared-memory bitsets (P-way reuse), so the algorithmic figure "
                   "exceeds the HBM peak and is not a bound; frac_actual is the DRAM traffic the kernel really moves"}
    if ctr is not None:
        inst = float(ctr["smsp__inst_executed.sum"])
        traffic = float(ctr.get("dram__bytes_read.sum", 0.0)) + float(ctr.get("dram__bytes_write.sum", 0.0))
        roofline.update({"bound": "issue (integer ALU)", "achieved": inst / (k1_ms * 1e-3), "peak": issue_peak,
                         "unit": "warp-inst/s", "frac": inst / (k1_ms * 1e-3) / issue_peak, "traffic": traffic,
                         "warp_instructions_per_launch": inst, "warp_instructions_per_row": inst / A / max(1, (P + 1023) // 1024),
                         "issue_active_pct": ctr.get("smsp__issue_active.avg.pct_of_peak_sustained_elapsed"),
                         "alu_pipe_active_pct": ctr.get("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed"),
                         "warps_active_pct": ctr.get("sm__warps_active.avg.pct_of_peak_sustained_active"),
                         "peak_definition": f"{eng_sms} SMs x 4 schedulers x {sm_hz / 1e6:.0f} MHz (clock sampled "
                                            "during the timed region)"})
        hbm["traffic"] = traffic
        hbm["frac_actual"] = traffic / (k1_ms * 1e-3) / 1e9 / peak
        # one pass over the length-bucketed 16-bit view (row_stride(m) u16 per row, kernels.cuh)
        m_all = np.diff(off.astype(np.int64))
        stride = np.where(m_all <= 0, 0, (2 * m_all - 1 + np.where(m_all > 2, (m_all - 1) // 2, 0) + 3) & ~3)
        view_bytes = float((stride * 2).sum())
        hbm["one_table_pass_bytes"] = view_bytes
        hbm["traffic_over_one_table_pass"] = traffic / view_bytes * max(1, 1) / max(1, (P + 1023) // 1024)
    else:
        roofline.update({"bound": "issue (integer ALU)", "achieved": None, "peak": issue_peak, "unit": "warp-inst/s",
                         "frac": None, "traffic": None})
    roofline["hbm"] = hbm
    roofline["checks_per_sm_clock"] = float(P) * A / (k1_ms * 1e-3) / (eng_sms * sm_hz)

    # single-path launches: the streaming (HBM/L2-bound) regime of the non-speculative driver
    d_c1 = torch.zeros(2, dtype=torch.int32, device=dev)
    one_off = d_po[:2].contiguous()
    n0 = int(po[1])
    eng.set_timing(True)
    for _ in range(3):
        flush.zero_()
        eng.score_paths_device(one_off.data_ptr(), d_ps.data_ptr(), 1, n0, d_c1.data_ptr(),
                               filter=True, stream=stream)
    eng.kernel_time_stats()
    for _ in range(10):
        flush.zero_()
        eng.score_paths_device(one_off.data_ptr(), d_ps.data_ptr(), 1, n0, d_c1.data_ptr(),
                               filter=True, stream=stream)
    ms1, n1 = eng.kernel_time_stats()
    eng.set_timing(False)
    p1_ms = ms1 / max(n1, 1)
    roofline["single_path_launch"] = {"bound": "hbm", "kernel_ms": p1_ms,
                                      "achieved": A * bytes_per_check / (p1_ms * 1e-3) / 1e9, "peak": peak,
                                      "unit": "GB/s", "frac": A * bytes_per_check / (p1_ms * 1e-3) / 1e9 / peak,
                                      "note": "P=1 (no reuse: here 4(m+1) B per check IS the traffic), L2 flushed "
                                              "before every launch"}

    # ---- e2e: the host C-ABI call with host buffers (H2D paths, D2H counts every step) ----------
    cx = px_e2e.c_exchange() if px_e2e is not None else None

    def e2e_call():
        if cx is not None:
            return eng.score_paths_sharded(cx, po, pst, filter=True)       # ONE C call: H2D, kernels, peer sum, D2H
        g = eng.score_paths(po, pst, filter=True)
        if dist is not None:
            tcount = torch.from_numpy(g.astype(np.int64)).to(dev)
            dist.all_reduce(tcount)
            g = tcount.cpu().numpy()
        return g

    for _ in range(3):
        e2e_call()
    c_e0 = eng.counters()
    barrier()
    e0 = time.perf_counter()
    for _ in range(args.steps):
        got_e = e2e_call()
    barrier()
    e_sec = time.perf_counter() - e0
    c_e1 = eng.counters()
    e_t = torch.tensor([e_sec], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(e_t, op=dist.ReduceOp.MAX)
    e_sec = float(e_t.item())
    e2e = {"value": checks_per_step * args.steps / e_sec, "unit": UNIT,
           "h2d_bytes_per_step": (c_e1["h2d_bytes"] - c_e0["h2d_bytes"]) // args.steps,
           "d2h_bytes_per_step": (c_e1["d2h_bytes"] - c_e0["d2h_bytes"]) // args.steps,
           "call": ("gfe_score_paths_sharded(host path CSR) -> summed host counts: kernels accumulate into the "
                    "exchange buffer, peer kernel sums the shards, one D2H; alignment shards resident"
                    if cx is not None else
                    "gfe_score_paths(host path CSR) -> host counts; alignment table resident"
                    + (" + NCCL all-reduce of the counts" if dist is not None else "")),
           "slow_pairs_per_step": int(c_e1["slow_pairs"] - c_e0["slow_pairs"]) // args.steps,
           "exact_pairs_per_step": int(c_e1["exact_pairs"] - c_e0["exact_pairs"]) // args.steps}
    assert np.array_equal(np.asarray(got_e).reshape(-1).astype(np.int64),
                          d_counts.cpu().numpy().astype(np.int64)), "host and device API disagree"
    parity["e2e_call_equals_device_path"] = True

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic",
            "config": arm_config(A, mbar, P, world, exchange),
            "clocks": clocks, "e2e": e2e, "roofline": roofline, "parity_check": parity,
            "gpu_launches": int(c1["kernel_launches"] - c0["kernel_launches"]) + (args.steps if px is not None else 0),
            "wall_ms_timed_region": wall * 1e3,
            "exact_dp_pairs_per_step": int(c1["exact_pairs"] - c0["exact_pairs"]) // args.steps}

    # ---- the same step through the multiplicity-compressed view (the product default) ----------------------
    if world == 1:
        eng.set_dedup(True)
        eng.set_universe(nodes)
        ds = eng.dedup_stats()
        for _ in range(3):
            flush.zero_()
            step()
        torch.cuda.synchronize()
        same = bool(torch.equal(ref_counts, d_counts))
        eng.set_timing(True)
        eng.kernel_time_stats()
        evd = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for k in range(args.steps):
            flush.zero_()
            evd[k][0].record()
            step()
            evd[k][1].record()
        torch.cuda.synchronize()
        dk_ms, dk_n = eng.kernel_time_stats()
        eng.set_timing(False)
        d_ms = float(sum(a.elapsed_time(b) for a, b in evd)) / args.steps
        line["dedup"] = {
            "note": "separate line, NOT the headline: alignments with equal walks (up to reverse complement) are scored "
                    "once and counted with their multiplicity (exact: the verdict depends on the walk only); the factor "
                    "is a property of the table — random walks of 3-8 steps on a 200-node graph repeat a lot",
            "rows": ds["rows"], "distinct_walks": ds["distinct_walks"], "view_rows": ds["view_rows"],
            "in_use": ds["in_use"], "row_reduction": ds["rows"] / max(ds["view_rows"], 1),
            "ms_per_step": d_ms, "kernel_ms": dk_ms / max(dk_n, 1),
            "effective_checks_per_s": float(P) * A / (d_ms * 1e-3),
            "brute_force_checks_per_s": value, "counts_equal_brute_force": same}
        if not same:
            raise SystemExit("dedup view counts differ from the brute-force counts")
        eng.set_dedup(False)
        eng.set_universe(nodes)

    # ---- the `gfalign search` driver itself: search paths/s -------------------------------------------
    if not args.no_search:
        m_steps = args.search_steps if world == 1 else min(args.search_steps, 20000)
        with gfalign_b200.Engine(local) as e2:
            barrier()
            s0 = time.perf_counter()
            e2.set_alignments(off, steps, t.n_nodes)
            up_s = time.perf_counter() - s0
            if cx is not None:
                rows, st = e2.search_sharded(cx, t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source,
                                             t.destination, nodes, counts, node_count, max_steps=m_steps)
            else:
                rows, st = e2.search(t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source,
                                     t.destination, nodes, counts, node_count, max_steps=m_steps)
            s_sec = time.perf_counter() - s0
        ts = torch.tensor([s_sec], dtype=torch.float64, device=dev)
        sig = torch.tensor([len(rows), sum(r.bad * 31 + r.good for r in rows) % (1 << 31)], dtype=torch.int64, device=dev)
        rows_agree = True
        if dist is not None:
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
            lo, hi = sig.clone(), sig.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            rows_agree = bool(torch.equal(lo, hi))
        s_sec = float(ts.item())
        srch = {"max_steps": m_steps, "seconds": s_sec, "seconds_table_upload": up_s,
                "seconds_search_loop": st["seconds_total"], "evaluatePath_calls": st["evaluations"],
                "paths_per_s": st["evaluations"] / st["seconds_total"],
                "paths_per_s_incl_upload": st["evaluations"] / s_sec,
                "effective_checks_per_s": st["evaluations"] * float(A) * world / st["seconds_total"],
                "kernel_batches": st["kernel_batches"], "paths_scored": st["paths_scored"],
                "speculation_hit_rate": st["evaluations"] / max(st["paths_scored"], 1),
                "rows": len(rows), "seconds_scoring": st["seconds_scoring"],
                "alignments_total": A * world, "rows_identical_on_all_ranks": rows_agree,
                "note": "search paths/s = evaluatePath calls the reference's dijkstra() would make (heap replay is "
                        "exact) / seconds of the search loop; reported separately from brute-force checks/s"}
        if not rows_agree:
            raise SystemExit("search rows differ between ranks")

        row_key = _row_key
        srch["rows_sha1"] = _rows_sha1(rows)
        if world == 1:
            # the same search with EVERY candidate scored by a full scan (gfe_score_paths, the kernel the oracle
            # check above covers): the incremental path must emit the same rows in the same order
            # (emission order is the contract: reference src/eval.cpp:176-183)
            os.environ["GFE_SEARCH_FULLSCAN"] = "1"
            try:
                with gfalign_b200.Engine(local) as e4:
                    e4.set_alignments(off, steps, t.n_nodes)
                    f0 = time.perf_counter()
                    rows_fs, st_fs = e4.search(t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source,
                                               t.destination, nodes, counts, node_count, max_steps=m_steps)
                    fs_sec = time.perf_counter() - f0
            finally:
                del os.environ["GFE_SEARCH_FULLSCAN"]
            same_rows = [row_key(r) for r in rows] == [row_key(r) for r in rows_fs]
            srch["rows_equal_fullscan_driver"] = same_rows
            srch["fullscan_driver_seconds"] = fs_sec
            srch["fullscan_evaluatePath_calls"] = st_fs["evaluations"]
            if not same_rows or st_fs["evaluations"] != st["evaluations"]:
                raise SystemExit("incremental search rows differ from the full-scan driver's")
        line["search"] = srch
        if cx is not None:
            # strong scaling of the N = 1 search (the cfg3 table of rank 0, replicated): candidate-batch sharding
            off0, steps0 = (off, steps) if rank == 0 else synth.random_walk_alignments(
                t.n_nodes, t.edge_from, t.edge_to, A, *synth.HIFI_LENGTHS, seed=11)
            with gfalign_b200.Engine(local) as e3:
                e3.set_alignments(off0, steps0, t.n_nodes)
                cx.set_batch_sharding(True)
                barrier()
                rows3, st3 = e3.search_sharded(cx, t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source,
                                               t.destination, nodes, counts, node_count, max_steps=args.search_steps)
                cx.set_batch_sharding(False)
            ts3 = torch.tensor([st3["seconds_total"]], dtype=torch.float64, device=dev)
            sig3 = torch.tensor([len(rows3), sum(r.bad * 31 + r.good for r in rows3) % (1 << 31)], dtype=torch.int64, device=dev)
            lo, hi = sig3.clone(), sig3.clone()
            dist.all_reduce(ts3, op=dist.ReduceOp.MAX)
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            if not torch.equal(lo, hi):
                raise SystemExit("batch-sharded search rows differ between ranks")
            line["search_strong_cfg3"] = {
                "sharding": "candidate batches (speculation subtrees dealt over the ranks), table replicated",
                "max_steps": args.search_steps, "alignments": A, "seconds_search_loop": float(ts3.item()),
                "evaluatePath_calls": st3["evaluations"], "paths_per_s": st3["evaluations"] / float(ts3.item()),
                "kernel_batches": st3["kernel_batches"], "paths_scored": st3["paths_scored"], "rows": len(rows3),
                "seconds_scoring": st3["seconds_scoring"], "rows_identical_on_all_ranks": True,
                "note": "same table and search as the N = 1 line's `search` block"}

    # ---- BASELINE configs[4]: cfg5, 50 M alignments split over the ranks -------------------------------
    eng.close()
    del flush
    torch.cuda.empty_cache()
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)
    if not args.no_cfg5:
        line["strong_cfg5"] = strong_cfg5(args, rank, world, local, dist, dev, orc, flush)
    # ---- BASELINE configs[3]: cfg4 kernels (N = 1) ---------------------------------------------------------
    if rank == 0 and world == 1 and not args.no_extras:
        import bench_extras
        line["extras"] = bench_extras.run_cfg4(local, args.extras_alignments, 5, not args.no_cpu_baseline, flush=flush)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline_block((off, steps, po, pst, 4))
        line["cpu_baseline"] = cb
        # BASELINE.md publishes no number for this metric: vs_baseline stays null (contract); the ratio to the
        # same-run reference CLI sample is given beside it, labelled
        line["vs_reference_sample"] = {"value": value / cb["value"], "note": "value / cpu_baseline.value; the reference "
                                       "ran `search -m 4` on the same 1 M-alignment table (see cpu_baseline.sample)"}
        if "search" in line and "search_paths_per_s" in cb:
            line["search"]["reference_paths_per_s"] = cb["search_paths_per_s"]
    if rank == 0:
        # the numbers a reader of the last ~1 KB of the line needs (records that keep only a tail of it)
        tail = {}
        if "search" in line:
            sr = line["search"]
            tail["search"] = {k: sr[k] for k in ("max_steps", "alignments_total", "evaluatePath_calls", "paths_per_s",
                                                 "seconds_search_loop", "seconds_scoring", "kernel_batches",
                                                 "speculation_hit_rate", "rows", "rows_sha1",
                                                 "rows_identical_on_all_ranks", "rows_equal_fullscan_driver")
                              if k in sr}
        if isinstance(line.get("strong_cfg5"), dict):
            c5 = line["strong_cfg5"]
            tail["strong_cfg5"] = {k: c5[k] for k in ("A_total", "n_gpus", "ms_per_step", "k1_ms", "exchange_ms",
                                                      "worklist_and_launch_ms", "checks_per_s") if k in c5}
            tail["strong_cfg5"]["parity_ok"] = (c5.get("parity_check") or {}).get("ok")
            if isinstance(c5.get("search"), dict):
                tail["search_cfg5"] = {k: c5["search"][k] for k in (
                    "max_steps", "n_gpus", "alignments_total", "evaluatePath_calls", "paths_per_s", "seconds_search_loop",
                    "seconds_scoring", "rows", "rows_sha1", "rows_equal_n1", "rows_equal_fullscan_driver",
                    "rows_identical_on_all_ranks") if k in c5["search"]}
        tail["headline"] = {"value": line.get("value"), "e2e": (line.get("e2e") or {}).get("value"),
                            "ms_per_step": line.get("ms_per_step"), "n_gpus": line.get("n_gpus"),
                            "parity_ok": (line.get("parity_check") or {}).get("ok")}
        line["tail_summary"] = tail
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
