#!/usr/bin/env python
"""bench.py — headline benchmark of the gfalign hot path on B200 (contract: see DESIGN.md §Measurement).

One "step" = one pass of the hot path over one batch: P candidate paths (the kind `gfalign search`
expands) scored against every alignment of the resident table with the id filter on
(evaluatePath x P, reference src/eval.cpp:67-108).  metric = path x alignment checks per second,
whole job.  N > 1 (torchrun): every rank holds its own alignment shard of the same shape (weak
scaling), scores the same batch, and the P x {bad, good} counts are combined with one NCCL
all-reduce per step inside the timed region.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3|cfg5] [--paths P]
  python bench.py --impl reference ...   # the reference's own CPU implementation, bounded sample
"""
from __future__ import annotations

import argparse
import json
import os
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

def reference_sample(workdir: str, A: int = 20000):
    """Text files of a bounded cfg3 sample for the reference CLI."""
    from gfalign_b200 import synth
    w = synth.make_workload("cfg3", A=A)
    t = w.tangle
    gfa, gaf, nl = (os.path.join(workdir, x) for x in ("s.gfa", "s.gaf", "s.tsv"))
    synth.write_gfa(gfa, t)
    synth.write_gaf(gaf, t.names, w.offsets, w.steps)
    synth.write_nodelist(nl, t)
    return w, gfa, gaf, nl


def time_reference_cli(binary: str, gfa: str, gaf: str, nl: str, t, m_steps: int):
    """Hot-path seconds of `gfalign search -m m_steps` = wall(-m m_steps) - wall(-m 0); the number
    of evaluatePath calls is read from the --verbose log (src/eval.cpp:145)."""
    base = [binary, "search", "-j", "1", "-f", gfa, "-g", gaf, "-n", nl, "-s", t.names[t.source],
            "-d", t.names[t.destination]]
    t0 = time.perf_counter()
    subprocess.run(base + ["-m", "0"], capture_output=True, check=True, timeout=600)
    load = time.perf_counter() - t0
    t0 = time.perf_counter()
    subprocess.run(base + ["-m", str(m_steps)], capture_output=True, check=True, timeout=600)
    full = time.perf_counter() - t0
    v = subprocess.run(base + ["-m", str(m_steps), "--verbose"], capture_output=True, text=True, timeout=600)
    calls = v.stderr.count("We can visit this node")
    return max(full - load, 1e-9), calls, load


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
    A_s, m_steps = 20000, 60
    rates = []
    with tempfile.TemporaryDirectory() as wd:
        w, gfa, gaf, nl = reference_sample(wd, A_s)
        t = w.tangle
        if binary:
            kind, cores = "reference", 1
            sample = (f"cfg3 subsample: 200-node tangle, A={A_s} alignments (2% of 1M), "
                      f"`gfalign search -j 1 -m {m_steps}` hot-path time = wall - wall(-m 0)")
            for i in range(args.warmup + args.steps):
                sec, calls, _ = time_reference_cli(binary, gfa, gaf, nl, t, m_steps)
                if i >= args.warmup:
                    rates.append(calls * A_s / sec)
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
            ms = 1e3 * 64 * A_s / np.mean(rates)
    value = float(np.mean(rates))
    line = {"metric": METRIC, "value": value, "unit": UNIT, "impl": "reference",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic",
            "config": {"workload": "cfg3 HiFi tangle (200 nodes, 1M alignments) — bounded sample",
                       "sample": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def cpu_baseline_block():
    """cpu_baseline of the default run: reference CLI on a bounded sample (kind=reference), with
    the oracle port's 1-thread and all-core rates on the same sample alongside."""
    from oracle import binding
    from gfalign_b200 import synth
    A_s, m_steps = 20000, 60
    out = {}
    with tempfile.TemporaryDirectory() as wd:
        w, gfa, gaf, nl = reference_sample(wd, A_s)
        t = w.tangle
        binary = binding.ref_binary()
        orc = binding.Oracle()
        po, ps = synth.candidate_paths(t, 32, seed=7)
        t0 = time.perf_counter()
        orc.evaluate_paths(po, ps, w.offsets, w.steps, True, 1)
        port1 = 32 * A_s / (time.perf_counter() - t0)
        cores = os.cpu_count() or 1
        t0 = time.perf_counter()
        orc.evaluate_paths(po, ps, w.offsets, w.steps, True, cores)
        portn = 32 * A_s / (time.perf_counter() - t0)
        if binary:
            sec, calls, load = time_reference_cli(binary, gfa, gaf, nl, t, m_steps)
            out = {"value": calls * A_s / sec, "unit": UNIT, "cores": 1, "kind": "reference",
                   "sample": f"cfg3 subsample A={A_s} (2% of 1M), unmodified reference CLI "
                             f"`search -j 1 -m {m_steps}`: {calls} evaluatePath calls in {sec:.2f} s "
                             f"(wall minus {load:.2f} s load); hot path is single-threaded"}
        else:
            out = {"value": portn, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"cfg3 subsample A={A_s}, 32 candidate paths, oracle port"}
        out["port_1core"] = port1
        out["port_allcores"] = {"value": portn, "cores": cores}
    return out


# ---- the GPU arm -------------------------------------------------------------------------------

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=["cfg3", "cfg5"])
    ap.add_argument("--paths", type=int, default=1024, help="candidate paths per step")
    ap.add_argument("--alignments", type=int, default=0, help="override A per GPU (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-search", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch
    import gfalign_b200
    from gfalign_b200 import synth

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

    # workload: every rank builds the same tangle and its own alignment shard (seed + rank)
    if args.workload == "cfg3":
        t = synth.make_tangle(200, 200, seed=1, minus_frac=0.05, count2_frac=0.05)
        A = args.alignments or 1_000_000
        ls, ps_ = synth.HIFI_LENGTHS
        desc = "cfg3 synthetic HiFi tangle: 200-node utig4 nodelist, 1M q60 alignments per GPU"
    else:
        t = synth.make_tangle(400, 800, seed=3, minus_frac=0.05, count2_frac=0.10)
        A = args.alignments or (50_000_000 // world)
        ls, ps_ = synth.ONT_LENGTHS
        desc = "cfg5 synthetic dense ONT tangle: 400 nodes, 50M alignments split over the GPUs"
    off, steps = synth.random_walk_alignments(t.n_nodes, t.edge_from, t.edge_to, A, ls, ps_,
                                              seed=11 + rank)
    nodes, counts, node_count = synth.node_table(t)
    po, pst = synth.candidate_paths(t, args.paths, seed=5)
    P = args.paths
    mbar = float(off[-1]) / A
    bytes_per_check = 4.0 * (mbar + 1.0)

    eng_sms = torch.cuda.get_device_properties(local).multi_processor_count
    eng = gfalign_b200.Engine(local)
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
    px = None
    exchange = "none"
    if dist is not None:
        exchange = "nccl all-reduce"
        if os.environ.get("GFE_EXCHANGE", "peer") == "peer":
            try:
                from gfalign_b200.sharded import PeerExchange
                px = PeerExchange(P * 2, dev)
                exchange = "peer-memory sum kernel (NVLink P2P loads)"
            except Exception as ex:   # no peer access between the GPUs of this box
                sys.stderr.write(f"[bench] peer exchange unavailable ({ex!r}); using NCCL\n")

    def step():
        if px is not None:
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

    # ---- roofline of the dominant kernel (score_paths) -----------------------------------------
    peak, peak_src = measured_peak_gbs()
    k1_ms = kern_ms / max(kern_n, 1)
    alg_bytes = float(P) * A * bytes_per_check
    achieved = alg_bytes / (k1_ms * 1e-3) / 1e9
    legacy = os.environ.get("GFE_FORCE_LEGACY") == "1"
    # dram__bytes_read+write of ONE launch from the committed ncu --set full capture of this command
    # (profiles/r01_score_bitset_ncu_full_summary.csv; cfg3, P = 1024): 20.41 MB read
    traffic = 20.41e6 if (args.workload == "cfg3" and P == 1024 and A == 1_000_000 and not legacy) else None
    # the limiter of the bitset kernel is the SM (integer ALU pipe 70 % busy, issue slots 61 %), not HBM:
    # warp instructions per launch from the same capture (1.1e8) vs 4 issue slots/clk/SM
    sm_clock = 1.965e9
    roofline = {"bound": "hbm", "kernel": "score_paths_kernel" if legacy else "score_bitset_kernel",
                "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": peak_src, "kernel_ms": k1_ms, "kernel_launches_timed": kern_n,
                "algorithmic_bytes_per_launch": alg_bytes,
                "note": "algorithmic bytes = P x A x 4(m+1) (SURVEY.md 8d); one pass over the table "
                        "serves every path of a chunk from shared-memory bitsets, so DRAM traffic "
                        "is ~1/1000 of it and frac > 1; the kernel is integer-ALU/issue bound "
                        "(DESIGN.md 4)",
                "issue_bound": {"checks_per_sm_clock": float(P) * A / (k1_ms * 1e-3) / (eng_sms * sm_clock),
                                "sms": eng_sms, "sm_clock_hz": sm_clock}}

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
    roofline["single_path_launch"] = {"kernel_ms": p1_ms,
                                      "achieved": A * bytes_per_check / (p1_ms * 1e-3) / 1e9,
                                      "frac": A * bytes_per_check / (p1_ms * 1e-3) / 1e9 / peak,
                                      "note": "P=1, L2 flushed before every launch"}

    # ---- e2e: the host C-ABI call with host buffers (H2D paths, D2H counts every step) ----------
    for _ in range(3):
        eng.score_paths(po, pst, filter=True)
    c_e0 = eng.counters()
    barrier()
    e0 = time.perf_counter()
    for _ in range(args.steps):
        got = eng.score_paths(po, pst, filter=True)
        if dist is not None:
            tcount = torch.from_numpy(got.astype(np.int64)).to(dev)
            dist.all_reduce(tcount)
            got = tcount.cpu().numpy()
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
           "call": "gfe_score_paths(host path CSR) -> host counts; alignment table resident",
           "slow_pairs_per_step": int(c_e1["slow_pairs"] - c_e0["slow_pairs"]) // args.steps,
           "exact_pairs_per_step": int(c_e1["exact_pairs"] - c_e0["exact_pairs"]) // args.steps}
    assert np.array_equal(got.reshape(-1).astype(np.int64),
                          d_counts.cpu().numpy().astype(np.int64)), "host and device API disagree"

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic",
            "config": {"workload": desc, "alignments_per_gpu": A, "mean_steps": mbar,
                       "candidate_paths_per_step": P, "filter": True,
                       "l2": "flushed (512 MiB write) before every timed step",
                       "parallelism": f"alignment shards x{world}, counts summed by {exchange}"
                       if world > 1 else "single GPU"},
            "clocks": sampler.summary(), "e2e": e2e, "roofline": roofline,
            "gpu_launches": int(c1["kernel_launches"] - c0["kernel_launches"]) + (args.steps if px is not None else 0),
            "wall_ms_timed_region": wall * 1e3,
            "exact_dp_pairs_per_step": int(c1["exact_pairs"] - c0["exact_pairs"]) // args.steps}

    if rank == 0 and not args.no_search and world == 1:
        # the `gfalign search` driver itself, table upload included: search paths/s
        with gfalign_b200.Engine(local) as e2:
            s0 = time.perf_counter()
            e2.set_alignments(off, steps, t.n_nodes)
            rows, st = e2.search(t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source,
                                 t.destination, nodes, counts, node_count, max_steps=2000)
            s_sec = time.perf_counter() - s0
        line["search"] = {"max_steps": 2000, "seconds": s_sec, "evaluatePath_calls": st["evaluations"],
                          "paths_per_s": st["evaluations"] / s_sec,
                          "checks_per_s": st["evaluations"] * A / s_sec,
                          "kernel_batches": st["kernel_batches"], "paths_scored": st["paths_scored"],
                          "rows": len(rows), "seconds_scoring": st["seconds_scoring"]}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_block()
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
