#!/usr/bin/env python
"""bench_cli.py — end to end through the drop-in CLI (text in, text out): `gfalign-b200` on the full
cfg3 text files next to the unmodified reference CLI (oracle/_ref/gfalign, when built) on a bounded
sample of them.  One JSON line per command.  Not the driver's contract (bench.py); run by hand on
the GPU box, results committed under profiles/.

  python bench_cli.py [--alignments 1000000] [--sample 20000]
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
CLI = os.path.join(ROOT, "gfalign_b200", "bin", "gfalign-b200")


def run(cmd, cwd, timeout=900):
    t0 = time.perf_counter()
    p = subprocess.run(cmd, cwd=cwd, capture_output=True, text=True, timeout=timeout)
    return time.perf_counter() - t0, p


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--alignments", type=int, default=1_000_000)
    ap.add_argument("--sample", type=int, default=20_000)
    args = ap.parse_args()
    from gfalign_b200 import synth
    from oracle import binding
    ref = binding.ref_binary()
    w = synth.make_workload("cfg3", A=args.alignments)
    t = w.tangle
    src, dst = t.names[t.source], t.names[t.destination]
    po, ps = synth.candidate_paths(t, 64, seed=3, max_len=150)
    k = int(np.argmax(np.diff(po)))
    path = ",".join(synth.step_str(t.names, int(s)) for s in ps[po[k]:po[k + 1]])
    with tempfile.TemporaryDirectory() as wd:
        synth.write_gfa(os.path.join(wd, "g.gfa"), t)
        synth.write_nodelist(os.path.join(wd, "n.tsv"), t)
        with open(os.path.join(wd, "keep.ls"), "w") as f:
            f.write("".join(nm + "\n" for nm in t.names[::2]))
        from bench_extras import _write_gaf_fast
        _write_gaf_fast(os.path.join(wd, "a.gaf"), t.names, w.offsets, w.steps)
        _write_gaf_fast(os.path.join(wd, "s.gaf"), t.names, w.offsets[:args.sample + 1], w.steps)
        gaf_mb = os.path.getsize(os.path.join(wd, "a.gaf")) / 1e6
        base = {"workload": "cfg3 text", "alignments": args.alignments, "gaf_mb": gaf_mb,
                "host_cores": os.cpu_count()}

        def emit(**kw):
            print(json.dumps({**base, **kw}), flush=True)

        run([CLI, "search", "-f", "g.gfa", "-g", "s.gaf", "-n", "n.tsv", "-s", src, "-d", dst, "-m", "10"], wd)
        for m in (2000, 20000, 100000):
            sec, p = run([CLI, "search", "-f", "g.gfa", "-g", "a.gaf", "-n", "n.tsv", "-s", src, "-d", dst,
                          "-m", str(m), "--timing"], wd)
            emit(tool="search", impl="gfalign-b200", max_steps=m, seconds=sec, rc=p.returncode,
                 rows=p.stdout.count("\n"), timing=p.stderr.strip().splitlines()[-1:] )
        sec, p = run([CLI, "evalGFA", "-f", "g.gfa", "-g", "a.gaf", "-o", "dec.gfa"], wd)
        emit(tool="evalGFA", impl="gfalign-b200", seconds=sec, rc=p.returncode)
        sec, p = run([CLI, "evalPath", "-f", "g.gfa", "-g", "a.gaf", "-p", path], wd)
        emit(tool="evalPath", impl="gfalign-b200", seconds=sec, rc=p.returncode, path_steps=path.count(",") + 1,
             stdout_mb=len(p.stdout) / 1e6, totals=p.stdout.strip().splitlines()[-1])
        sec, p = run([CLI, "filter", "-g", "a.gaf", "-n", "keep.ls", "--min-nodes", "2", "-o", "f.gaf"], wd)
        emit(tool="filter", impl="gfalign-b200", seconds=sec, rc=p.returncode,
             kept=sum(1 for _ in open(os.path.join(wd, "f.gaf"))))
        # the same tools on the bounded sample, both implementations (stdout compared)
        cmds = {
            "search": ["search", "-f", "g.gfa", "-g", "s.gaf", "-n", "n.tsv", "-s", src, "-d", dst, "-m", "60"],
            "evalGFA": ["evalGFA", "-f", "g.gfa", "-g", "s.gaf", "-o", "dec_s.gfa"],
            "evalPath": ["evalPath", "-f", "g.gfa", "-g", "s.gaf", "-p", path],
            "filter": ["filter", "-g", "s.gaf", "-n", "keep.ls", "--min-nodes", "2", "-o", "f_s.gaf"],
        }
        for tool, cmd in cmds.items():
            sec, p = run([CLI] + cmd, wd)
            rec = {"tool": tool, "impl": "gfalign-b200", "sample_alignments": args.sample, "seconds": sec,
                   "rc": p.returncode}
            if ref:
                try:
                    rsec, rp = run([ref] + cmd[:1] + ["-j", "1"] + cmd[1:], wd, timeout=600)
                    rec.update(reference_seconds=rsec, reference_rc=rp.returncode,
                               stdout_identical=(rp.stdout == p.stdout))
                except subprocess.TimeoutExpired:
                    rec.update(reference_seconds=None, reference_timeout=600)
            emit(**rec)


if __name__ == "__main__":
    main()
