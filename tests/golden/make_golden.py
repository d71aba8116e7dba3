#!/usr/bin/env python
"""Generates the golden fixtures of tests/golden/ from the UNMODIFIED reference, compiled by
oracle/Makefile into oracle/_ref/ (run in the build container, where /root/reference exists):

    python tests/golden/make_golden.py

  pairs.npz   per-pair traceback scores T_fw / T_rc of the reference's own alignPaths()
              (src/alignments.cpp:556-561) through oracle/_ref/libgfalign_ref.so
  cli.json.gz stdout (and the decorated GFA's L lines) of the reference CLI `gfalign
              search | evalPath | evalGFA -j 1` on the reference's testFiles and on small seeded
              synthetic graphs; the input texts ride along so the tests need no /root/reference
"""
import gzip
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from gfalign_b200 import synth  # noqa: E402
from oracle import binding  # noqa: E402

REF_TESTFILES = "/root/reference/testFiles"


def make_pairs(shim):
    rng = np.random.default_rng(20261104)
    A_off, A_st, B_off, B_st, tf, tr = [0], [], [0], [], [], []

    def add(a, b):
        f, r = shim.pair_scores(np.array(a, dtype=np.uint32), np.array(b, dtype=np.uint32))
        A_st.extend(a); A_off.append(len(A_st))
        B_st.extend(b); B_off.append(len(B_st))
        tf.append(f); tr.append(r)

    for alpha, n_hi, m_hi, count in [(2, 8, 9, 600), (3, 12, 8, 600), (6, 40, 17, 600),
                                     (12, 60, 10, 300), (4, 30, 45, 150)]:
        for _ in range(count):
            n = int(rng.integers(1, n_hi + 1))
            m = int(rng.integers(0, m_hi + 1))
            a = (rng.integers(0, alpha, n) * 2 + rng.integers(0, 2, n)).tolist()
            b = (rng.integers(0, alpha, m) * 2 + rng.integers(0, 2, m)).tolist()
            add(a, b)
    # planted: exact sub-walks, reverse complements, overhangs at either end of the path
    for _ in range(600):
        n = int(rng.integers(2, 30))
        a = (rng.permutation(40)[:n] * 2 + rng.integers(0, 2, n)).tolist()
        i = int(rng.integers(0, n)); j = int(rng.integers(i + 1, n + 1))
        b = a[i:j]
        kind = int(rng.integers(0, 5))
        if kind == 1:
            b = [x ^ 1 for x in reversed(b)]
        elif kind == 2:      # overhang before the path start
            b = [int(rng.integers(40, 50)) * 2 for _ in range(int(rng.integers(1, 4)))] + a[:j]
        elif kind == 3:      # overhang past the path end
            b = a[i:] + [int(rng.integers(40, 50)) * 2 for _ in range(int(rng.integers(1, 4)))]
        elif kind == 4:      # one substituted step
            b = list(b); b[int(rng.integers(0, len(b)))] ^= 1
        add(a, b)
    return dict(A_off=np.array(A_off, np.uint32), A_st=np.array(A_st, np.uint32),
                B_off=np.array(B_off, np.uint32), B_st=np.array(B_st, np.uint32),
                t_fw=np.array(tf, np.int32), t_rc=np.array(tr, np.int32))


def run_ref(binary, args, cwd):
    # always bounded: the reference can hang on its error paths (thread pool never joined)
    p = subprocess.run([binary] + args, cwd=cwd, capture_output=True, text=True, timeout=120)
    return p.stdout, p.returncode


def cli_case(binary, name, files, args, outfile=None):
    """Run the reference CLI in a scratch dir holding `files`; keep stdout (+ L lines of outfile)."""
    with tempfile.TemporaryDirectory() as wd:
        for fn, txt in files.items():
            with open(os.path.join(wd, fn), "w") as f:
                f.write(txt)
        out, rc = run_ref(binary, args, wd)
        case = {"name": name, "files": files, "args": args, "stdout": out, "rc": rc}
        if outfile:
            with open(os.path.join(wd, outfile)) as f:
                text = f.read()
            case["out_L_lines"] = [l for l in text.split("\n") if l.startswith("L\t")]
            case["out_text"] = text
        return case


def tangle_texts(t, off, st):
    with tempfile.TemporaryDirectory() as wd:
        synth.write_gfa(os.path.join(wd, "g.gfa"), t)
        synth.write_gaf(os.path.join(wd, "a.gaf"), t.names, off, st)
        synth.write_nodelist(os.path.join(wd, "n.tsv"), t)
        return {fn: open(os.path.join(wd, fn)).read() for fn in ("g.gfa", "a.gaf", "n.tsv")}


def make_cli(binary):
    cases = []
    rd = lambda fn: open(os.path.join(REF_TESTFILES, fn)).read()  # noqa: E731
    r3 = {"random3.gfa": rd("random3.gfa"), "random3.gaf": rd("random3.gaf"),
          "random3.search_nodelist.tsv": rd("random3.search_nodelist.tsv")}
    base = ["search", "-j", "1", "-f", "random3.gfa", "-n", "random3.search_nodelist.tsv"]
    cases.append(cli_case(binary, "random3_search_noaln", r3, base + ["-s", "1", "-d", "4"]))
    cases.append(cli_case(binary, "random3_search", r3, base + ["-g", "random3.gaf", "-s", "1", "-d", "4"]))
    cases.append(cli_case(binary, "random3_search_all", r3,
                          base + ["-g", "random3.gaf", "-s", "1", "-d", "4", "--return-all-paths"]))
    cases.append(cli_case(binary, "random3_search_minnodes", r3,
                          base + ["-g", "random3.gaf", "-s", "1", "-d", "4", "--min-nodes", "3"]))
    cases.append(cli_case(binary, "random3_search_m1", r3,
                          base + ["-g", "random3.gaf", "-s", "1", "-d", "4", "-m", "1"]))
    for i, p in enumerate(["1+,2+,3+,4+", "1+,4+", "4-,3-,2-,1-", "1+,3+,4+", "5-", "1+;2+,4+"]):
        cases.append(cli_case(binary, f"random3_evalpath_{i}", r3,
                              ["evalPath", "-j", "1", "-f", "random3.gfa", "-g", "random3.gaf", "-p", p]))
    r2 = {"random2.gfa": rd("random2.gfa"), "random2.gaf": rd("random2.gaf")}
    cases.append(cli_case(binary, "random2_evalgfa", r2,
                          ["evalGFA", "-j", "1", "-g", "random2.gaf", "-f", "random2.gfa", "-o", "dec.gfa"],
                          outfile="dec.gfa"))
    r1 = {"random1.gfa": rd("random1.gfa"), "random1.gaf": rd("random1.gaf")}
    cases.append(cli_case(binary, "random1_evalgfa", r1,
                          ["evalGFA", "-j", "1", "-g", "random1.gaf", "-f", "random1.gfa", "-o", "dec.gfa"],
                          outfile="dec.gfa"))
    # hairpin probe
    hp = {"h.gfa": "H\tVN:Z:1.2\nS\ta\tACGT\nS\tb\tACGT\nL\ta\t+\ta\t-\t0M\nL\ta\t+\tb\t+\t0M\n"
                   "L\ta\t+\ta\t+\t0M\nL\tb\t-\ta\t-\t0M\n",
          "h.gaf": "".join(f"r{i}\t20\t0\t20\t+\t>a<a>a>a>b\t20\t0\t20\t20\t20\t60\n" for i in range(3))
                   + "r3\t8\t0\t8\t+\t<b<a\t8\t0\t8\t8\t8\t60\n"}
    cases.append(cli_case(binary, "hairpin_evalgfa", hp,
                          ["evalGFA", "-j", "1", "-g", "h.gaf", "-f", "h.gfa", "-o", "dec.gfa"],
                          outfile="dec.gfa"))
    # filter mode (reference testFiles + a ':'-bearing tag, a numeric column with leading zeros)
    r3f = dict(r3)
    r3f["random3.filter_nodelist.ls"] = rd("random3.filter_nodelist.ls")
    fb = ["filter", "-j", "1", "-g", "random3.gaf", "-n", "random3.filter_nodelist.ls"]
    cases.append(cli_case(binary, "random3_filter", r3f, fb + ["-o", "out.gaf"], outfile="out.gaf"))
    cases.append(cli_case(binary, "random3_filter_min3", r3f, fb + ["--min-nodes", "3", "-o", "out.gaf"],
                          outfile="out.gaf"))
    odd = {"n.ls": "a\nb\nc\n",
           "a.gaf": "r1\t010\t0\t10\t+\t>a<b\t10\t0\t10\t10\t10\t60\ttp:A:P\tzz:Z:x:y\n"
                    "r2\t12\t1\t9\t-\t>a>d\t10\t0\t10\t9\t10\t0\n"
                    "r3\t7\t0\t7\t+\t<c\t7\t0\t7\t7\t7\t30\n"}
    cases.append(cli_case(binary, "odd_filter", odd, ["filter", "-j", "1", "-g", "a.gaf", "-n", "n.ls",
                                                      "-o", "out.gaf"], outfile="out.gaf"))
    # seeded synthetic tangles: search (driver order + scoring), evalPath (per-pair scores), evalGFA
    for k, (n_nodes, extra, A, c2, m_steps) in enumerate([(12, 14, 300, 0.3, 400), (30, 40, 1500, 0.1, 300),
                                                         (60, 70, 4000, 0.05, 150), (8, 16, 200, 0.5, 500)]):
        t = synth.make_tangle(n_nodes, extra, seed=100 + k, minus_frac=0.15, count2_frac=c2, prefix="u")
        off, st = synth.random_walk_alignments(t.n_nodes, t.edge_from, t.edge_to, A,
                                               [1, 2, 3, 4, 5, 6, 8, 20], [.05, .1, .3, .25, .15, .08, .05, .02],
                                               seed=200 + k)
        files = tangle_texts(t, off, st)
        sd = ["-s", t.names[t.source], "-d", t.names[t.destination]]
        b = ["search", "-j", "1", "-f", "g.gfa", "-g", "a.gaf", "-n", "n.tsv"] + sd + ["-m", str(m_steps)]
        cases.append(cli_case(binary, f"tangle{k}_search", files, b))
        cases.append(cli_case(binary, f"tangle{k}_search_all", files, b + ["--return-all-paths"]))
        po, ps = synth.candidate_paths(t, 3, seed=300 + k)
        for i in range(3):
            p = ",".join(synth.step_str(t.names, int(s)) for s in ps[po[i]:po[i + 1]])
            cases.append(cli_case(binary, f"tangle{k}_evalpath_{i}", files,
                                  ["evalPath", "-j", "1", "-f", "g.gfa", "-g", "a.gaf", "-p", p]))
        keep = "".join(nm + "\n" for nm in t.names[: max(3, (2 * n_nodes) // 3)])
        ff = dict(files)
        ff["keep.ls"] = keep
        cases.append(cli_case(binary, f"tangle{k}_filter", ff,
                              ["filter", "-j", "1", "-g", "a.gaf", "-n", "keep.ls", "--min-nodes", "2",
                               "-o", "out.gaf"], outfile="out.gaf"))
        cases.append(cli_case(binary, f"tangle{k}_evalgfa", files,
                              ["evalGFA", "-j", "1", "-g", "a.gaf", "-f", "g.gfa", "-o", "dec.gfa"],
                              outfile="dec.gfa"))
    return cases


def main():
    binding.build(ref=True)
    binary = binding.ref_binary()
    assert binary, "oracle/_ref/gfalign not built"
    np.savez_compressed(os.path.join(HERE, "pairs.npz"), **make_pairs(binding.RefShim()))
    with gzip.open(os.path.join(HERE, "cli.json.gz"), "wt") as f:
        json.dump({"generator": "tests/golden/make_golden.py", "reference": "vgl-hub/gfalign @ 187b1de",
                   "cases": make_cli(binary)}, f, indent=0)
    print("golden fixtures written")


if __name__ == "__main__":
    main()
