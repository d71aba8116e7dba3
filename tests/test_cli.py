"""The drop-in CLI `gfalign_b200/bin/gfalign-b200` against the stdout (and output files) of the
unmodified reference CLI captured in tests/golden/cli.json.gz.  CPU part: the host-only paths
(alignment summary of evalGFA, argument / input errors, the loud no-GPU failure).  GPU part: every
golden case of search / evalPath / evalGFA / filter, byte for byte."""
import os
import subprocess

import pytest

from test_gpu_parity import _cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "gfalign_b200", "bin", "gfalign-b200")


def _run(case, tmp, drop_out=False):
    for fn, txt in case["files"].items():
        with open(os.path.join(tmp, fn), "w") as f:
            f.write(txt)
    args = list(case["args"])
    if drop_out and "-o" in args:
        i = args.index("-o")
        del args[i:i + 2]
    p = subprocess.run([CLI] + args, cwd=tmp, capture_output=True, text=True, timeout=300)
    return p


def test_cli_built():
    assert os.access(CLI, os.X_OK), "run __graft_entry__.build()"


def test_evalgfa_alignment_summary_host_only(tmp_path):
    """evalGFA without -o prints only the alignment summary (sort, markDuplicates, printStats:
    reference src/input-gfalign.cpp:84-90) — host code, no GPU needed."""
    seen = 0
    for c in _cases():
        if not c["name"].endswith("_evalgfa"):
            continue
        p = _run(c, str(tmp_path), drop_out=True)
        assert p.returncode == 0, p.stderr
        assert p.stdout == c["stdout"], c["name"]
        seen += 1
    assert seen >= 7


def test_argument_and_input_errors(tmp_path):
    c = next(x for x in _cases() if x["name"] == "random3_evalpath_0")
    tmp = str(tmp_path)
    _run(c, tmp)   # writes the files
    r = subprocess.run([CLI, "frobnicate"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "mode 'frobnicate' does not exist. Terminating." in r.stderr
    r = subprocess.run([CLI, "evalPath", "-f", "nope.gfa"], cwd=tmp, capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and r.stdout == "Error - file does not exist: nope.gfa\n"
    r = subprocess.run([CLI, "evalPath", "-f", "random3.gfa", "-g", "random3.gaf", "-p", "1+,9+"],
                       cwd=tmp, capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and r.stderr == "Error: cannot find node (9). Terminating.\n"
    with open(os.path.join(tmp, "bad.tsv"), "w") as f:
        f.write("2\t1\nnot-a-node\t1\n")
    r = subprocess.run([CLI, "search", "-f", "random3.gfa", "-g", "random3.gaf", "-n", "bad.tsv", "-s", "1",
                        "-d", "4"], cwd=tmp, capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and r.stderr == "Error: node not in graph (pIUd: not-a-node)\n"


def test_no_gpu_fails_loudly(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    c = next(x for x in _cases() if x["name"] == "random3_search")
    p = _run(c, str(tmp_path))
    assert p.returncode != 0 and "no CPU fallback" in p.stderr and p.stdout == ""


@pytest.mark.gpu
def test_cli_matches_reference_stdout(tmp_path):
    counts = {}
    for c in _cases():
        d = os.path.join(str(tmp_path), c["name"])
        os.makedirs(d)
        p = _run(c, d)
        assert p.returncode == c["rc"], (c["name"], p.stderr)
        assert p.stdout == c["stdout"], c["name"]
        tool = c["args"][0]
        counts[tool] = counts.get(tool, 0) + 1
        if "-o" in c["args"]:
            out = open(os.path.join(d, c["args"][c["args"].index("-o") + 1])).read()
            if tool == "filter":
                assert out == c["out_text"], c["name"]
            else:
                assert [l for l in out.split("\n") if l.startswith("L\t")] == c["out_L_lines"], c["name"]
                if c["name"].startswith(("tangle", "hairpin", "random2")):
                    assert out == c["out_text"], c["name"]     # H / S / L / J / P records: whole file
    assert counts == {"search": 13, "evalPath": 18, "evalGFA": 7, "filter": 7}


@pytest.mark.gpu
def test_cli_search_options(tmp_path):
    """--speculate (batch size) never changes the rows; -m 0 prints only the step-limit line."""
    c = next(x for x in _cases() if x["name"] == "tangle1_search_all")
    base = None
    for spec in ("1", "64", "4096"):
        d = os.path.join(str(tmp_path), "s" + spec)
        os.makedirs(d)
        cc = dict(c)
        cc["args"] = c["args"] + ["--speculate", spec]
        p = _run(cc, d)
        assert p.returncode == 0 and p.stdout == c["stdout"]
        base = p.stdout
    assert base
