"""CPU suite: the C++ text packers (gfe_text_*: GFA1 / GAF -> packed arrays, the host side of the
drop-in CLI) against the Python restatement in tests/textio.py — which the golden search /
evalPath / evalGFA tests pin to the reference — on every golden case, on odd walks, and on a GAF
large enough to be split across parser threads."""
import os

import numpy as np
import pytest

from gfalign_b200 import capi, synth
import textio
from test_gpu_parity import _cases


def _write(tmp_path, files):
    out = {}
    for fn, txt in files.items():
        p = os.path.join(tmp_path, fn)
        with open(p, "w") as f:
            f.write(txt)
        out[fn] = p
    return out


def _check_against_textio(gfa_text, gaf_text, tp):
    names, ef, et = textio.parse_gfa(gfa_text)
    assert tp.names == names
    assert np.array_equal(tp.edge_from, ef) and np.array_equal(tp.edge_to, et)
    ao, at, o0, o1 = synth.build_adjacency(len(names), ef, et)
    assert np.array_equal(tp.adj_offsets, ao) and np.array_equal(tp.adj_target, at)
    assert np.array_equal(tp.adj_or0, o0) and np.array_equal(tp.adj_or1, o1)
    if gaf_text is not None:
        qn, off, st = textio.parse_gaf(gaf_text, names)
        assert tp.A == len(qn)
        assert np.array_equal(tp.offsets, off) and np.array_equal(tp.steps, st)
        assert tp.qnames == qn


def test_golden_cases_pack_like_textio(tmp_path):
    seen = set()
    for c in _cases():
        gfa = next((k for k in c["files"] if k.endswith(".gfa")), None)
        if gfa is None:      # filter mode takes no graph
            continue
        gaf = next((k for k in c["files"] if k.endswith(".gaf")), None)
        key = (c["files"][gfa], c["files"].get(gaf))
        if key in seen:
            continue
        seen.add(key)
        paths = _write(str(tmp_path), c["files"])
        tp = capi.TextPack(paths[gfa], paths[gaf] if gaf else None, threads=3)
        _check_against_textio(c["files"][gfa], c["files"].get(gaf), tp)
    assert len(seen) >= 7


def test_odd_walks_and_records(tmp_path):
    gfa = "S\ta\t*\nS\tb\t*\nL\ta\t+\tc\t-\t0M\nL\tc\t-\tb\t+\t0M\tXX:i:3\nL\ta\t+\tc\t-\t5M\n"
    rec = "\t9\t0\t9\t+\t{}\t9\t0\t9\t9\t9\t60\n"
    walks = [">a<c>b", "<b>c<a", ">zz>a", "*", "", "><a", "abc>b", ">a"]
    gaf = "".join(f"q{i}" + rec.format(w) for i, w in enumerate(walks)) + "\n"   # trailing empty line
    paths = _write(str(tmp_path), {"g.gfa": gfa, "a.gaf": gaf})
    tp = capi.TextPack(paths["g.gfa"], paths["a.gaf"], threads=1)
    assert tp.names == ["a", "b", "c"]          # S records first, then names first seen on L
    _check_against_textio(gfa, gaf, tp)
    # unknown names alias to node 0, a piece's first character is its orientation mark
    assert tp.steps[tp.offsets[2]:tp.offsets[3]].tolist() == [0, 0]
    assert tp.steps[tp.offsets[3]:tp.offsets[4]].tolist() == [1]
    assert tp.offsets[5] - tp.offsets[4] == 0


def test_errors(tmp_path):
    paths = _write(str(tmp_path), {"g2.gfa": "H\tVN:Z:2.0\nS\ta\t4\tACGT\n", "g.gfa": "S\ta\t*\n",
                                   "bad.gaf": "q\t1\t2\n"})
    with pytest.raises(capi.GfeError):
        capi.TextPack(paths["g2.gfa"])
    with pytest.raises(capi.GfeError):
        capi.TextPack(os.path.join(str(tmp_path), "missing.gfa"))
    with pytest.raises(capi.GfeError):
        capi.TextPack(paths["g.gfa"], paths["bad.gaf"])


def test_threaded_parse_is_order_preserving(tmp_path):
    t = synth.make_tangle(300, 400, seed=5, minus_frac=0.1, count2_frac=0.0, prefix="utig4-")
    off, st = synth.random_walk_alignments(t.n_nodes, t.edge_from, t.edge_to, 120_000,
                                           [1, 3, 5, 9, 40], [.1, .4, .3, .15, .05], seed=6)
    gfa_p, gaf_p = os.path.join(str(tmp_path), "g.gfa"), os.path.join(str(tmp_path), "a.gaf")
    synth.write_gfa(gfa_p, t)
    synth.write_gaf(gaf_p, t.names, off, st)
    assert os.path.getsize(gaf_p) > 4 << 20     # several 1 MiB parser chunks
    one = capi.TextPack(gfa_p, gaf_p, threads=1)
    many = capi.TextPack(gfa_p, gaf_p, threads=7)
    for tp in (one, many):
        assert tp.A == 120_000
        assert np.array_equal(tp.offsets, off) and np.array_equal(tp.steps, st)
    assert one.qnames[:5] == [f"read{i}" for i in range(5)] and one.qnames == many.qnames
