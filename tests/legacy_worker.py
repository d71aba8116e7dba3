"""Worker of tests/test_gpu_legacy.py: scores a cfg3 sample with GFE_FORCE_LEGACY=1 (set by the
parent) so that score_paths_kernel — the kernel behind rows > 16 steps and universes without a pair
dictionary — carries every verdict, and checks it against the CPU oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import gfalign_b200  # noqa: E402
from gfalign_b200 import synth  # noqa: E402
from oracle.binding import Oracle  # noqa: E402

assert os.environ.get("GFE_FORCE_LEGACY") == "1"
w = synth.make_workload("cfg3", A=40000)
t = w.tangle
nodes, _, _ = synth.node_table(t)
po, ps = synth.candidate_paths(t, 80, seed=17)
with gfalign_b200.Engine(0) as eng:
    eng.set_alignments(w.offsets, w.steps, t.n_nodes)
    eng.set_universe(nodes)
    eng.set_timing(True)
    for filt in (True, False):
        got = eng.score_paths(po, ps, filter=filt)
        want = Oracle().evaluate_paths(po, ps, w.offsets, w.steps, filt, os.cpu_count() or 1)
        assert np.array_equal(got, want), filt
    print("legacy ok", eng.last_kernel_ms())
