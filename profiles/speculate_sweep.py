"""Search driver: wall time of `gfe_search` (cfg3, 1 M alignments resident) vs the speculation width
(paths scored per batch).  Rows are identical for every width (the heap replay defines them)."""
import json
import sys
import time

sys.path.insert(0, ".")
import numpy as np                      # noqa: E402
from gfalign_b200 import capi, synth     # noqa: E402

w = synth.make_workload("cfg3", A=1_000_000)
t = w.tangle
nodes, counts, node_count = synth.node_table(t)
eng = capi.Engine(0)
eng.set_alignments(w.offsets, w.steps, t.n_nodes)
ref = None
for max_steps in (20000, 100000):
    for spec in (128, 256, 512, 1024, 2048, 4096):
        args = (t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination, nodes, counts, node_count)
        eng.search(*args, max_steps=200, speculate=spec)      # warm-up
        t0 = time.perf_counter()
        rows, st = eng.search(*args, max_steps=max_steps, speculate=spec)
        dt = time.perf_counter() - t0
        key = [r.columns(t.names) for r in rows]
        if spec == 128:
            ref = key
        print(json.dumps({"max_steps": max_steps, "speculate": spec, "seconds": round(dt, 4), "rows": len(rows),
                          "same_rows": key == ref, "batches": st.get("kernel_batches"),
                          "paths_scored": st.get("paths_scored"), "calls": st.get("evaluations"), "seconds_scoring": st.get("seconds_scoring")}), flush=True)
eng.close()
