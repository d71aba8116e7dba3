"""Search driver on cfg3 (1 M alignments resident): seconds of `gfe_search -m 100000` vs the speculation
parameters of the incremental path (width = candidates per batch, roots = speculation roots taken from
the heap's best frontier paths), with the full-scan driver (GFE_SEARCH_FULLSCAN=1) as the row checker.
usage: python profiles/search_sweep.py [max_steps]"""
import json
import os
import sys
import time

sys.path.insert(0, ".")
from gfalign_b200 import capi, synth     # noqa: E402

max_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
w = synth.make_workload("cfg3", A=1_000_000)
t = w.tangle
nodes, counts, node_count = synth.node_table(t)
eng = capi.Engine(0)
eng.set_alignments(w.offsets, w.steps, t.n_nodes)
args = (t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination, nodes, counts, node_count)


def run(label, **env):
    for k, v in env.items():
        os.environ[k] = str(v)
    eng.search(*args, max_steps=200)      # warm-up (index build, buffers)
    t0 = time.perf_counter()
    rows, st = eng.search(*args, max_steps=max_steps)
    dt = time.perf_counter() - t0
    for k in env:
        del os.environ[k]
    key = [r.columns(t.names) for r in rows]
    print(json.dumps({"label": label, **env, "max_steps": max_steps, "seconds": round(dt, 4),
                      "loop_seconds": round(st["seconds_total"], 4), "seconds_scoring": round(st["seconds_scoring"], 4),
                      "rows": len(rows), "batches": st["kernel_batches"], "paths_scored": st["paths_scored"],
                      "calls": st["evaluations"], "paths_per_s": round(st["evaluations"] / st["seconds_total"]),
                      "hit_rate": round(st["evaluations"] / max(st["paths_scored"], 1), 4)}), flush=True)
    return key


ref = run("full-scan driver (round 1 policy)", GFE_SEARCH_FULLSCAN=1)
for width in (256, 1024, 2048, 4096):
    for roots in (1, 4, 8, 16, 32):
        key = run("incremental", GFE_SPEC_WIDTH=width, GFE_SPEC_ROOTS=roots)
        if key != ref:
            print(json.dumps({"ERROR": "rows differ from the full-scan driver", "width": width, "roots": roots}))
eng.close()
