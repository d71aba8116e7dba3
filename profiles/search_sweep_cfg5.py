"""Search driver on the cfg5 table of bench.py (400-node tangle, 50 M alignments drawn on the device in 8 pieces,
one GPU): seconds of `gfe_search -m <max_steps>` vs the speculation parameters, the full-scan driver
(GFE_SEARCH_FULLSCAN=1) as the row checker.  usage: python profiles/search_sweep_cfg5.py [max_steps] [A_total] [one]"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, ".")
from gfalign_b200 import capi, synth     # noqa: E402

max_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
A_total = int(sys.argv[2]) if len(sys.argv) > 2 else 50_000_000
dev = torch.device("cuda", 0)
t = synth.make_tangle(400, 800, seed=3, minus_frac=0.05, count2_frac=0.10)
nodes, counts, node_count = synth.node_table(t)
d_off, d_st = synth.table_shard_device(t.n_nodes, t.edge_from, t.edge_to, A_total, *synth.ONT_LENGTHS, seed=1300,
                                       device=dev)
eng = capi.Engine(0)
eng.set_alignments_device(d_off.data_ptr(), d_st.data_ptr(), A_total, int(d_off[-1].item()), t.n_nodes)
args = (t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination, nodes, counts, node_count)


def run(label, **env):
    for k, v in env.items():
        os.environ[k] = str(v)
    eng.search(*args, max_steps=50)      # warm-up (universe, index build, buffers)
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


if len(sys.argv) > 3 and sys.argv[3] == "one":      # ncu captures: the default policy only
    run("incremental (default policy)")
    eng.close()
    sys.exit(0)
ref = run("full-scan driver", GFE_SEARCH_FULLSCAN=1)
for width, roots in ((4096, 256), (4096, 16), (1024, 256), (1024, 16), (1024, 1), (256, 16), (256, 1), (64, 8), (64, 1),
                     (16, 1)):
    key = run("incremental", GFE_SPEC_WIDTH=width, GFE_SPEC_ROOTS=roots)
    if key != ref:
        print(json.dumps({"ERROR": "rows differ from the full-scan driver", "width": width, "roots": roots}))
eng.close()
