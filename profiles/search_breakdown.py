import json, os, sys, time
sys.path.insert(0, ".")
from gfalign_b200 import capi, synth
w = synth.make_workload("cfg3", A=1_000_000)
t = w.tangle
nodes, counts, node_count = synth.node_table(t)
args = (t.adj_offsets, t.adj_target, t.adj_or0, t.adj_or1, t.source, t.destination, nodes, counts, node_count)
for width, roots in ((4096,256),):
    os.environ["GFE_SPEC_WIDTH"]=str(width); os.environ["GFE_SPEC_ROOTS"]=str(roots)
    eng = capi.Engine(0)
    eng.set_alignments(w.offsets, w.steps, t.n_nodes)
    eng.search(*args, max_steps=200)
    t0=time.perf_counter(); rows, st = eng.search(*args, max_steps=100000); dt=time.perf_counter()-t0
    print(width, roots, round(dt,4), st["kernel_batches"], st["paths_scored"], round(st["seconds_scoring"],4), flush=True)
    eng.close()
