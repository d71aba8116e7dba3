#!/bin/bash
# 2-GPU `search` through gfalign_b200.multigpu on the cfg3 text files: peer-memory exchange vs NCCL.
set -e
D=$(mktemp -d)
python - "$D" <<'P'
import sys, os
sys.path.insert(0, '.')
from gfalign_b200 import synth
from bench_extras import _write_gaf_fast
d = sys.argv[1]
w = synth.make_workload("cfg3", A=400000); t = w.tangle
synth.write_gfa(os.path.join(d, "g.gfa"), t); synth.write_nodelist(os.path.join(d, "n.tsv"), t)
_write_gaf_fast(os.path.join(d, "a.gaf"), t.names, w.offsets, w.steps)
open(os.path.join(d, "sd"), "w").write(t.names[t.source] + " " + t.names[t.destination] + "\n")
P
read S T < "$D/sd"
for X in peer nccl; do
  T0=$(date +%s%N)
  env GFE_EXCHANGE=$X GFE_DEBUG_TIMING=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
    --master-addr 127.0.0.1 --master-port 2966$((RANDOM % 10)) -m gfalign_b200.multigpu search -f "$D/g.gfa" -g "$D/a.gaf" \
    -n "$D/n.tsv" -s "$S" -d "$T" -m 20000 > "$D/out_$X.txt" 2> "$D/err_$X.txt" || true
  echo "$X wall $(( ($(date +%s%N) - T0) / 1000000 )) ms"
  grep -h "multigpu" "$D/err_$X.txt" || tail -3 "$D/err_$X.txt"
done
cmp "$D/out_peer.txt" "$D/out_nccl.txt" && echo "stdout identical ($(wc -l < "$D/out_peer.txt") rows)"
gfalign_b200/bin/gfalign-b200 search -f "$D/g.gfa" -g "$D/a.gaf" -n "$D/n.tsv" -s "$S" -d "$T" -m 20000 > "$D/out_1.txt" && cmp "$D/out_1.txt" "$D/out_peer.txt" && echo "identical to the single-GPU CLI"
