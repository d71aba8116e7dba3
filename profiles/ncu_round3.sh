#!/bin/bash
# Round-3 ncu captures (run on the GPU box through gpurun; outputs under gpurun_out/, summaries are made in
# the build container with profiles/summarize_ncu.py and committed under profiles/).
set -x
NCU="ncu --clock-control none"
# 1. launch list of the bench command (per-launch durations are cold-cache and serialised: shares, not absolutes)
$NCU --metrics gpu__time_duration.sum -c 600 --csv --log-file gpurun_out/r03_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-extras --no-cfg5 --no-cpu-baseline --no-ncu --search-steps 3000 > gpurun_out/r03_launches_bench.out 2>&1
# 2. score_incr_kernel in the regime of the cfg5 search (50 M rows: ~1.2 M rows per node list)
$NCU --set full --import-source on -k regex:score_incr -s 6 -c 2 -f -o gpurun_out/r03_score_incr_cfg5 \
    python profiles/search_sweep_cfg5.py 400 50000000 one > gpurun_out/r03_score_incr_cfg5.out 2>&1
ls -la gpurun_out/*.ncu-rep
