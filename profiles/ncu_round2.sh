#!/bin/bash
# Round-2 ncu captures (run on the GPU box through gpurun; outputs under gpurun_out/, summaries are
# made in the build container with profiles/summarize_ncu.py and committed under profiles/).
set -x
NCU="ncu --clock-control none"
# 1. launch list of the bench command (per-launch durations are cold-cache and serialised: shares, not absolutes)
$NCU --metrics gpu__time_duration.sum -c 600 --csv --log-file gpurun_out/r02_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-extras --no-cfg5 --no-cpu-baseline --no-ncu --search-steps 3000 > gpurun_out/r02_launches_bench.out 2>&1
# 2. full captures of the dominant kernels
$NCU --set full --import-source on -k regex:score_bitset -s 3 -c 3 -f -o gpurun_out/r02_score_bitset \
    python bench.py --counters-child > /dev/null 2>&1
$NCU --set full --import-source on -k regex:score_incr -s 30 -c 3 -f -o gpurun_out/r02_score_incr \
    python profiles/search_breakdown.py > /dev/null 2>&1
$NCU --set full --import-source on -k "regex:edge_stream|pair_exact_kernel|pair_classify|pair_trace" -c 8 -f -o gpurun_out/r02_cfg4 \
    python bench_extras.py --reps 1 --no-cpu --trace-pairs 100000 > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep
