#!/bin/bash
# GPU-side sweep of the wave_trace_bvh warp-scheduler thresholds (default leaf size, dense exact batches).
W=${1:-cfg5-box}
for cfg in "16 8 12" "16 4 12" "16 12 12" "12 8 12" "20 8 12" "16 8 16" "16 8 8" "12 4 16" "20 12 16" "16 16 12"; do
  set -- $cfg
  RT_WAVE_MIN_T=$1 RT_WAVE_DRAIN=$2 RT_WAVE_REFILL=$3 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-brute-leg --workload $W 2>&1 | python tools/brief.py "min_t $1 drain $2 refill $3"
done
