#!/bin/bash
# GPU-side sweep of the wave_trace_bvh warp-scheduler thresholds (default leaf size).
W=${1:-cfg5-box}
for cfg in "20 12 8" "16 12 8" "24 12 8" "20 8 8" "20 16 8" "20 12 12" "20 12 4" "16 8 8" "12 8 8" "16 8 12"; do
  set -- $cfg
  RT_WAVE_MIN_T=$1 RT_WAVE_DRAIN=$2 RT_WAVE_REFILL=$3 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-brute-leg --workload $W 2>&1 | python tools/brief.py "min_t $1 drain $2 refill $3"
done
