#!/bin/bash
# GPU-side A/B of the wave_trace_filter configurations (RT_FILTER_CFG) on the 1080p cfg5 frame, brute-force search.
for C in 0 1 2 3 4 5 6 7; do
  RT_FILTER_CFG=$C python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-brute-leg --workload cfg5-small --accel 1 2>&1 | python tools/brief.py "filter cfg $C"
done
