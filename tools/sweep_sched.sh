#!/bin/bash
# GPU-side sweep of the warp-scheduler thresholds of the persistent trace kernels (min lanes for a node / list step, lanes
# with a queued candidate that trigger an exact round, idle lanes that trigger a refill).
# usage: tools/sweep_sched.sh [bvh|grid] [workload]
K=${1:-bvh}; W=${2:-cfg5-box}
if [ "$K" = bvh ]; then P=RT_WAVE; else P=RT_GRID; fi
for cfg in "12 4 16" "8 4 16" "16 4 16" "12 2 16" "12 8 16" "12 4 8" "12 4 24" "8 2 12" "16 8 24" "6 4 16" "12 6 20"; do
  set -- $cfg
  env ${P}_MIN_T=$1 ${P}_DRAIN=$2 ${P}_REFILL=$3 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-brute-leg --no-configs --workload $W 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$K min_t $1 drain $2 refill $3: kernel %.2f ms e2e %.2f'%(d['ms_per_step'],d['e2e']['ms_per_step']))"
done
