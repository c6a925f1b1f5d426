#!/bin/bash
# GPU-side profiling pass (B200_PROFILING.md recipe): launch list of the default bench command, then one
# `--set full` capture of the two trace kernels.  Every ncu run follows a plain run of the same command that exited 0.
# usage: tools/gpu_profile.sh <tag>      -> gpurun_out/<tag>_*
set -u
TAG=${1:-prof}
O=gpurun_out
mkdir -p $O
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-brute-leg"
$B > $O/${TAG}_bench_plain.json 2> $O/${TAG}_bench_plain.err || { echo "plain bench failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $O/${TAG}_launches_cfg5.csv $B > $O/${TAG}_ncu_launches.log 2>&1
S="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-brute-leg --workload cfg5-small"
$S > $O/${TAG}_bench_small.json 2>&1 || { echo "small bench failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:wave_trace_bvh -c 4 -f -o $O/${TAG}_trace_bvh $S > $O/${TAG}_ncu_bvh.log 2>&1
F="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-brute-leg --workload cfg5-small --accel 1"
$F > $O/${TAG}_bench_filter.json 2>&1 || { echo "filter bench failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:wave_trace_filter -c 1 -f -o $O/${TAG}_trace_filter $F > $O/${TAG}_ncu_filter.log 2>&1
ls -la $O | tail -20
# the dominant kernel on the headline workload itself (8K frame): first two generations (primary rays, shadow rays)
ncu --set full --clock-control none --import-source on -k regex:wave_trace_bvh -c 2 -f -o $O/${TAG}_trace_bvh_8k $B > $O/${TAG}_ncu_bvh_8k.log 2>&1
ls -la $O | tail -8
