#!/bin/bash
# GPU-side A/B of wave_trace_bvh variants on a cfg5 frame size (default: the full 8K frame).
W=${1:-cfg5-box}
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-brute-leg --workload $W"
for L in 2 1; do for F in 1 0; do
  RT_WAVE_LEAF_FILTER=$F $B --leaf $L 2>&1 | python tools/brief.py "leaf $L leaf_filter $F"
  RT_B200_LIB=build/librt_b200_l1.so RT_WAVE_LEAF_FILTER=$F $B --leaf $L 2>&1 | python tools/brief.py "L1 leaf $L leaf_filter $F"
done; done
