#!/bin/bash
# GPU-side sweep of BVH leaf size / pipeline on a cfg5 frame size (default cfg5-small).
W=${1:-cfg5-small}
for L in 1 2 4; do
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-brute-leg --leaf $L --workload $W 2>&1 | python tools/brief.py leaf $L
done
for L in 1 4; do
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-brute-leg --leaf $L --pipeline 0 --workload $W 2>&1 | python tools/brief.py leaf $L
done
