#!/bin/bash
# Staggered readback plan of rt_render (RT_STAGGER_K sequences, sizes shrinking by RT_STAGGER_Q, round-robin over RT_STAGGER_S
# streams) on the headline frame: kernel / end-to-end ms per setting.  Run on the GPU box.
out=${1:-gpurun_out/sweep_stagger.txt}
: > $out
run() {
  python bench.py --steps 6 --warmup 4 --no-configs --no-brute-leg --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'kernel %.2f e2e %.2f dev %s' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e'].get('device_ms')))" >> $out
}
for s in 2 3 4; do for k in 5 6 8 10; do for q in 0.6 0.7 0.8; do
  RT_STAGGER_S=$s RT_STAGGER_K=$k RT_STAGGER_Q=$q run "stagger S $s K $k q $q"
done; done; done
cat $out
