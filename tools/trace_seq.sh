run() { python bench.py --steps 3 --warmup 3 --no-configs --no-brute-leg --no-cpu-baseline 2>gpurun_out/tr.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'kernel %.2f e2e %.2f' % (d['ms_per_step'], d['e2e']['ms_per_step']))"; grep -A12 "seq trace" gpurun_out/tr.err | tail -13; }
export RT_SEQ_TRACE=1
run default
RT_STAGGER_K=5 RT_STAGGER_Q=0.65 run "K5 q.65"
RT_STAGGER_K=4 RT_STAGGER_Q=0.65 run "K4 q.65"
RT_STAGGER_K=6 RT_STAGGER_Q=0.65 RT_STAGGER_S=3 run "K6 q.65 S3"
RT_STAGGER_K=5 RT_STAGGER_Q=0.65 RT_STAGGER_S=3 run "K5 q.65 S3"
RT_STAGGER_K=8 RT_STAGGER_Q=0.7 RT_STAGGER_S=3 run "K8 q.7 S3"
