# round-2 sweeps of the frame-level knobs (gpurun -- bash tools/sweep_r2.sh)
for h in -1 2 4 6; do for b in 2 4 6 8; do
RT_HEAD_GEN=$h python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-brute-leg --bands $b --lanes $b 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('head $h bands $b kernel %.2f e2e %.2f dev %s launches %d'%(d['ms_per_step'],d['e2e']['ms_per_step'],d['e2e']['device_ms'],d['gpu_launches']))"
done; done
