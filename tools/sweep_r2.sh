run() { python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-brute-leg --no-configs $2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1 kernel %.2f e2e %.2f dev %s'%(d['ms_per_step'],d['e2e']['ms_per_step'],d['e2e']['device_ms']))"; }
for b in 2 3 4; do for r in 2.5 3.5; do RT_BAND_RATIO=$r run "sequential $b ratio $r" "--bands $b"; done; done
RT_COPY_MODE=2 RT_BAND_RATIO=3.5 run "sequential 4 NO COPY" "--bands 4"
for b in 2 3 4 6; do RT_BAND_MODE=0 run "concurrent $b" "--bands $b"; done
RT_COPY_MODE=2 RT_BAND_MODE=0 run "concurrent 4 NO COPY" "--bands 4"
