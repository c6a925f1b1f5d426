run() { python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-brute-leg --no-configs $2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1 kernel %.2f e2e %.2f'%(d['ms_per_step'],d['e2e']['ms_per_step']))"; }
run base
for v in adv4 adv5 grid5 grid6; do RT_B200_LIB=build/librt_$v.so run "variant $v"; done
for a in 3 4 5 6 8; do RT_ADV_GRID=$a RT_B200_LIB=build/librt_adv4.so run "adv4 grid mult $a"; done
