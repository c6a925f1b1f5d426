run() { python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-brute-leg --no-configs --grid3 400 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1 kernel %.2f e2e %.2f'%(d['ms_per_step'],d['e2e']['ms_per_step']))"; }
for v in build/librt_dda4.so build/librt_dda5.so; do RT_B200_LIB=$v run "lib $v"; done
for cfg in "12 4 16" "8 4 16" "16 4 16" "12 2 16" "12 8 16" "12 4 8" "12 4 24" "20 8 16" "24 4 16" "16 8 24"; do set -- $cfg; RT_DDA_MIN_T=$1 RT_DDA_DRAIN=$2 RT_DDA_REFILL=$3 run "dda min_t $1 drain $2 refill $3"; done
