# round-2 sweeps (gpurun -- bash tools/sweep_r2.sh)
for d in 0 100 200 400 800 1600; do
python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-brute-leg --no-configs --grid3 $d 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('grid3 $d kernel %.2f e2e %.2f pair %d exact %d cells/nodes %d'%(d['ms_per_step'],d['e2e']['ms_per_step'],d['rays']['pair_tests'],d['rays']['exact_tests'],d['rays']['node_visits']))"
done
