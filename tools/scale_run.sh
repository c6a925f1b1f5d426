#!/bin/bash
# bench.py under torchrun at N ranks with optional environment: tools/scale_run.sh N tag [VAR=value ...]
n=$1; tag=$2; shift 2
env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
  bench.py --gpus $n --steps 8 --warmup 4 --no-configs --no-brute-leg > gpurun_out/scale_${tag}.json 2> gpurun_out/scale_${tag}.err
python - <<PY
import json
d=json.loads(open("gpurun_out/scale_${tag}.json").read().strip().splitlines()[-1])
nv=d["e2e"].get("nvlink_gather_variant") or {}
print("${tag}: kernel %.2f e2e %.2f dev %s nvlink %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"].get("device_ms"), nv.get("ms_per_step")))
PY
