#!/bin/bash
# scaling evidence on one box: scripts/scale_run.sh "2 4 8"   (run under gpurun --gpus 8)
mkdir -p gpurun_out
for n in $1; do
  for spec in "C2 weak 30" "C2 strong 30" "C4 strong 5"; do
    set -- $spec
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --workload $1 --scaling $2 --steps $3 --warmup 3 --no-cpu-baseline 2> gpurun_out/scale_${1}_n${n}_$2.err | grep '^{' > gpurun_out/scale_${1}_n${n}_$2.json
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/scale_${1}_n${n}_$2.json").read())
    print("$1 n=$n $2", "value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"].get("ms_per_step"))
except Exception as e:
    print("$1 n=$n $2 FAILED", e)
PY
  done
done
