#!/bin/bash
# the bench lines and launch lists committed under profiles/ (run on the GPU box): scripts/round_bench.sh <tag>
tag=$1; mkdir -p gpurun_out
python bench.py > gpurun_out/${tag}_bench_C2_n1.json 2> gpurun_out/${tag}_bench_C2.err
for w in C1 C3 C4; do python bench.py --workload $w --no-cpu-baseline > gpurun_out/${tag}_bench_${w}_n1.json 2> gpurun_out/${tag}_bench_${w}.err; done
for w in C1 C2 C3 C4; do python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_bench_${w}_n1.json").read()); r=d["roofline"]
print("$w", d["config"]["kernel"], "kernel_ms", r["kernel_ms"], "shade", r["shade_kernel_ms"], "frac", r["frac"], "step", d["ms_per_step"], d["value"], "e2e", d["e2e"]["ms_per_step"], d["e2e"]["value"], "launches", d["gpu_launches"], "clocks", d["clocks"])
PY
done
for w in C2 C1; do
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_bench_${w}.csv python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu_${w}.log 2>&1
done
