#!/bin/bash
# the bench lines and launch lists committed under profiles/ (run on the GPU box): scripts/round_bench.sh <tag>
tag=$1; mkdir -p gpurun_out
python bench.py > gpurun_out/${tag}_bench_C2_n1.json 2> gpurun_out/${tag}_bench_C2.err
for w in C1 C3 C4 C5; do python bench.py --workload $w --no-cpu-baseline > gpurun_out/${tag}_bench_${w}_n1.json 2> gpurun_out/${tag}_bench_${w}.err; done
python bench.py --impl reference > gpurun_out/${tag}_bench_C2_reference_arm.json 2> gpurun_out/${tag}_bench_ref.err
for w in C1 C2 C3 C4; do python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_bench_${w}_n1.json").read()); r=d["roofline"]
print("$w", d["run"]["kernel"], "kernel_ms", r["kernel_ms"], "shade", r["shade_kernel_ms"], "frac", r["frac"], "frame", d["ms_per_frame"], d["value"], "e2e", d["e2e"]["ms_per_frame"], d["e2e"]["value"], "pinned", d["e2e_pinned"]["ms_per_frame"], "launches", d["gpu_launches"], "clocks", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
PY
done
python -c "
import json
d=json.loads(open('gpurun_out/${tag}_bench_C5_n1.json').read()); print('C5', d['value'], d.get('frames_per_s'), d['ms_per_step'])
d=json.loads(open('gpurun_out/${tag}_bench_C2_reference_arm.json').read()); print('reference arm', d['value'], d['cpu_baseline'])"
for w in C1; do
  python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/${tag}_plain_${w}.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_bench_${w}.csv python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/${tag}_ncu_${w}.log 2>&1
done
