#!/bin/bash
# device-resident and e2e ms per frame with / without the streaming finisher, and for slab heights / finisher CTAs per SM
out=$1; wl=$2; steps=$3; shift 3
for s in $@; do
  IFS=: read mode rows ctas <<< "$s"
  r=$(FRB_TUNE_STREAM=$mode ${rows:+FRB_TUNE_STREAMROWS=$rows} ${ctas:+FRB_TUNE_STREAMCTAS=$ctas} python bench.py --workload $wl --steps $steps --warmup 3 --no-cpu-baseline 2>/dev/null \
      | python -c 'import json,sys; d=json.loads(sys.stdin.read()); r=d["roofline"]; print(r["kernel_ms"], r["shade_kernel_ms"], d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["e2e"]["value"])')
  echo "$wl stream=$mode rows=${rows:-auto} ctas=${ctas:-1} kernel_ms,exposed_finish_ms,step_ms,value,e2e_ms,e2e_value: $r" | tee -a "$out"
done
