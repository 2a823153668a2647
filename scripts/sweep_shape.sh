#!/bin/bash
# usage: sweep_shape.sh out workload steps "shape:db:nb:stream ..."   (shape = index into kDeferShapes of frb_api.cu)
out=$1; wl=$2; steps=$3; shift 3
for s in $@; do
  IFS=: read shape db nb stream <<< "$s"
  r=$(env FRB_TUNE_SHAPE=$shape FRB_TUNE_DB=$db FRB_TUNE_NB=$nb FRB_TUNE_STREAM=${stream:-2} python bench.py --workload $wl --steps $steps --warmup 3 --no-cpu-baseline 2>/dev/null \
      | python -c 'import json,sys; d=json.loads(sys.stdin.read()); r=d["roofline"]; print(r["kernel_ms"], r["shade_kernel_ms"], r["frac"], d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["e2e"]["value"])')
  echo "$wl shape=$shape B=$db nb=$nb stream=${stream:-2} kernel_ms,finish_ms,frac,step_ms,value,e2e_ms,e2e_value: $r" | tee -a "$out"
done
