#!/bin/bash
# usage: sweep_some.sh out workload steps "minb:db:nb ..."  — bench.py at the given deferred-kernel shapes
out=$1; wl=$2; steps=$3; shift 3
for s in $@; do
  IFS=: read minb db nb <<< "$s"
  r=$(FRB_TUNE_DB=$db FRB_TUNE_NB=$nb FRB_TUNE_MINB=$minb python bench.py --workload $wl --steps $steps --warmup 3 --no-cpu-baseline --kernel defer6 2>/dev/null \
      | python -c 'import json,sys; d=json.loads(sys.stdin.read()); r=d["roofline"]; print(r["kernel_ms"], r["shade_kernel_ms"], r["frac"], d["ms_per_step"], d["e2e"]["ms_per_step"], r["lane_utilisation"])')
  echo "$wl minb=$minb B=$db nb=$nb kernel_ms,shade_ms,frac,step_ms,e2e_ms,lane_util: $r" | tee -a "$out"
done
