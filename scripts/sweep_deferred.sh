#!/bin/bash
# Sweep of the deferred kernel's shape on one B200: FRB_TUNE_DB (iterations per block) x FRB_TUNE_NB (blocks per trip) x
# FRB_TUNE_MINB (resident CTAs per SM).  Prints workload, shape, iterate-kernel ms (incl. resolve) and fraction of FP64 FMA peak.
out=${1:-gpurun_out/sweep_deferred.txt}
wl=${2:-C2}
steps=${3:-8}
: > "$out"
for minb in 4 5; do for db in 8 16 32; do for nb in 1 2 4 8; do
  [ "$db" = 32 ] && [ "$nb" = 8 ] && continue
  r=$(FRB_TUNE_DB=$db FRB_TUNE_NB=$nb FRB_TUNE_MINB=$minb python bench.py --workload $wl --steps $steps --warmup 3 --no-cpu-baseline --kernel defer6 2>/dev/null \
      | python -c 'import json,sys; d=json.loads(sys.stdin.read()); r=d["roofline"]; print(r["kernel_ms"], r["frac"], d["ms_per_step"], d["e2e"]["ms_per_step"], r["lane_utilisation"])')
  echo "$wl minb=$minb B=$db nb=$nb kernel_ms,frac,step_ms,e2e_ms,lane_util: $r" | tee -a "$out"
done; done; done
