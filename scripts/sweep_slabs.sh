#!/bin/bash
# e2e (host-buffer) ms per frame for different slab counts / deferred-kernel occupancy / equal vs shrinking slabs
out=$1; wl=$2; shift 2
for s in $@; do
  IFS=: read minb slabs eq <<< "$s"
  r=$(FRB_TUNE_MINB=$minb FRB_TUNE_SLABS=$slabs FRB_TUNE_SLABEQ=${eq:-1} python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null \
      | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["value"])')
  echo "$wl minb=$minb slabs=$slabs equal=${eq:-1} step_ms,e2e_ms,e2e_value: $r" | tee -a "$out"
done
