#!/bin/bash
# one summary line per (workload, kernel): scripts/bench_line.sh C1 tile [extra bench.py args]
w=$1; k=$2; shift 2
python bench.py --workload $w --steps 20 --no-cpu-baseline --kernel $k "$@" 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$w $k', d['config']['kernel'], 'kernel_ms', r['kernel_ms'], 'shade', r['shade_kernel_ms'], 'frac', r['frac'], 'frame', d['ms_per_frame'], d['value'], 'e2e', d['e2e']['ms_per_frame'], d['e2e']['value'], 'lane', r['lane_utilisation'])"
