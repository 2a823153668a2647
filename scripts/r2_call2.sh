#!/bin/bash
# GPU call: test suite, bench N=1 and N=2 (strong), in-process scaling
mkdir -p gpurun_out
df -h /dev/shm /tmp | cat
nvidia-smi -L
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench_C2_n1.json 2> gpurun_out/r2_bench_C2_n1.err; echo "bench n1 rc=$?"; tail -3 gpurun_out/r2_bench_C2_n1.err
N=$(nvidia-smi -L | wc -l)
if [ "$N" -ge 2 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2_bench_C2_n2.json 2> gpurun_out/r2_bench_C2_n2.err; echo "bench n2 rc=$?"; tail -5 gpurun_out/r2_bench_C2_n2.err
  python scripts/inprocess_scale.py C2 C4 C1 C3 > gpurun_out/r2_inprocess_n2.txt 2>&1; tail -12 gpurun_out/r2_inprocess_n2.txt
fi
python - <<'PY'
import json
for n in (1,2):
    try:
        d=json.loads([l for l in open(f"gpurun_out/r2_bench_C2_n{n}.json") if l.startswith("{")][-1])
        print(n, "value", d["value"], "ms/frame", d["ms_per_frame"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_frame"], "pinned", d.get("e2e_pinned",{}).get("ms_per_frame"), "frac", d["roofline"]["frac"], d["roofline"]["step_frac"], d["roofline"]["e2e_frac"])
        print("  extra", json.dumps(d["extra"]))
        print("  clocks", d["clocks"])
    except Exception as e: print(n, "FAILED", e)
PY
