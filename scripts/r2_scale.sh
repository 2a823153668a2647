#!/bin/bash
# round-2 multi-GPU evidence on one box: scripts/r2_scale.sh "1 2 4 8" <tag>     (run under gpurun --gpus 8)
# 1. in-process: frb_render(device = -1) on 1..N devices, every workload, full log (no traceback allowed)
# 2. one process per GPU (torchrun): bench.py --gpus N, default = north_star's partition (one frame, interleaved bands, one host image)
# 3. C5 animation by frame over all GPUs
tag=${2:-r2}; mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${tag}_gpus.txt 2>&1
python scripts/inprocess_scale.py C2 C4 C1 C3 > gpurun_out/${tag}_inprocess_scale.txt 2>&1; echo "inprocess rc=$?" >> gpurun_out/${tag}_inprocess_scale.txt
tail -3 gpurun_out/${tag}_inprocess_scale.txt | cut -c1-400
for n in $1; do
  if [ "$n" = 1 ]; then launcher="python"; else launcher="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n))"; fi
  timeout 600 $launcher bench.py --gpus $n --no-cpu-baseline 2> gpurun_out/${tag}_bench_C2_n${n}.err | grep '^{' > gpurun_out/${tag}_bench_C2_n${n}.json
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_bench_C2_n${n}.json").read()); x=d.get("extra",{})
    print("C2 n=$n", d["scaling"], "value", d["value"], "ms/step", d.get("ms_per_frame"), "e2e", d["e2e"]["value"], d["e2e"].get("ms_per_frame"),
          "eff", x.get("strong_efficiency"), "c4", (x.get("c4_strong") or {}).get("efficiency"), "mdc", x.get("multidevice_check"), "inproc", x.get("inprocess"))
except Exception as e:
    print("C2 n=$n FAILED", e)
PY
done
nmax=$(echo $1 | awk '{print $NF}')
if [ "$nmax" = 1 ]; then launcher="python"; else launcher="python -m torch.distributed.run --nnodes=1 --nproc-per-node $nmax --master-addr 127.0.0.1 --master-port 29600"; fi
timeout 600 $launcher bench.py --gpus $nmax --workload C5 --no-cpu-baseline 2> gpurun_out/${tag}_bench_C5_n${nmax}.err | grep '^{' > gpurun_out/${tag}_bench_C5_n${nmax}.json
python -c "
import json
d=json.loads(open('gpurun_out/${tag}_bench_C5_n${nmax}.json').read()); print('C5 n=$nmax value', d['value'], 'frames/s', d.get('frames_per_s'), 'ms/step', d['ms_per_step'])"
