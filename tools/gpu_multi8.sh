#!/bin/bash
# weak-scaling bench at the GPU counts given as arguments (one box), default transport
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
for N in "$@"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus $N --steps 10 --warmup 3 --no-e2e > gpurun_out/bench_n${N}.json 2> gpurun_out/bench_n${N}.err
  echo "N=$N rc=$?"; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_n${N}.json').read().strip().splitlines()[-1])
    print(d['n_gpus'], round(d['value'],2), 'GDOF/s', round(d['ms_per_step'],3), 'ms', d['launches_per_step'], d['config']['domain'], d['max_res_after'])
except Exception as e:
    print('FAILED', e)
PY
  grep -v "OMP_NUM_THREADS\|^\*\*\*\|destroy_process_group" gpurun_out/bench_n${N}.err | tail -3
done
