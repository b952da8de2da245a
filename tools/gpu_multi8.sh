#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_8.txt
for N in 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus $N --steps 10 --warmup 3 --no-e2e > gpurun_out/bench_n${N}_graph.json 2> gpurun_out/bench_n${N}_graph.err
  echo "N=$N rc=$?"; python -c "
import json,sys
d=json.loads(open('gpurun_out/bench_n${N}_graph.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['launches_per_step'], d['config']['domain'])"; tail -3 gpurun_out/bench_n${N}_graph.err
done
