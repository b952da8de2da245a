#!/bin/bash
# round-1 pass g (N GPUs): multi-rank parity on GPU 0 (emulated ranks, overlap path), then NCCL weak-scaling bench with / without overlap
N=${1:-2}
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q -k "uniform or fast" > gpurun_out/pytest_multirank.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multirank.log
tail -5 gpurun_out/pytest_multirank.log
for extra in "" "--opt overlap=0" "--opt tail_smem=0"; do
  tag=$(echo "$extra" | tr -d ' -=')
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus $N --steps 10 --warmup 3 --no-e2e $extra > gpurun_out/bench_n${N}_${tag:-overlap}.json 2> gpurun_out/bench_n${N}_${tag:-overlap}.err
  echo "rc=$?"; tail -c 600 gpurun_out/bench_n${N}_${tag:-overlap}.json | head -c 300; echo; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_n${N}_${tag:-overlap}.json').read().strip().splitlines()[-1])
    print('N=$N ${tag:-overlap}', round(d['value'],2), 'GDOF/s', round(d['ms_per_step'],3), 'ms', d['launches_per_step'])
except Exception as e:
    print('FAILED', e)
PY
  tail -3 gpurun_out/bench_n${N}_${tag:-overlap}.err
done
