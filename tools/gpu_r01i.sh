#!/bin/bash
# round-1 pass i (N GPUs): emulated-rank parity on GPU 0 (peer-mapped put path), overlap probe, weak-scaling bench put vs send/recv
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q -k "transports or fast" > gpurun_out/pytest_multirank.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multirank.log
tail -5 gpurun_out/pytest_multirank.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 tools/probe_overlap.py > gpurun_out/probe_n$N.log 2> gpurun_out/probe_n$N.err
echo "probe rc=$?"; cat gpurun_out/probe_n$N.log; tail -3 gpurun_out/probe_n$N.err
for extra in "" "--opt fused=0 --opt overlap=0" "--opt p2p=0 --opt overlap=0"; do
  tag=$(echo "$extra" | tr -d ' -=')
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 \
    bench.py --gpus $N --steps 10 --warmup 3 --no-e2e $extra > gpurun_out/bench_n${N}_${tag:-default}.json 2> gpurun_out/bench_n${N}_${tag:-default}.err
  echo "rc=$?"; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_n${N}_${tag:-default}.json').read().strip().splitlines()[-1])
    print('N=$N ${tag:-default}', round(d['value'],2), 'GDOF/s', round(d['ms_per_step'],3), 'ms', d['launches_per_step'], d['max_res_after'])
except Exception as e:
    print('FAILED', e)
PY
  grep -v "OMP_NUM_THREADS\|^\*\*\*\|destroy_process_group" gpurun_out/bench_n${N}_${tag:-default}.err | tail -3
done
