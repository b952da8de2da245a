#!/bin/bash
# quick multi-GPU comparison of options: bash tools/gpu_multi_quick.sh N tag "opt1" "opt2" ...   (each opt string = bench.py flags)
N=$1; T=$2; shift 2
mkdir -p gpurun_out
i=0
for OPTS in "$@"; do
  i=$((i+1))
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$i \
    bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --no-strong --no-parity $OPTS > gpurun_out/bench_${T}_n${N}_$i.json 2> gpurun_out/bench_${T}_n${N}_$i.err
  echo "[$i] rc=$? opts: $OPTS"
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/bench_${T}_n${N}_$i.json') if l.startswith('{')][-1])
    print('   ', d['n_gpus'], round(d['value'],2), 'GDOF/s', round(d['ms_per_step'],3), 'ms launches', d['launches_per_step'], d['config']['domain'], d['max_res_after'], d['roofline']['finest_level_ms'])
except Exception as e:
    print('    FAILED', e)
PY
done
