#!/bin/bash
# weak (512^3 per GPU) and strong (1024^3 in total, BASELINE configs[4]) scaling at the GPU counts given as arguments
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
for N in "$@"; do
  for MODE in weak strong; do
    EXTRA=""; [ $MODE = strong ] && EXTRA="--strong --cells 1024"
    timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
      bench.py --gpus $N --steps 10 --warmup 3 $EXTRA --trace gpurun_out/trace_n${N}_${MODE} > gpurun_out/bench_n${N}_${MODE}.json 2> gpurun_out/bench_n${N}_${MODE}.err
    echo "N=$N $MODE rc=$?"; python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/bench_n${N}_${MODE}.json') if l.startswith('{')][-1])
    print(d['n_gpus'], d['scaling'], round(d['value'],2), 'GDOF/s', round(d['ms_per_step'],3), 'ms', d['launches_per_step'], d['config']['domain'], d['max_res_after'], 'e2e', d['e2e'] and round(d['e2e']['value'],2))
except Exception as e:
    print('FAILED', e)
PY
    python - <<PY
import json
try:
    t=json.load(open('gpurun_out/trace_n${N}_${MODE}.rank0.json'))
    print('  trace rank 0: %.3f ms per cycle;' % (t['total_ms'] / t['cycles']), ', '.join('%s %.3f' % (k, v[0] / t['cycles']) for k, v in list(t['by_step_ms'].items())[:8]))
except Exception as e:
    print('  no trace', e)
PY
    grep -v "OMP_NUM_THREADS\|^\*\*\*\|destroy_process_group" gpurun_out/bench_n${N}_${MODE}.err | tail -3
  done
done

# Teardown experiment: does a multi-rank run with the end-to-end leg leave on its own when the
# library releases its peer mappings in order?  Short timeout: round 1 lost 80 GPU-minutes to runs that never exited.
N=${1:-2}
for ORD in 1 0; do
  t0=$(date +%s)
  OMG_ORDERED_TEARDOWN=$ORD OMG_BENCH_PLAIN_EXIT=1 timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29613 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/teardown_n${N}_ord$ORD.json 2> gpurun_out/teardown_n${N}_ord$ORD.err
  echo "teardown N=$N ordered=$ORD rc=$? after $(( $(date +%s) - t0 )) s (124 = never exited)"
done
