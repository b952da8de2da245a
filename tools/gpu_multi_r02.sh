#!/bin/bash
# round-2 multi-GPU pass: the default bench line at N GPUs (weak 512^3 per GPU + strong 1024^3 record + bitwise parity check
# of the real NCCL / IPC data plane + end-to-end leg) with the per-rank step timeline; the run must exit on its own.
# usage: gpurun --gpus N -- bash tools/gpu_multi_r02.sh N [tag]
N=${1:-2}; T=${2:-r02m}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_${T}_n$N.txt
nvidia-smi topo -m >> gpurun_out/gpus_${T}_n$N.txt 2>&1
t0=$(date +%s)
timeout ${BENCH_TIMEOUT:-300} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
  bench.py --gpus $N --steps 10 --warmup 3 --trace gpurun_out/trace_${T}_n$N > gpurun_out/bench_${T}_n$N.json 2> gpurun_out/bench_${T}_n$N.err
echo "N=$N rc=$? after $(( $(date +%s) - t0 )) s (124 = never exited)"
python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/bench_${T}_n$N.json') if l.startswith('{')][-1])
    print(d['n_gpus'], d['scaling'], round(d['value'],2), 'GDOF/s', round(d['ms_per_step'],3), 'ms', 'launches', d['launches_per_step'], d['config']['domain'], d['max_res_after'])
    print(' e2e', d.get('e2e') and (round(d['e2e']['value'],2), round(d['e2e']['ms_per_step'],1)), 'strong', d.get('strong') and (d['strong']['ms_per_step'], d['strong']['value']), 'parity', d.get('parity_check') and (d['parity_check'].get('bitwise'), d['parity_check'].get('fused_smooths'), d['parity_check'].get('put_phases')))
    print(' finest', d['roofline']['finest_level_ms'], 'timers', d.get('timers_min_max_s'))
except Exception as e:
    print('FAILED', e)
try:
    for r in range($N):
        t=json.load(open('gpurun_out/trace_${T}_n$N.rank%d.json' % r))
        print('  trace rank %d: %.3f ms per cycle;' % (r, t['total_ms'] / t['cycles']), ', '.join('%s %.3f' % (k, v[0] / t['cycles']) for k, v in list(t['by_step_ms'].items())[:10]))
except Exception as e:
    print('  no trace', e)
PY
grep "bench rank\|bench.py\|rror\|File \"/root" gpurun_out/bench_${T}_n$N.err | tail -60
