#!/bin/bash
# round-1 pass e: parity of the re-tuned kernels, variant timings, full ncu captures of the finest-level kernels
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e"
run() { tag=$1; shift; env "$@" timeout 300 $B $EXTRA > gpurun_out/var_$tag.json 2> gpurun_out/var_$tag.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/var_$tag.json').read().strip().splitlines()[-1])
    print('$tag', round(d['value'],2), 'GDOF/s', round(d['ms_per_step'],3), 'ms', {k: round(v,4) for k,v in d['roofline']['finest_level_ms'].items()})
except Exception as e:
    print('$tag FAILED', e)
PY
}
EXTRA="" run base X=1
EXTRA="" run stencil_mb2 OMG_STENCIL_MB=2
EXTRA="" run stencil_mb4 OMG_STENCIL_MB=4
EXTRA="" run prolong_mb2 OMG_PROLONG_MB=2
EXTRA="" run prolong_mb4 OMG_PROLONG_MB=4
EXTRA="" run smooth_kt8 OMG_SMOOTH_KT=8
EXTRA="--opt tail_smem=0" run tail_global X=1
EXTRA="--max-res" run maxres X=1
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_smooth_lpl_fast -c 1 -o gpurun_out/prof_smooth $CMD > gpurun_out/ncu_full1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_stencil_fast -c 1 -o gpurun_out/prof_stencil $CMD > gpurun_out/ncu_full2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_corr_prolong_fast -s 3 -c 1 -o gpurun_out/prof_prolong $CMD > gpurun_out/ncu_full3.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tail -c 1 -o gpurun_out/prof_tail $CMD > gpurun_out/ncu_full4.log 2>&1
ls -la gpurun_out
