#!/bin/bash
# round-1 pass f: parity incl. mg_smoother_gs and the variable-coefficient fast kernels; config-4 operator timings
set -x
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e"
run() { tag=$1; shift; timeout 600 $B "$@" > gpurun_out/var_$tag.json 2> gpurun_out/var_$tag.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/var_$tag.json').read().strip().splitlines()[-1])
    print('$tag', round(d['value'],2), 'GDOF/s', round(d['ms_per_step'],3), 'ms', {k: round(v,4) for k,v in d['roofline']['finest_level_ms'].items()})
except Exception as e:
    print('$tag FAILED', e)
PY
tail -2 gpurun_out/var_$tag.err
}
run base
run vlpl --op vlaplacian
run vlpl_nofast --op vlaplacian --no-fast
run helmh --op helmholtz
run vhelmh --op vhelmholtz
run box32 --box 32
run lpl_tail512 --tail-cells 512
run lpl_tail64 --tail-cells 64
