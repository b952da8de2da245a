#!/bin/bash
# round-1 pass o: m_diffusion on the device + branch-free batched fp64 division in the variable-coefficient kernels
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_diffusion.py tests/test_gpu_steps.py -m gpu -x -q > gpurun_out/pytest_o1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_o1.log
tail -15 gpurun_out/pytest_o1.log
timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_diffusion.py --deselect tests/test_gpu_steps.py > gpurun_out/pytest_o2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_o2.log
tail -6 gpurun_out/pytest_o2.log
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
run vlpl --op vlaplacian
run vhelmh --op vhelmholtz
run vlpl32 --op vlaplacian --box 32
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --op vlaplacian"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_smooth -c 1 -o gpurun_out/prof_vsmooth2 $CMD > gpurun_out/ncu_o1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_stencil -c 1 -o gpurun_out/prof_vstencil2 $CMD > gpurun_out/ncu_o2.log 2>&1
ls gpurun_out | head -50
