#!/bin/bash
# round-2 pass zc (1 GPU): halo put of the corrected faces from inside the prolongation kernel (option fused_prolong):
# the multi-rank suites first (emulated ranks on one GPU), then the whole GPU suite
set -x
mkdir -p gpurun_out
T=r02zc
( time timeout 900 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_standalone_ops.py tests/test_gpu_regrid.py tests/test_diffusion.py tests/test_free_space.py -m gpu -q -x ) > gpurun_out/pytest_${T}_multirank.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}_multirank.log
tail -8 gpurun_out/pytest_${T}_multirank.log
( time timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -10 gpurun_out/pytest_${T}.log
