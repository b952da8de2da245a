#!/bin/bash
# round-2 pass zk (1 GPU, the round's last GPU minutes): warp-level coarse solve on coarsest levels of up to four boxes --
# the per-step suite (tail fast path on (64,32,32) grids: two-box coarsest level), the multi-rank suite, smoke
set -x
mkdir -p gpurun_out
T=r02zk
( time timeout 150 python -m pytest tests/test_gpu_steps.py tests/test_gpu_multirank.py -m gpu -q -x --durations=3 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -9 gpurun_out/pytest_${T}.log
( timeout 40 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" ) > gpurun_out/smoke_${T}.log 2>&1; tail -2 gpurun_out/smoke_${T}.log
