#!/bin/bash
# round-2 pass y (1 GPU): the allocation-free entry point omg_debug_ddiv_static -- its test alone in a fresh process, then
# the whole GPU suite (which it closes), then smoke()
set -x
mkdir -p gpurun_out
T=r02y
( time timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x ) > gpurun_out/pytest_${T}_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}_alone.log
tail -6 gpurun_out/pytest_${T}_alone.log
( time timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -16 gpurun_out/pytest_${T}.log
( time timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" ) > gpurun_out/smoke_${T}.log 2>&1; tail -4 gpurun_out/smoke_${T}.log
