#!/bin/bash
# round-2 pass z (1 GPU): HEAD after a container restore -- the default bench line (incl. the configs[2] AMR record, never run
# before), the reference arm as the driver calls it, launch list of the default command, the whole GPU suite, smoke()
set -x
mkdir -p gpurun_out
T=r02z
( time timeout 900 python bench.py ) > gpurun_out/bench_${T}.json 2> gpurun_out/bench_${T}.err; echo "bench rc=$?" >> gpurun_out/bench_${T}.err
tail -c 1500 gpurun_out/bench_${T}.json; tail -5 gpurun_out/bench_${T}.err
( time timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -16 gpurun_out/pytest_${T}.log
( time timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" ) > gpurun_out/smoke_${T}.log 2>&1; tail -4 gpurun_out/smoke_${T}.log
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/bench_${T}_reference.json 2> gpurun_out/bench_${T}_reference.err; echo "ref rc=$?" >> gpurun_out/bench_${T}_reference.err
tail -c 800 gpurun_out/bench_${T}_reference.json
