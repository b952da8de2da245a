#!/bin/bash
# round-2 pass zh (1 GPU): whole GPU suite at HEAD with put_fence = 2 by default (emulated-rank tests run the in-kernel puts), smoke, default bench line
set -x
mkdir -p gpurun_out
T=r02zh
( time timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -10 gpurun_out/pytest_${T}.log
( time timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" ) > gpurun_out/smoke_${T}.log 2>&1; tail -3 gpurun_out/smoke_${T}.log
( time timeout 900 python bench.py ) > gpurun_out/bench_${T}.json 2> gpurun_out/bench_${T}.err; echo "bench rc=$?" >> gpurun_out/bench_${T}.err
tail -3 gpurun_out/bench_${T}.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_r02zh.json").read().strip().splitlines()[0])
print(d["value"], d["ms_per_step"], d["roofline"]["frac"], d["cycle_hbm_frac"], d["e2e"]["value"], d["strong"]["ms_per_step"], d["parity_check"]["bitwise"], d["cpu_baseline"]["value"])
PY
