#!/bin/bash
# round-2 pass zb (1 GPU): tail_simple = 2 by default + warp-level coarse solve: whole GPU suite, default bench line, launch list
set -x
mkdir -p gpurun_out
T=r02zb
( time timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -14 gpurun_out/pytest_${T}.log
( time timeout 900 python bench.py ) > gpurun_out/bench_${T}.json 2> gpurun_out/bench_${T}.err; echo "bench rc=$?" >> gpurun_out/bench_${T}.err
tail -4 gpurun_out/bench_${T}.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_r02zb.json").read().strip().splitlines()[0])
print(d["value"], d["ms_per_step"], d["roofline"]["frac"], d["cycle_hbm_frac"], d["e2e"]["value"], d["strong"]["ms_per_step"], d["parity_check"]["bitwise"])
print({n: (c["ms_per_vcycle"], c.get("fmg_ms")) for n, c in d.get("configs", {}).items()})
PY
Q="--no-cpu-baseline --no-amr --no-e2e --no-parity --no-strong --steps 4 --warmup 3"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${T}.csv python bench.py $Q > gpurun_out/ncu_${T}.log 2>&1; echo "ncu rc=$?"
python tools/summarize_launches.py gpurun_out/launches_${T}.csv > gpurun_out/launches_${T}.txt 2>&1; head -12 gpurun_out/launches_${T}.txt
