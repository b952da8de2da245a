#!/bin/bash
# round-2 pass za (1 GPU): warp-level coarse solve (option warp_coarse): A/B on tail-dominated problems, the default bench line with it
# on and off, the whole GPU suite with it on (default)
set -x
mkdir -p gpurun_out
T=r02za
( time timeout 300 python tools/quick_warp_coarse.py ) > gpurun_out/warp_coarse_${T}.txt 2>&1; echo "rc=$?" >> gpurun_out/warp_coarse_${T}.txt
cat gpurun_out/warp_coarse_${T}.txt | tail -20
Q="--no-cpu-baseline --no-amr --no-e2e --no-parity"
( timeout 600 python bench.py $Q ) > gpurun_out/bench_${T}_on.json 2> gpurun_out/bench_${T}_on.err; echo "rc=$?" >> gpurun_out/bench_${T}_on.err
( timeout 600 python bench.py $Q --opt warp_coarse=0 ) > gpurun_out/bench_${T}_off.json 2> gpurun_out/bench_${T}_off.err; echo "rc=$?" >> gpurun_out/bench_${T}_off.err
python - <<'PY'
import json
for k in ("on", "off"):
    d = json.loads(open("gpurun_out/bench_r02za_%s.json" % k).read().strip().splitlines()[0])
    print(k, d["value"], d["ms_per_step"], {n: (c["ms_per_vcycle"], c["fmg_ms"]) for n, c in d.get("configs", {}).items()})
PY
( time timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -12 gpurun_out/pytest_${T}.log
