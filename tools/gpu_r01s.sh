#!/bin/bash
# round-1 pass s: tail fast path for single-box levels between constant physical boundaries (smooth, residual, coarse rhs, correction)
set -x
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_s.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_s.log
tail -8 gpurun_out/pytest_s.log
timeout 600 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/bench_s.json 2> gpurun_out/bench_s.err; echo "bench rc=$?"
timeout 600 python bench.py --no-cpu-baseline --no-e2e --opt tail_simple=0 > gpurun_out/bench_s_nosimple.json 2> gpurun_out/bench_s_ns.err
timeout 600 python bench.py --no-cpu-baseline --no-e2e --tail-cells 32768 > gpurun_out/bench_s_tail32k.json 2> gpurun_out/bench_s_t32.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_s.csv $CMD > gpurun_out/ncu_list_s.log 2>&1
python - <<'PY'
import json
for f in ("bench_s", "bench_s_nosimple", "bench_s_tail32k"):
    try:
        d = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        print(f, round(d["value"], 2), round(d["ms_per_step"], 3), d["max_res_after"])
    except Exception as e:
        print(f, "FAILED", e)
PY
python tools/summarize_launches.py gpurun_out/launches_s.csv 2>/dev/null | grep -E "k_tail|total"
