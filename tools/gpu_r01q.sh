#!/bin/bash
# round-1 pass q: all GPU tests at HEAD (diffusion, copy_to/from_tree, ddiv), tail with face tables staged in shared memory
set -x
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/pytest_q.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_q.log
tail -8 gpurun_out/pytest_q.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"
timeout 600 python bench.py --no-cpu-baseline --no-e2e --opt tail_smem=0 > gpurun_out/bench_q_tailglobal.json 2> gpurun_out/bench_q_tg.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_q.csv $CMD > gpurun_out/ncu_list_q.log 2>&1
python - <<'PY'
import json
for f in ("bench_q", "bench_q_tailglobal"):
    try:
        d = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        print(f, round(d["value"], 2), round(d["ms_per_step"], 3), d["e2e"] and round(d["e2e"]["value"], 3))
    except Exception as e:
        print(f, "FAILED", e)
PY
grep k_tail gpurun_out/launches_q.csv | head -8
