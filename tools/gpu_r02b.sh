#!/bin/bash
# round-2 pass b (1 GPU): persistent ring smoother (k_smooth_ring) and option store_fine_residual: parity, bench, ncu
set -x
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -q -x --durations=5 ) > gpurun_out/pytest_r02b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_r02b.log
tail -12 gpurun_out/pytest_r02b.log
Q="--no-cpu-baseline --no-e2e --no-strong --no-parity"
timeout 300 python bench.py $Q > gpurun_out/bench_r02b_ring1.json 2> gpurun_out/bench_r02b_ring1.err
timeout 300 python bench.py $Q --opt ring=0 > gpurun_out/bench_r02b_ring0.json 2> gpurun_out/bench_r02b_ring0.err
timeout 300 python bench.py $Q --opt ring=2 > gpurun_out/bench_r02b_ring2.json 2> gpurun_out/bench_r02b_ring2.err
timeout 300 python bench.py $Q --opt store_fine_residual=0 > gpurun_out/bench_r02b_nores.json 2> gpurun_out/bench_r02b_nores.err
timeout 300 python bench.py $Q --op helmholtz > gpurun_out/bench_r02b_helmholtz.json 2> gpurun_out/bench_r02b_helmholtz.err
CMD="python bench.py --steps 2 --warmup 3 $Q"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_ring -c 1 -o gpurun_out/prof_r02b_ring $CMD > gpurun_out/ncu_r02b_1.log 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r02b*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 2), "GDOF/s", round(d["ms_per_step"], 3), "ms", "roof", d.get("roofline") and round(d["roofline"]["frac"] or 0, 3),
              d["roofline"]["finest_level_ms"], "fast/gen", d.get("fast_launches_per_step"), d.get("general_launches_per_step"), d["max_res_after"])
    except Exception as e:
        print(f, "FAILED", e)
PY
tail -3 gpurun_out/*r02b*.err
