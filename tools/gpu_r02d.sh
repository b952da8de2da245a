#!/bin/bash
# round-2 pass d (1 GPU): column-wise update loop in the smoothers; ring-kernel shape sweep (OMG_RING_CFG)
set -x
mkdir -p gpurun_out
T=r02d
( time timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multirank.py -q -x ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -5 gpurun_out/pytest_${T}.log
Q="--no-cpu-baseline --no-e2e --no-strong --no-parity"
timeout 300 python bench.py $Q --opt ring=0 > gpurun_out/bench_${T}_ring0.json 2> gpurun_out/bench_${T}_ring0.err
for CFG in 51222 25651 25622 51251 51231 25613; do
  OMG_RING_CFG=$CFG timeout 300 python bench.py $Q --opt ring=1 > gpurun_out/bench_${T}_ring_$CFG.json 2> gpurun_out/bench_${T}_ring_$CFG.err
done
timeout 300 python bench.py $Q --opt ring=0 --opt store_fine_residual=0 > gpurun_out/bench_${T}_ring0_nores.json 2> gpurun_out/bench_${T}_ring0_nores.err
CMD="python bench.py --steps 2 --warmup 3 $Q"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_lpl_fast -c 1 -o gpurun_out/prof_${T}_smooth $CMD --opt ring=0 > gpurun_out/ncu_${T}_1.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_ring -c 1 -o gpurun_out/prof_${T}_ring $CMD --opt ring=1 > gpurun_out/ncu_${T}_2.log 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r02d*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 2), "GDOF/s", round(d["ms_per_step"], 3), "ms", "roof", d.get("roofline") and round(d["roofline"]["frac"] or 0, 3), d.get("roofline") and d["roofline"].get("finest_level_ms"),
              "fast/gen", d.get("fast_launches_per_step"), d.get("general_launches_per_step"), d.get("max_res_after"))
    except Exception as e:
        print(f, "FAILED", e)
PY
for f in gpurun_out/*${T}*.err; do echo "== $f"; tail -3 $f; done 2>&1 | tail -40
