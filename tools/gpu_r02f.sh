#!/bin/bash
# round-2 pass f (1 GPU): ring kernel with several producer warps; TMA load/store version of the correction + prolongation kernel
set -x
mkdir -p gpurun_out
T=r02f
( time timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_steps.py tests/test_gpu_multirank.py -q -x ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -5 gpurun_out/pytest_${T}.log
Q="--no-cpu-baseline --no-e2e --no-strong --no-parity"
for TMA in 0 3 4 5; do
  OMG_PROLONG_TMA=$TMA timeout 300 python bench.py $Q --opt ring=0 > gpurun_out/bench_${T}_prolong_tma$TMA.json 2> gpurun_out/bench_${T}_prolong_tma$TMA.err
done
for CFG in 256514 256512 512514 512414 256222 512222; do
  OMG_RING_CFG=$CFG timeout 300 python bench.py $Q --opt ring=1 > gpurun_out/bench_${T}_ring_$CFG.json 2> gpurun_out/bench_${T}_ring_$CFG.err
done
CMD="python bench.py --steps 2 --warmup 3 $Q"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_corr_prolong_tma -c 1 -o gpurun_out/prof_${T}_prolong $CMD --opt ring=0 > gpurun_out/ncu_${T}_1.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_ring -c 1 -o gpurun_out/prof_${T}_ring $CMD --opt ring=1 > gpurun_out/ncu_${T}_2.log 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r02f*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 2), "GDOF/s", round(d["ms_per_step"], 3), "ms", "roof", d.get("roofline") and round(d["roofline"]["frac"] or 0, 3), d.get("roofline") and d["roofline"].get("finest_level_ms"),
              "fast/gen", d.get("fast_launches_per_step"), d.get("general_launches_per_step"), d.get("max_res_after"))
    except Exception as e:
        print(f, "FAILED", e)
PY
