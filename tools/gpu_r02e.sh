#!/bin/bash
# round-2 pass e (1 GPU): stand-alone level operations across emulated ranks (f2), ring kernel with table prefetch
set -x
mkdir -p gpurun_out
T=r02e
( time timeout 600 python -m pytest tests/test_gpu_standalone_ops.py -q ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -30 gpurun_out/pytest_${T}.log
( time timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "ring_smoother or config0" ) > gpurun_out/pytest_${T}_ring.log 2>&1; echo "ring rc=$?"
Q="--no-cpu-baseline --no-e2e --no-strong --no-parity"
for CFG in 51222 25651 25622 51251 51231; do
  OMG_RING_CFG=$CFG timeout 300 python bench.py $Q --opt ring=1 > gpurun_out/bench_${T}_ring_$CFG.json 2> gpurun_out/bench_${T}_ring_$CFG.err
done
CMD="python bench.py --steps 2 --warmup 3 $Q"
OMG_RING_CFG=25651 timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_ring -c 1 -o gpurun_out/prof_${T}_ring $CMD --opt ring=1 > gpurun_out/ncu_${T}_2.log 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r02e*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 2), "GDOF/s", round(d["ms_per_step"], 3), "ms", "roof", d.get("roofline") and round(d["roofline"]["frac"] or 0, 3), d.get("roofline") and d["roofline"].get("finest_level_ms"),
              "fast/gen", d.get("fast_launches_per_step"), d.get("general_launches_per_step"), d.get("max_res_after"))
    except Exception as e:
        print(f, "FAILED", e)
PY
