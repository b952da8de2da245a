#!/bin/bash
# round-2 pass x (1 GPU): the shared-memory staged lexicographic Gauss-Seidel sweep (k_gs_sweep_fast): parity tests with
# mg_smoother_gs, bench with --smoother gs (tuned / general path), launch list, one full ncu capture of the sweep
set -x
mkdir -p gpurun_out
T=${1:-r02x}
( time timeout 900 python -m pytest tests -m gpu -q -x -k "gs or config0 or smoke" --durations=5 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -12 gpurun_out/pytest_${T}.log
Q="--no-cpu-baseline --no-e2e --no-strong --no-configs"
timeout 300 python bench.py $Q --smoother gs > gpurun_out/bench_${T}_gs.json 2> gpurun_out/bench_${T}_gs.err
OMG_GS_FAST=0 timeout 300 python bench.py $Q --smoother gs > gpurun_out/bench_${T}_gs_general.json 2> gpurun_out/bench_${T}_gs_general.err
CMD="python bench.py --steps 2 --warmup 3 $Q --no-parity --smoother gs"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${T}.csv $CMD > gpurun_out/ncu_list_${T}.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_${T}.csv > gpurun_out/launches_${T}.txt; head -14 gpurun_out/launches_${T}.txt
mkdir -p /tmp/ncu
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_gs_sweep_fast -c 1 -o /tmp/ncu/gs_${T} $CMD > gpurun_out/ncu_full_${T}.log 2>&1
python tools/ncu_summary.py /tmp/ncu/gs_${T}.ncu-rep > gpurun_out/ncu_gs_${T}.txt 2>&1; head -40 gpurun_out/ncu_gs_${T}.txt
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r02*_gs*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 2), "GDOF/s", round(d["ms_per_step"], 3), "ms launches", d.get("launches_per_step"), "fast/gen", d.get("fast_launches_per_step"), d.get("general_launches_per_step"), d.get("max_res_after"), d["roofline"].get("finest_level_ms"), (d.get("parity_check") or {}).get("bitwise"))
    except Exception as e:
        print(f, "FAILED", e)
PY
