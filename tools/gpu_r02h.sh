#!/bin/bash
# round-2 pass h (1 GPU): whole GPU suite at HEAD (free-space front end, ring launch order of fused levels, teardown order),
# smoke, ahelmholtz bench with lambda set, launch list + ncu captures (kept small: only summaries leave the box)
set -x
mkdir -p gpurun_out
T=r02h
( time timeout 1700 python -m pytest tests -m gpu -q --durations=8 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -30 gpurun_out/pytest_${T}.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_${T}.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_${T}.log
Q="--no-cpu-baseline --no-e2e --no-strong --no-parity"
timeout 300 python bench.py $Q --op ahelmholtz > gpurun_out/bench_${T}_ahelmholtz.json 2> gpurun_out/bench_${T}_ahelmholtz.err
timeout 300 python bench.py $Q > gpurun_out/bench_${T}_quick.json 2> gpurun_out/bench_${T}_quick.err
CMD="python bench.py --steps 2 --warmup 3 $Q"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${T}.csv $CMD > gpurun_out/ncu_list_${T}.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_${T}.csv > gpurun_out/launches_${T}.txt
mkdir -p /tmp/ncu
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_ring -c 1 -o /tmp/ncu/smooth $CMD > gpurun_out/ncu_${T}_1.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_stencil_fast -c 1 -o /tmp/ncu/stencil $CMD > gpurun_out/ncu_${T}_2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_corr_prolong_fast -s 4 -c 1 -o /tmp/ncu/prolong $CMD > gpurun_out/ncu_${T}_3.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_lpl_fast -c 1 -o /tmp/ncu/asmooth $CMD --op ahelmholtz > gpurun_out/ncu_${T}_4.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_stencil_fast -c 1 -o /tmp/ncu/astencil $CMD --op ahelmholtz > gpurun_out/ncu_${T}_5.log 2>&1
python tools/ncu_summary.py /tmp/ncu/smooth.ncu-rep > gpurun_out/${T}_ncu_smooth_finest.txt
python tools/ncu_summary.py /tmp/ncu/stencil.ncu-rep > gpurun_out/${T}_ncu_stencil.txt
python tools/ncu_summary.py /tmp/ncu/prolong.ncu-rep > gpurun_out/${T}_ncu_prolong.txt
python tools/ncu_summary.py /tmp/ncu/asmooth.ncu-rep > gpurun_out/${T}_ncu_ahelmholtz_smooth.txt
python tools/ncu_summary.py /tmp/ncu/astencil.ncu-rep > gpurun_out/${T}_ncu_ahelmholtz_stencil.txt
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r02h*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 2), "GDOF/s", round(d["ms_per_step"], 3), "ms", "roof", d.get("roofline") and round(d["roofline"]["frac"] or 0, 3), d.get("roofline") and d["roofline"].get("finest_level_ms"),
              "fast/gen", d.get("fast_launches_per_step"), d.get("general_launches_per_step"), d.get("max_res_after"))
    except Exception as e:
        print(f, "FAILED", e)
PY
ls -la gpurun_out | head -40; du -sh gpurun_out
