#!/bin/bash
# round-2 pass j (1 GPU): cluster tail (k_tail_cluster: the levels up to 64 boxes of 16^3 + the old tail in one launch of a
# cluster of 8 CTAs): parity suite, then bench with the cluster tail on / off / other level limits
set -x
mkdir -p gpurun_out
T=r02j
( time timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -12 gpurun_out/pytest_${T}.log
Q="--no-cpu-baseline --no-e2e --no-strong --no-parity"
timeout 300 python bench.py $Q > gpurun_out/bench_${T}_cluster8.json 2> gpurun_out/bench_${T}_cluster8.err
timeout 300 python bench.py $Q --opt tail_cluster=0 > gpurun_out/bench_${T}_cluster0.json 2> gpurun_out/bench_${T}_cluster0.err
timeout 300 python bench.py $Q --opt tail_cluster_cells=32768 > gpurun_out/bench_${T}_cells32k.json 2> gpurun_out/bench_${T}_cells32k.err
timeout 300 python bench.py $Q --opt tail_cluster_cells=2097152 > gpurun_out/bench_${T}_cells2m.json 2> gpurun_out/bench_${T}_cells2m.err
timeout 300 python bench.py $Q --opt tail_cluster=4 > gpurun_out/bench_${T}_cluster4.json 2> gpurun_out/bench_${T}_cluster4.err
timeout 300 python bench.py $Q --opt tail_cluster_always=1 --opt tail_cluster_cells=4096 > gpurun_out/bench_${T}_cluster_tailonly.json 2> gpurun_out/bench_${T}_cluster_tailonly.err
CMD="python bench.py --steps 2 --warmup 3 $Q"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${T}.csv $CMD > gpurun_out/ncu_list_${T}.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_${T}.csv > gpurun_out/launches_${T}.txt; head -14 gpurun_out/launches_${T}.txt
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r02j*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 2), "GDOF/s", round(d["ms_per_step"], 3), "ms launches", d.get("launches_per_step"), "fast/gen", d.get("fast_launches_per_step"), d.get("general_launches_per_step"), d.get("max_res_after"))
    except Exception as e:
        print(f, "FAILED", e)
PY
tail -3 gpurun_out/bench_${T}_*.err | tail -30
