#!/bin/bash
# round-2 pass g (1 GPU): whole GPU suite at HEAD (ring smoother on by default, ahelmholtz on the TMA kernels), default
# bench line + reference arm, operator sweep, two more ring shapes, launch list, ncu captures of the finest-level kernels
set -x
mkdir -p gpurun_out
T=r02g
( time timeout 1500 python -m pytest tests -m gpu -q --durations=5 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -12 gpurun_out/pytest_${T}.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_${T}.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_${T}.log
( time timeout 900 python bench.py ) > gpurun_out/bench_${T}.json 2> gpurun_out/bench_${T}.err; echo "bench rc=$?"; tail -5 gpurun_out/bench_${T}.err
Q="--no-cpu-baseline --no-e2e --no-strong --no-parity"
timeout 300 python bench.py $Q --store-res > gpurun_out/bench_${T}_storeres.json 2> gpurun_out/bench_${T}_storeres.err
for CFG in 256224 256132; do
  OMG_RING_CFG=$CFG timeout 300 python bench.py $Q > gpurun_out/bench_${T}_ring_$CFG.json 2> gpurun_out/bench_${T}_ring_$CFG.err
done
for OP in helmholtz vlaplacian vhelmholtz ahelmholtz; do
  timeout 300 python bench.py $Q --op $OP > gpurun_out/bench_${T}_$OP.json 2> gpurun_out/bench_${T}_$OP.err
done
timeout 300 python bench.py $Q --box 32 > gpurun_out/bench_${T}_box32.json 2> gpurun_out/bench_${T}_box32.err
CMD="python bench.py --steps 2 --warmup 3 $Q"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${T}.csv $CMD > gpurun_out/ncu_list_${T}.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_ring -c 1 -o gpurun_out/prof_${T}_smooth $CMD > gpurun_out/ncu_${T}_1.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_stencil_fast -c 1 -o gpurun_out/prof_${T}_stencil $CMD > gpurun_out/ncu_${T}_2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_corr_prolong_fast -s 4 -c 1 -o gpurun_out/prof_${T}_prolong $CMD > gpurun_out/ncu_${T}_3.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_lpl_fast -c 1 -o gpurun_out/prof_${T}_asmooth $CMD --op ahelmholtz > gpurun_out/ncu_${T}_4.log 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r02g*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 2), "GDOF/s", round(d["ms_per_step"], 3), "ms", "e2e", d.get("e2e") and round(d["e2e"]["value"], 2),
              "roof", d.get("roofline") and round(d["roofline"]["frac"] or 0, 3), d.get("roofline") and d["roofline"].get("finest_level_ms"),
              "strong", d.get("strong") and round(d["strong"]["ms_per_step"], 2), "parity", d.get("parity_check") and d["parity_check"].get("bitwise"),
              "fast/gen", d.get("fast_launches_per_step"), d.get("general_launches_per_step"), d.get("max_res_after"))
    except Exception as e:
        print(f, "FAILED", e)
PY
