#!/bin/bash
# round-2 pass c (1 GPU), after a container restore that lost gpurun_out/ of pass a: ring-smoother sanity first (short
# limit: the kernel has only run in the emulator), then the whole GPU suite, smoke, the default bench line, the reference
# arm, option and operator sweeps, launch list, full ncu captures of the finest-level kernels.
set -x
mkdir -p gpurun_out
T=r02c
( time timeout 240 python -m pytest tests/test_gpu_parity.py -q -x -k "ring_smoother or config0 or store_fine" ) > gpurun_out/pytest_${T}_ring.log 2>&1; RC=$?
echo "ring sanity rc=$RC"; tail -5 gpurun_out/pytest_${T}_ring.log
RINGOPT=""
if [ $RC -ne 0 ]; then RINGOPT="--opt ring=0"; export OMG_RING_DEFAULT=0; echo "RING FAILED: the rest of the pass runs with ring=0"; fi
( time timeout 1500 python -m pytest tests -m gpu -q --durations=8 ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -25 gpurun_out/pytest_${T}.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_${T}.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_${T}.log
( time timeout 900 python bench.py $RINGOPT ) > gpurun_out/bench_${T}.json 2> gpurun_out/bench_${T}.err; echo "bench rc=$?"; tail -5 gpurun_out/bench_${T}.err
( time timeout 600 python bench.py --impl reference --steps 20 --warmup 5 ) > gpurun_out/bench_${T}_reference.json 2> gpurun_out/bench_${T}_reference.err; echo "ref rc=$?"
Q="--no-cpu-baseline --no-e2e --no-strong --no-parity"
timeout 300 python bench.py $Q --opt ring=0 > gpurun_out/bench_${T}_ring0.json 2> gpurun_out/bench_${T}_ring0.err
if [ $RC -eq 0 ]; then
  timeout 300 python bench.py $Q --opt ring=1 > gpurun_out/bench_${T}_ring1.json 2> gpurun_out/bench_${T}_ring1.err
  timeout 300 python bench.py $Q --opt ring=2 > gpurun_out/bench_${T}_ring2.json 2> gpurun_out/bench_${T}_ring2.err
  timeout 300 python bench.py $Q --op helmholtz > gpurun_out/bench_${T}_helmholtz.json 2> gpurun_out/bench_${T}_helmholtz.err
fi
timeout 300 python bench.py $Q $RINGOPT --opt store_fine_residual=0 > gpurun_out/bench_${T}_nores.json 2> gpurun_out/bench_${T}_nores.err
timeout 300 python bench.py $Q $RINGOPT --opt pdl=1 > gpurun_out/bench_${T}_pdl.json 2> gpurun_out/bench_${T}_pdl.err
timeout 300 python bench.py $Q $RINGOPT --opt tail_cells=32768 > gpurun_out/bench_${T}_tc32k.json 2> gpurun_out/bench_${T}_tc32k.err
for OP in vlaplacian vhelmholtz ahelmholtz; do
  timeout 300 python bench.py $Q --op $OP > gpurun_out/bench_${T}_$OP.json 2> gpurun_out/bench_${T}_$OP.err
done
timeout 300 python bench.py $Q --op vlaplacian --opt tail_simple=2 > gpurun_out/bench_${T}_vlaplacian_ts2.json 2> gpurun_out/bench_${T}_vlaplacian_ts2.err
CMD="python bench.py --steps 2 --warmup 3 $Q $RINGOPT"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${T}.csv $CMD > gpurun_out/ncu_list_${T}.log 2>&1
if [ $RC -eq 0 ]; then
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_ring -c 1 -o gpurun_out/prof_${T}_ring $CMD > gpurun_out/ncu_${T}_1.log 2>&1
fi
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_stencil_fast -c 1 -o gpurun_out/prof_${T}_stencil $CMD --opt store_fine_residual=0 > gpurun_out/ncu_${T}_2.log 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r02c*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 2), "GDOF/s", round(d["ms_per_step"], 3), "ms", "e2e", d.get("e2e") and round(d["e2e"]["value"], 2),
              "roof", d.get("roofline") and round(d["roofline"]["frac"] or 0, 3), d.get("roofline") and d["roofline"].get("finest_level_ms"),
              "strong", d.get("strong") and round(d["strong"]["ms_per_step"], 2),
              "parity", d.get("parity_check") and d["parity_check"].get("bitwise"), "fast/gen", d.get("fast_launches_per_step"), d.get("general_launches_per_step"), d.get("max_res_after"))
    except Exception as e:
        print(f, "FAILED", e)
PY
tail -3 gpurun_out/*${T}*.err | tail -60
ls -la gpurun_out | tail -40
