#!/bin/bash
# First GPU call of the next round (1 GPU, ~6 min): everything that was written after round 1's GPU budget ran out.
#   1. the whole GPU suite incl. the opt-in tests (tail_simple = 2 variants, option pdl, the three C++ examples)
#   2. bitwise quick check + tail timing with the gated variants switched on
#   3. default bench line, launch list
set -x
mkdir -p gpurun_out
OMG_TEST_UNVERIFIED=1 timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/pytest_r2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_r2.log
tail -8 gpurun_out/pytest_r2.log
timeout 60 python tools/quick_check.py > gpurun_out/quick_check_r2.log 2>&1; tail -7 gpurun_out/quick_check_r2.log
timeout 60 python tools/quick_tail_timing.py > gpurun_out/quick_tail_timing_r2.log 2>&1; tail -6 gpurun_out/quick_tail_timing_r2.log
timeout 600 python bench.py > gpurun_out/bench_r2.json 2> gpurun_out/bench_r2.err; echo "bench rc=$?"
for OP in vlaplacian vhelmholtz; do
  for TS in 1 2; do
    timeout 600 python bench.py --no-cpu-baseline --no-e2e --op $OP --opt tail_simple=$TS > gpurun_out/bench_r2_${OP}_ts$TS.json 2> gpurun_out/bench_r2_${OP}_ts$TS.err
  done
done
# programmatic dependent launch of the TMA-staged kernels (option pdl, gated): effect on the default bench
timeout 600 python bench.py --no-cpu-baseline --no-e2e --opt pdl=1 > gpurun_out/bench_r2_pdl.json 2> gpurun_out/bench_r2_pdl.err
# the tail got 2.4x faster after the tail_cells sweep of pass c: does a larger tail pay now?
for TC in 32768; do
  timeout 600 python bench.py --no-cpu-baseline --no-e2e --opt tail_cells=$TC > gpurun_out/bench_r2_tc$TC.json 2> gpurun_out/bench_r2_tc$TC.err
  timeout 600 python bench.py --no-cpu-baseline --no-e2e --opt tail_cells=$TC --opt pdl=1 > gpurun_out/bench_r2_tc${TC}_pdl.json 2> gpurun_out/bench_r2_tc${TC}_pdl.err
done
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r2.csv $CMD > gpurun_out/ncu_list_r2.log 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r2*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["value"], 2), round(d["ms_per_step"], 3), d.get("e2e") and round(d["e2e"]["value"], 2))
    except Exception as e:
        print(f, "FAILED", e)
PY
# Multi-GPU afterwards, in its own call and with a short timeout per run (round 1 lost 80 GPU-minutes to two runs that
# printed their line and then blocked in teardown):
#   gpurun --gpus 4 --timeout 900 -- 'bash tools/gpu_multi_r.sh 2 4'
