#!/bin/bash
# round-1 pass n: state of HEAD on a fresh box: parity, default bench line (e2e + cpu arm), reference arm, 1024^3 on one GPU,
# launch list and full ncu captures of the finest-level kernels (Laplacian + vlaplacian smoother)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
timeout 600 python bench.py --n 1024 --no-cpu-baseline --no-e2e > gpurun_out/bench_1024.json 2> gpurun_out/bench_1024.err; echo "1024 rc=$?"
timeout 600 python bench.py --n 1024 --max-res --no-cpu-baseline --no-e2e > gpurun_out/bench_1024_maxres.json 2> gpurun_out/bench_1024_maxres.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_smooth_lpl_fast -c 1 -o gpurun_out/prof_smooth $CMD > gpurun_out/ncu_full1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_stencil_fast -c 2 -o gpurun_out/prof_stencil $CMD > gpurun_out/ncu_full2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_corr_prolong_fast -s 3 -c 1 -o gpurun_out/prof_prolong $CMD > gpurun_out/ncu_full3.log 2>&1
timeout 300 $CMD --op vlaplacian > gpurun_out/plain_v.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_smooth -c 1 -o gpurun_out/prof_vsmooth $CMD --op vlaplacian > gpurun_out/ncu_full5.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_stencil -c 1 -o gpurun_out/prof_vstencil $CMD --op vlaplacian > gpurun_out/ncu_full6.log 2>&1
ls -la gpurun_out
