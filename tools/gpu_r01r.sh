#!/bin/bash
# round-1 pass r: source-level profile of the single-CTA tail
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tail -s 3 -c 1 -o gpurun_out/prof_tail_r $CMD > gpurun_out/ncu_r.log 2>&1
tail -3 gpurun_out/ncu_r.log
