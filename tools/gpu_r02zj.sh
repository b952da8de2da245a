#!/bin/bash
# round-2 pass zj (1 GPU): ncu --set full of the finest-level ring half-sweep and of the single-CTA tail at HEAD
set -x
mkdir -p gpurun_out /tmp/ncu
T=r02zj
Q="--no-cpu-baseline --no-amr --no-e2e --no-parity --no-strong --steps 2 --warmup 3"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_smooth_ring --launch-skip 16 -c 1 -o /tmp/ncu/ring_${T} python bench.py $Q > gpurun_out/ncu_ring_${T}.log 2>&1; echo "ncu rc=$?"
python tools/ncu_summary.py /tmp/ncu/ring_${T}.ncu-rep > gpurun_out/${T}_ncu_smooth_finest.txt 2>&1; head -8 gpurun_out/${T}_ncu_smooth_finest.txt
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_tail --launch-skip 3 -c 1 -o /tmp/ncu/tail_${T} python bench.py $Q > gpurun_out/ncu_tail_${T}.log 2>&1; echo "ncu rc=$?"
python tools/ncu_summary.py /tmp/ncu/tail_${T}.ncu-rep > gpurun_out/${T}_ncu_tail.txt 2>&1; head -8 gpurun_out/${T}_ncu_tail.txt
