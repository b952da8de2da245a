#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 tools/probe_overlap.py > gpurun_out/probe_n$N.log 2> gpurun_out/probe_n$N.err
echo "rc=$?"; cat gpurun_out/probe_n$N.log; tail -5 gpurun_out/probe_n$N.err
