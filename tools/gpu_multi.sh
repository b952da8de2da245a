#!/bin/bash
# usage: gpu_multi.sh N  -- weak-scaling bench on N GPUs of one box, with and without CUDA-graph replay
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_$N.txt
for extra in "" "--no-graph"; do
  tag=$(echo "$extra" | tr -d ' -')
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus $N --steps 10 --warmup 3 --no-e2e $extra > gpurun_out/bench_n${N}_${tag:-graph}.json 2> gpurun_out/bench_n${N}_${tag:-graph}.err
  echo "rc=$?"; tail -c 1500 gpurun_out/bench_n${N}_${tag:-graph}.json; tail -5 gpurun_out/bench_n${N}_${tag:-graph}.err
done
