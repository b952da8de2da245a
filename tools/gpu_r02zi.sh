#!/bin/bash
# round-2 pass ze (4 GPUs): the default bench line as the driver launches it at HEAD (in-kernel puts of the smoother and of the
# prolongation kernel on 8 real ranks, parity check on the NCCL / CUDA-IPC data plane, clean teardown)
set -x
mkdir -p gpurun_out
T=r02zi
nvidia-smi -L > gpurun_out/gpus_${T}.txt
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29651 \
  bench.py --gpus 4 --no-cpu-baseline > gpurun_out/bench_${T}_n4.json 2> gpurun_out/bench_${T}_n4.err; echo "default rc=$?"
tail -4 gpurun_out/bench_${T}_n4.err
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/bench_r02zi_n4.json") if l.startswith("{")][-1])
print(d["value"], d["ms_per_step"], d["launches_per_step"], d["parity_check"]["bitwise"], d["parity_check"].get("fused_prolongs"), d["strong"]["ms_per_step"], d["e2e"]["value"])
PY
