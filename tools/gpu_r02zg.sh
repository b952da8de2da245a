#!/bin/bash
# round-2 pass zg (2 GPUs): in-kernel puts with one system-scope fence per CTA (option put_fence = 2) against a fence per thread,
# natural box order against partition-face boxes first; the last run repeats the parity check on the real data plane with put_fence = 2
set -x
mkdir -p gpurun_out
T=r02zg
bash tools/gpu_multi_quick.sh 2 ${T} "--opt put_fence=1" "--opt put_fence=2" "--opt put_fence=1 --opt bnd_first_boxes=100000000" "--opt put_fence=2 --opt bnd_first_boxes=100000000" "--opt put_fence=2" "--opt put_fence=1"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29641 \
  bench.py --gpus 2 --no-cpu-baseline --no-e2e --no-strong --opt put_fence=2 > gpurun_out/bench_${T}_n2_parity.json 2> gpurun_out/bench_${T}_n2_parity.err; echo "parity run rc=$?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/bench_r02zg_n2_parity.json") if l.startswith("{")][-1])
print(d["value"], d["ms_per_step"], d["parity_check"]["bitwise"], d["parity_check"]["fused_smooths"], d["parity_check"]["fused_prolongs"])
PY
