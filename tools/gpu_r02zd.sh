#!/bin/bash
# round-2 pass zd (2 GPUs): the default bench line as the driver launches it (parity check on the NCCL / CUDA-IPC data plane
# incl. the prolongation kernel's in-kernel put), then the A/B of option fused_prolong
set -x
mkdir -p gpurun_out
T=r02zd
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 \
  bench.py --gpus 2 --no-cpu-baseline > gpurun_out/bench_${T}_n2.json 2> gpurun_out/bench_${T}_n2.err; echo "default rc=$?"
tail -3 gpurun_out/bench_${T}_n2.err
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/bench_r02zd_n2.json") if l.startswith("{")][-1])
print(d["value"], d["ms_per_step"], d["launches_per_step"], d["parity_check"], d["strong"]["ms_per_step"], d["e2e"]["value"])
PY
bash tools/gpu_multi_quick.sh 2 ${T} "--opt fused_prolong=1" "--opt fused_prolong=0" "--opt fused_prolong=1" "--opt fused_prolong=0"
