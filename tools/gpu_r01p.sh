#!/bin/bash
# round-1 pass p: diffusion + copy_to/from_tree + ddiv self-test parity, default bench line with the new end-to-end leg
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_diffusion.py tests/test_gpu_steps.py -m gpu -q > gpurun_out/pytest_p1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_p1.log
tail -25 gpurun_out/pytest_p1.log
timeout 600 python bench.py > gpurun_out/bench_default_p.json 2> gpurun_out/bench_default_p.err; echo "bench rc=$?"
timeout 600 python bench.py --e2e-cc --no-cpu-baseline > gpurun_out/bench_e2e_cc_p.json 2> gpurun_out/bench_e2e_cc_p.err; echo "bench rc=$?"
python - <<'PY'
import json
for f in ("bench_default_p", "bench_e2e_cc_p"):
    try:
        d = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        print(f, round(d["value"], 2), d["e2e"])
    except Exception as e:
        print(f, "FAILED", e)
PY
tail -3 gpurun_out/bench_default_p.err
