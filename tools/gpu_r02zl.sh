#!/bin/bash
# round-2 pass zl (1 GPU, the last GPU seconds of the round): the GPU suite at HEAD without its one 80 s case
set -x
mkdir -p gpurun_out
T=r02zl
( time timeout 160 python -m pytest tests -m gpu -q -x --deselect "tests/test_gpu_parity.py::test_shapes_and_box_sizes[n1-8]" ) > gpurun_out/pytest_${T}.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${T}.log
tail -7 gpurun_out/pytest_${T}.log
