#!/bin/bash
set -u
mkdir -p gpurun_out
for r in 1 2 3; do
timeout 900 python -m pytest tests/test_gpu_tc_varlen.py -q -m gpu -x -k "empty_query" > gpurun_out/pytest_l$r.log 2>&1; echo "run $r: pytest exit $? $(grep -E 'passed|failed' gpurun_out/pytest_l$r.log | tail -1)"
done
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_gpu.log | tail -8
