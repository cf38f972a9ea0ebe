#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_mask.py -q -m gpu -x > gpurun_out/pytest_mask.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_mask.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_mask.log | tail -15
timeout 600 python tests/cuda/det_bench.py > gpurun_out/det_bench.log 2>&1; cat gpurun_out/det_bench.log
