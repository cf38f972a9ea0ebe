#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tc_bwd.py tests/test_gpu_tc_mask.py tests/test_gpu_random_sweep.py tests/test_gpu_generic.py tests/test_gpu_xcheck_wheel.py -q -m gpu -x > gpurun_out/pytest_wide2.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_wide2.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_wide2.log | tail -15
python tests/cuda/wide_bench.py det > gpurun_out/wide_bench2.log 2>&1; cat gpurun_out/wide_bench2.log
