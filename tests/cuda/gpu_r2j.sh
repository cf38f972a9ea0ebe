#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tc_bwd.py tests/test_gpu_tc_varlen.py tests/test_gpu_tc_mask.py tests/test_gpu_random_sweep.py tests/test_gpu_xcheck_wheel.py -q -m gpu -x -k "deterministic or mask or sweep or xcheck or wheel" > gpurun_out/pytest_m2.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_m2.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_m2.log | tail -15
python tests/cuda/det_bench.py > gpurun_out/det_bench.log 2>&1; cat gpurun_out/det_bench.log
