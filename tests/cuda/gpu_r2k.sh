#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_tc_bwd.py tests/test_gpu_tc_varlen.py tests/test_gpu_tc_mask.py tests/test_gpu_dropout_exact.py tests/test_gpu_random_sweep.py tests/test_gpu_xcheck_wheel.py tests/test_zz_no_alloc.py -q -m gpu -x > gpurun_out/pytest_k.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_k.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_k.log | tail -15
python tests/cuda/det_bench.py > gpurun_out/det_bench.log 2>&1; cat gpurun_out/det_bench.log
