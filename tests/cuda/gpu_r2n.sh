#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tc_varlen.py tests/test_gpu_random_sweep.py tests/test_gpu_generic.py tests/test_gpu_dropout_exact.py tests/test_gpu_fullsize.py tests/test_gpu_tc_mask.py -q -m gpu -x > gpurun_out/pytest_n.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_n.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_n.log | tail -10
python tests/cuda/short_end_det.py 2>&1 | head -4
python tests/cuda/quick_bench.py varlen 2>&1 | tail -5
