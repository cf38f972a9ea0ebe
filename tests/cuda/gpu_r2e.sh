#!/bin/bash
# Round-2 GPU pass E: the wide-head (136..256) tcgen05 backward: parity, exact masks, timing
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tc_bwd.py -q -m gpu -x -k "d256 or d192 or d160 or d136 or d224 or d200" > gpurun_out/pytest_wide.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_wide.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_wide.log | tail -25
timeout 600 python -m pytest tests/test_gpu_dropout_exact.py -q -m gpu -x -k "d256 or d192" > gpurun_out/pytest_wide_mask.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_wide_mask.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_wide_mask.log | tail -15
timeout 300 python tests/cuda/wide_bench.py > gpurun_out/wide_bench.log 2>&1; cat gpurun_out/wide_bench.log
