#!/bin/bash
# Round-2 GPU pass G: deterministic dQ (num_splits == 1) as two launches (dK/dV, then mirrored dQ): parity, bit
# reproducibility, timing; and the fused kernel's own suites (the kernel source is shared).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_bwd.py tests/test_gpu_tc_varlen.py -q -m gpu -x -k "deterministic" > gpurun_out/pytest_det.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_det.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_det.log | tail -15
timeout 600 python tests/cuda/quick_bench.py det > gpurun_out/quick_det.log 2>&1; cat gpurun_out/quick_det.log
timeout 900 python -m pytest tests/test_gpu_tc_bwd.py tests/test_gpu_tc_varlen.py tests/test_gpu_dropout_exact.py tests/test_gpu_tc_mask.py -q -m gpu -x -k "not deterministic" > gpurun_out/pytest_fused.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_fused.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_fused.log | tail -8
