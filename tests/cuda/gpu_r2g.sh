#!/bin/bash
# Round-2 GPU pass G: ordered dQ (num_splits == 1) with the new turn order: parity + bit reproducibility + timing
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_bwd.py tests/test_gpu_tc_varlen.py -q -m gpu -x -k "deterministic" > gpurun_out/pytest_det.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_det.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_det.log | tail -15
timeout 600 python tests/cuda/quick_bench.py det > gpurun_out/quick_det.log 2>&1; cat gpurun_out/quick_det.log
