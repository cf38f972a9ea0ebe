#!/bin/bash
# Round-2 GPU pass D: packed attn_mask parity; per-step series of the bench workload after a cold start
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_tc_mask.py -q -m gpu -x > gpurun_out/pytest_mask.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_mask.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_mask.log | tail -15
python tests/cuda/step_series.py > gpurun_out/series_plain.log 2>&1; cat gpurun_out/series_plain.log
python tests/cuda/step_series.py smi > gpurun_out/series_smi.log 2>&1; cat gpurun_out/series_smi.log
python bench.py --steps 20 --warmup 5 --no-e2e > gpurun_out/bench2.log 2>gpurun_out/bench2.err; cut -c1-400 gpurun_out/bench2.log
