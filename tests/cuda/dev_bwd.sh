#!/bin/bash
# Developer GPU pass for the backward kernel: parity tests, timings, trace.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "bwd or varlen" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for v in "" $*; do echo "== variant '$v'"; FA_B200_VARIANT=$v python tests/cuda/quick_bench.py bwd; done 2>&1 | tee gpurun_out/quick_bwd.log
