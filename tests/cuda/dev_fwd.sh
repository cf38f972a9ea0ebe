#!/bin/bash
# Developer GPU pass for the forward kernel: parity tests, timings, traces.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "fwd or varlen or generic or shard" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for v in "" $*; do echo "== variant '$v'"; FA_B200_VARIANT=$v python tests/cuda/quick_bench.py fwd; done 2>&1 | tee gpurun_out/quick_fwd.log
echo "=== trace fwd" > gpurun_out/trace_fwd.log; FA_B200_VARIANT=trace python tests/cuda/trace_run.py fwd 0.0 0 >> gpurun_out/trace_fwd.log 2>&1
echo "=== trace fwd dropout" >> gpurun_out/trace_fwd.log; FA_B200_VARIANT=trace python tests/cuda/trace_run.py fwd 0.1 0 >> gpurun_out/trace_fwd.log 2>&1
