#!/bin/bash
# Developer GPU pass for the forward kernel: parity tests, tuning variants, traces.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "fwd or varlen or generic" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for v in "" e0 e2 e6; do echo "== variant '$v'"; FA_B200_VARIANT=$v python tests/cuda/quick_bench.py fwd; done 2>&1 | tee gpurun_out/quick_fwd.log
for v in trace trace0; do for c in 0; do
  echo "=== $v fwd causal=$c"; FA_B200_VARIANT=$v python tests/cuda/trace_run.py fwd 0.0 $c; done; done > gpurun_out/trace_fwd.log 2>&1
echo "=== trace fwd dropout"; FA_B200_VARIANT=trace python tests/cuda/trace_run.py fwd 0.1 0 >> gpurun_out/trace_fwd.log 2>&1
