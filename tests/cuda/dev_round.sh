#!/bin/bash
# Developer GPU pass: parity tests, quick timings with a clock log, in-kernel traces.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_event_reasons.active --format=csv,noheader -lms 100 > gpurun_out/clocks_quick.csv &
SMI=$!
python tests/cuda/quick_bench.py > gpurun_out/quick_bench.log 2>&1; cat gpurun_out/quick_bench.log
kill $SMI
sort gpurun_out/clocks_quick.csv | uniq -c | sort -rn | head -8
for w in fwd bwd; do for c in 0 1; do
  echo "=== trace $w causal=$c"; FA_B200_TRACE=1 python tests/cuda/trace_run.py $w 0.0 $c; done; done > gpurun_out/trace.log 2>&1
echo "=== trace fwd dropout"; FA_B200_TRACE=1 python tests/cuda/trace_run.py fwd 0.1 0 >> gpurun_out/trace.log 2>&1
echo "=== trace bwd dropout"; FA_B200_TRACE=1 python tests/cuda/trace_run.py bwd 0.1 0 >> gpurun_out/trace.log 2>&1
tail -5 gpurun_out/trace.log
