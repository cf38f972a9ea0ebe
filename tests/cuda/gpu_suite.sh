#!/bin/bash
# The whole -m gpu suite + smoke + the contract bench, as the round-end driver runs them.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_gpu.log | tail -12
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.log 2>&1; cut -c1-300 gpurun_out/bench_ref.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2>gpurun_out/bench.err; cat gpurun_out/bench.log; tail -3 gpurun_out/bench.err
python tests/cuda/quick_bench.py mask > gpurun_out/quick_mask.log 2>&1; cat gpurun_out/quick_mask.log
