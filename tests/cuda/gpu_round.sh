#!/bin/bash
# One GPU-box pass: parity tests, quick timings, contract bench (+ reference arm), ncu launch list.
# The `ncu --set full` capture is a separate call (tests/cuda/gpu_ncu_full.sh).
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python tests/cuda/quick_bench.py > gpurun_out/quick_bench.log 2>&1; cat gpurun_out/quick_bench.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; cat gpurun_out/bench.log
timeout 60 tests/cuda/red_probe > gpurun_out/red_probe.log 2>&1; tail -3 gpurun_out/red_probe.log
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1; cat gpurun_out/bench_ref.log
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_bench.log 2>&1
tail -2 gpurun_out/ncu_bench.log
