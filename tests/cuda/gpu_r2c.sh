#!/bin/bash
# Round-2 GPU pass C (one GPU): the no-allocation entry point's test alone in a fresh process, then the whole
# parity suite (which runs it last), the contract bench, the reference arm, the per-shape table.
set -u
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "pytest(alone) exit $?" >> gpurun_out/pytest_noalloc_alone.log
tail -4 gpurun_out/pytest_noalloc_alone.log
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|error|pytest exit" gpurun_out/pytest_gpu.log | tail -5
grep -E "^FAILED|^ERROR|^E  " gpurun_out/pytest_gpu.log | head -40
python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2>gpurun_out/bench.err; cat gpurun_out/bench.log; tail -3 gpurun_out/bench.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1; cat gpurun_out/bench_ref.log
python tests/cuda/quick_bench.py > gpurun_out/quick_bench.log 2>&1; cat gpurun_out/quick_bench.log
