#!/bin/bash
# Round-2 GPU pass A (one box, one GPU): full parity suite, A/B of tuning variants, contract bench + reference arm,
# cross-check against the flash_attn wheel, compute-sanitizer memcheck on the small cases (one tool per call).
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|error|pytest exit" gpurun_out/pytest_gpu.log | tail -5
grep -E "^FAILED|^ERROR" gpurun_out/pytest_gpu.log | head -40
python tests/cuda/ab_bench.py fwd p10 stageo sg1 sg4 aig draw > gpurun_out/ab_fwd.log 2>&1; cat gpurun_out/ab_fwd.log
FA_B200_VARIANT=draw python -m pytest tests/test_gpu_dropout_exact.py tests/test_gpu_tc_fwd.py tests/test_gpu_tc_varlen.py -q -k "forward or dropout or many_items or d192" > gpurun_out/pytest_draw.log 2>&1; echo "pytest(draw) exit $?" >> gpurun_out/pytest_draw.log
grep -E "passed|failed|error|exit" gpurun_out/pytest_draw.log | tail -3; grep -E "^FAILED|^ERROR" gpurun_out/pytest_draw.log | head -20
AB_CAUSAL=0 python tests/cuda/ab_bench.py fwd draw > gpurun_out/ab_fwd_full.log 2>&1; cat gpurun_out/ab_fwd_full.log
AB_D=64 python tests/cuda/ab_bench.py fwd draw > gpurun_out/ab_fwd_d64.log 2>&1; cat gpurun_out/ab_fwd_d64.log
python tests/cuda/ab_bench.py bwd p10 sg1 sg4 > gpurun_out/ab_bwd.log 2>&1; cat gpurun_out/ab_bwd.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2>gpurun_out/bench.err; cat gpurun_out/bench.log; tail -3 gpurun_out/bench.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1; cat gpurun_out/bench_ref.log
python tests/cuda/xcheck_flash_attn.py > gpurun_out/xcheck_flash_attn.log 2>&1; cat gpurun_out/xcheck_flash_attn.log | tail -20
python tests/cuda/quick_bench.py > gpurun_out/quick_bench.log 2>&1; cat gpurun_out/quick_bench.log
timeout 900 compute-sanitizer --tool memcheck python tests/cuda/sanitize_cases.py > gpurun_out/sanitize_memcheck.log 2>&1
echo "memcheck exit $?" >> gpurun_out/sanitize_memcheck.log; tail -8 gpurun_out/sanitize_memcheck.log
