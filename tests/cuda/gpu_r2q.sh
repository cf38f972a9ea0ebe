#!/bin/bash
# return_softmax on the tcgen05 forward: the new tests, the neighbours they could disturb, one timing.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tc_return_softmax.py tests/test_gpu_generic.py tests/test_gpu_tc_fwd.py tests/test_abi.py -x -q -m gpu > gpurun_out/r2q_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2q_pytest.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/r2q_pytest.log | tail -25
timeout 300 python tests/cuda/softmax_ret_bench.py > gpurun_out/r2q_softmax_ret_bench.log 2>&1; cat gpurun_out/r2q_softmax_ret_bench.log | tail -8
