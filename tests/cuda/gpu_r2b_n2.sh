#!/bin/bash
# Round-2 GPU pass B (one box, TWO GPUs): one-process multi-device test, the parity files whose references
# changed, bench under torchrun (weak line + strong + ring_c5 keys).
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi_device.py tests/test_gpu_fullsize.py tests/test_gpu_tc_bwd.py -q -m gpu > gpurun_out/pytest_n2.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_n2.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/pytest_n2.log | tail -15
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.log 2> gpurun_out/bench_n2.err
cat gpurun_out/bench_n2.log; tail -5 gpurun_out/bench_n2.err
