#!/bin/bash
# Round-2 GPU pass F (one box, EIGHT GPUs): ring context parallel through the C entry points (parity at S=16384,
# timing at S=131072), one-process multi-device test, world-1 ring test, bench under torchrun.
set -u
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 tests/cuda/ring_check.py 16384 1 > gpurun_out/ring_check_n8.log 2>&1; echo "ring_check exit $?"; grep -E "parity|workload|Error|error" gpurun_out/ring_check_n8.log | tail -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 tests/cuda/ring_check.py 131072 0 > gpurun_out/ring_c5_n8.log 2>&1; echo "ring_c5 exit $?"; grep -E "workload|Error|error" gpurun_out/ring_c5_n8.log | tail -5
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_n8.log 2> gpurun_out/bench_n8.err
cat gpurun_out/bench_n8.log; tail -5 gpurun_out/bench_n8.err
