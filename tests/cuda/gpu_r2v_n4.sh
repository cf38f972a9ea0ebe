#!/bin/bash
# FOUR GPUs: dropout in ring mode and the plain ring parity check on the current library.
set -u
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29514 tests/cuda/ring_dropout_check.py > gpurun_out/r2v_ring_dropout_n4.log 2>&1; echo "ring_dropout_check exit $?"
grep -E "parity|Error|error" gpurun_out/r2v_ring_dropout_n4.log | cut -c1-330 | tail -10
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29515 tests/cuda/ring_check.py 16384 1 > gpurun_out/r2v_ring_check_n4.log 2>&1; echo "ring_check exit $?"
grep -E "parity|workload|Error|error" gpurun_out/r2v_ring_check_n4.log | cut -c1-400 | tail -6
