#!/bin/bash
# N = 8: ring forward / forward+backward at S = 131072 against the number of SMs the persistent kernels leave to NCCL.
set -u
mkdir -p gpurun_out
for r in 8 0 4 16; do
  FA_B200_RING_SM_RESERVE=$r timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2951$((r % 10)) tests/cuda/ring_check.py 131072 0 > gpurun_out/ring_reserve_$r.log 2>&1
  echo "reserve $r: $(grep workload gpurun_out/ring_reserve_$r.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('fwd %.2f ms  fwd+bwd %.2f ms' % (d['fwd_ms'], d['fwd_bwd_ms']))")"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 tests/cuda/ring_check.py 16384 1 > gpurun_out/ring_check_n8.log 2>&1; echo "parity run exit $?: $(grep -c 'parity OK' gpurun_out/ring_check_n8.log) ranks OK"
