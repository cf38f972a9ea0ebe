#!/bin/bash
# Round-2 GPU pass S (one box, TWO GPUs): dropout in ring mode (C entry points against the written per-step masks and the
# Python schedule), the ring tests of the suite, then the bench under torchrun (its ring_c5 key rides the same C code).
set -u
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 tests/cuda/ring_dropout_check.py > gpurun_out/r2s_ring_dropout_n2.log 2>&1; echo "ring_dropout_check exit $?"
grep -E "parity|Error|error" gpurun_out/r2s_ring_dropout_n2.log | cut -c1-400 | tail -8
timeout 600 python -m pytest tests/test_gpu_shard_ring.py -q -m gpu > gpurun_out/r2s_pytest_n2.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2s_pytest_n2.log
grep -E "passed|failed|error|pytest exit|^FAILED|^ERROR|^E  " gpurun_out/r2s_pytest_n2.log | tail -15
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2s_bench_n2.log 2> gpurun_out/r2s_bench_n2.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r2s_bench_n2.log").read().strip().splitlines()[-1])
    print("bench n2:", j["value"], j["unit"], "ms/step", j["ms_per_step"], "strong", j.get("strong"), "ring_c5", j.get("ring_c5"))
except Exception as e:
    print("bench n2 unreadable:", e)
PY
tail -3 gpurun_out/r2s_bench_n2.err
