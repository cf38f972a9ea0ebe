#!/bin/bash
# Developer GPU pass: dropout parity + timing of both kernels.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "dropout or harness or c2_small" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python tests/cuda/quick_bench.py 2>&1 | grep -E "p=0.1|s4096 h32 d128 causal=True" | tee gpurun_out/quick_drop.log
