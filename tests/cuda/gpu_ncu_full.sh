#!/bin/bash
# One `ncu --set full` capture of the config-3 kernels (after the same command ran plainly).
set -u
mkdir -p gpurun_out
python tests/cuda/prof_case.py 16 2 > gpurun_out/plain_prof.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'tc_fwd_kernel|tc_bwd_kernel|dq_convert|preprocess' -s 4 -c 4 \
    -f -o gpurun_out/prof_fwd_bwd python tests/cuda/prof_case.py 16 2 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
