#!/bin/bash
# Runs the UMMA probe over a list of descriptor encodings; one process each so a
# faulting variant cannot poison the next.  Output: gpurun_out/probe.log
mkdir -p gpurun_out
P=tests/cuda/umma_probe
LOG=gpurun_out/probe.log
: > $LOG
nvidia-smi --query-gpu=name,driver_version --format=csv >> $LOG 2>&1
run() { echo "--- $*" >> $LOG; timeout 60 $P "$@" >> $LOG 2>&1; echo "exit=$?" >> $LOG; }
run d=128
run d=64
run d=128 k_lbo=0
run d=128 p_col=0
run d=128 p_col=64
run d=128 mn_lbo=1024 mn_sbo=16384
run d=128 mn_step=32
run d=64 mn_lbo=1024 mn_sbo=8192
cat $LOG
