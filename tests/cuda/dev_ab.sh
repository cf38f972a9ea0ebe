#!/bin/bash
# Developer A/B: quick timings of tuning variants (build.py --variant=x), '' = the product library.
#   bash tests/cuda/dev_ab.sh fwd|bwd variant...
set -u
mkdir -p gpurun_out
which=$1; shift
for v in "" $* ""; do echo "== variant '$v'"; FA_B200_VARIANT=$v python tests/cuda/quick_bench.py $which; done 2>&1 | tee gpurun_out/quick_ab.log
