#!/bin/bash
# Round-2 evidence pass (one GPU): parity suite, per-shape timings, contract bench + reference arm, launch list,
# ncu --set full of the C3 kernels and of the wide-head / deterministic kernels.  profiles/collect.py r02e afterwards.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
python tests/cuda/quick_bench.py > gpurun_out/quick_bench.log 2>&1; cat gpurun_out/quick_bench.log
python tests/cuda/wide_bench.py > gpurun_out/wide_bench.log 2>&1; cat gpurun_out/wide_bench.log
python tests/cuda/det_bench.py > gpurun_out/det_bench.log 2>&1; cat gpurun_out/det_bench.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2>gpurun_out/bench.err; cat gpurun_out/bench.log; tail -3 gpurun_out/bench.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1; cat gpurun_out/bench_ref.log
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-sustained > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-sustained > gpurun_out/ncu_bench.log 2>&1
tail -2 gpurun_out/ncu_bench.log
python tests/cuda/prof_case.py 16 2 > gpurun_out/plain_prof.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'tc_fwd_kernel|tc_bwd_kernel|dq_convert|preprocess' -s 4 -c 4 \
    -f -o gpurun_out/prof_fwd_bwd python tests/cuda/prof_case.py 16 2 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
# the reports together exceed what gpurun copies back (64 MiB): summarise on the box, keep the text and the stall table
python profiles/summarize.py gpurun_out/prof_fwd_bwd.ncu-rep > gpurun_out/c3_fwd_bwd_ncu.txt 2>&1

rm -f gpurun_out/prof_fwd_bwd.ncu-rep
python tests/cuda/prof_case2.py > gpurun_out/plain_prof2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'tc_bwd_wide_kernel|tc_bwd_kernel' -s 3 -c 3 \
    -f -o gpurun_out/prof_wide_det python tests/cuda/prof_case2.py > gpurun_out/ncu_full2.log 2>&1
tail -3 gpurun_out/ncu_full2.log
python profiles/summarize.py gpurun_out/prof_wide_det.ncu-rep > gpurun_out/wide_det_ncu.txt 2>&1
rm -f gpurun_out/prof_wide_det.ncu-rep
python tests/cuda/short_end_case.py > gpurun_out/plain_short.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/short_end_launches.csv \
    python tests/cuda/short_end_case.py > gpurun_out/ncu_short.log 2>&1
python tests/cuda/host_latency.py > gpurun_out/host_latency.log 2>&1; cat gpurun_out/host_latency.log
ls -la gpurun_out/ | tail -12
