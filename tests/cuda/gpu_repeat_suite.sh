#!/bin/bash
# The whole GPU suite several times over (a timing-dependent hang shows as one slow failing test): reliability evidence.
set -u
mkdir -p gpurun_out
for r in 1 2 3 4; do
  python -m pytest tests -m gpu -x -q --durations=4 > gpurun_out/suite_$r.log 2>&1; rc=$?
  echo "== suite run $r rc $rc: $(grep -E ' passed| failed' gpurun_out/suite_$r.log | tail -1) | slowest: $(grep -E 's call' gpurun_out/suite_$r.log | head -2 | awk '{print $1, $3}' | tr '\n' ' ')"
done
