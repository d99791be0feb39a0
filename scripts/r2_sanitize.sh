#!/bin/bash
# compute-sanitizer sign-off (memcheck + racecheck) over scripts/sanitize_workload.py; logs -> gpurun_out/r2_sanitizer_*.log
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
timeout -s KILL 120 python scripts/sanitize_workload.py > gpurun_out/r2_sanitizer_plain.log 2>&1; tail -2 gpurun_out/r2_sanitizer_plain.log
for part in small prop tiled steps shard; do
  timeout -s KILL 900 $CS --tool memcheck --error-exitcode 9 --print-limit 20 python scripts/sanitize_workload.py $part > gpurun_out/r2_sanitizer_memcheck_$part.log 2>&1
  echo "memcheck $part rc=$?"; tail -3 gpurun_out/r2_sanitizer_memcheck_$part.log
done
for part in small prop tiled steps shard; do
  timeout -s KILL 900 $CS --tool racecheck --error-exitcode 9 --print-limit 20 python scripts/sanitize_workload.py $part > gpurun_out/r2_sanitizer_racecheck_$part.log 2>&1
  echo "racecheck $part rc=$?"; tail -3 gpurun_out/r2_sanitizer_racecheck_$part.log
done
