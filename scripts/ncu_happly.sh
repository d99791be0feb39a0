#!/bin/bash
# One H-apply of the tiled path at N=28 under ncu --set full (run only after the plain command exited 0).
set -e
CMD="python scripts/time_happly.py --n 28 --attempts 1 --order 4"
$CMD > gpurun_out/ncu_plain.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pass_ -s 14 -c 5 -f -o gpurun_out/r1i_happly $CMD > gpurun_out/ncu_run.log 2>&1
tail -3 gpurun_out/ncu_run.log
