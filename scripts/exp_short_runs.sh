#!/bin/bash
# Short-run strided passes at N = 28: time the H-apply's kernel classes with explicit group sizes,
# i.e. tiles made of 256- (baseline 8,8), 128-, 64-, 32- and 16-byte runs, in the mid and in the centre position.
# usage: scripts/exp_short_runs.sh OUT.jsonl
OUT=${1:-gpurun_out/short_runs.jsonl}
: > $OUT
for G in 8,8 9,7 7,9 10,6 6,10 11,5 5,11 12,4 4,12; do
  echo "groups $G" >&2
  QSB200_GROUPS=$G timeout -s KILL 300 python scripts/time_happly.py --n 28 --attempts 2 --profile | sed "s/^{/{\"groups\": \"$G\", /" >> $OUT
done
cat $OUT
