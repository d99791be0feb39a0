#!/bin/bash
# One GPU-box pass: (optionally) the full GPU test suite, the default bench line, the ncu launch list and --set full
# captures of one H-apply of a DP5 step (+ the next stage's pass_first) and of the error sweep at N=28.
# usage: scripts/gpu_round.sh TAG [notest]
TAG=${1:-rX}
mkdir -p gpurun_out
if [ "$2" != "notest" ]; then
  timeout -s KILL 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -3 gpurun_out/${TAG}_pytest.log
fi
timeout -s KILL 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || tail -5 gpurun_out/${TAG}_bench.err
cut -c1-300 gpurun_out/${TAG}_bench.json
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-sharded"
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/${TAG}_ncu_list.log 2>&1
# launches matching the filter: prime 5, then 31 per step attempt (6 x [first, mid, centre, mid, last] + error_sums)
timeout -s KILL 900 ncu --set full --clock-control none --import-source on -k regex:'pass_' -s 115 -c 6 -f -o gpurun_out/${TAG}_dp5_apply $B > gpurun_out/${TAG}_ncu_full.log 2>&1
timeout -s KILL 900 ncu --set full --clock-control none -k regex:'error_sums' -s 3 -c 1 -f -o gpurun_out/${TAG}_error_sums $B >> gpurun_out/${TAG}_ncu_full.log 2>&1
for r in dp5_apply error_sums; do
  ncu -i gpurun_out/${TAG}_$r.ncu-rep --page raw --csv > gpurun_out/${TAG}_${r}_raw.csv 2>/dev/null
done
du -sh gpurun_out; ls -la gpurun_out/${TAG}_*
