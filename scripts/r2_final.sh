#!/bin/bash
# Round-2 evidence pass on one GPU: full GPU test suite, smoke, default bench line, ncu launch list, ncu --set full of the
# 26 launches of one DP5 step attempt at N = 28.   usage: scripts/r2_final.sh TAG [notest]
TAG=${1:-r2z}
mkdir -p gpurun_out
if [ "$2" == "shardtests" ]; then
  timeout -s KILL 900 python -m pytest tests/test_gpu_shard.py -x -q > gpurun_out/${TAG}_pytest_shard.log 2>&1; tail -3 gpurun_out/${TAG}_pytest_shard.log
elif [ "$2" != "notest" ]; then
  timeout -s KILL 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; tail -3 gpurun_out/${TAG}_pytest_gpu.log
  timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -4 gpurun_out/${TAG}_smoke.log
fi
timeout -s KILL 300 python scripts/sanitize_workload.py > gpurun_out/${TAG}_selfcheck_workload.log 2>&1; tail -3 gpurun_out/${TAG}_selfcheck_workload.log
timeout -s KILL 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || tail -5 gpurun_out/${TAG}_bench.err
cut -c1-300 gpurun_out/${TAG}_bench.json
timeout -s KILL 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; cut -c1-200 gpurun_out/${TAG}_bench_reference.json
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-sharded --no-grid"
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench_steps1.csv $B > gpurun_out/${TAG}_ncu_list.log 2>&1
# pass_* launches: prime 5, warm-up 3 x 26, then the timed step's 26 (four stage pairs are fused: pass_last_first)
timeout -s KILL 1500 ncu --set full --clock-control none --import-source on -k regex:'pass_' -s 83 -c 26 -f -o /tmp/${TAG}_dp5_step $B > gpurun_out/${TAG}_ncu_full.log 2>&1
# the .ncu-rep (126 MB) stays on the box: gpurun_out is capped at 64 MiB; the raw page and the source page of pass_centre / pass_last come back as CSV
ncu -i /tmp/${TAG}_dp5_step.ncu-rep --page raw --csv > gpurun_out/${TAG}_dp5_step_raw.csv 2>/dev/null
ncu -i /tmp/${TAG}_dp5_step.ncu-rep --page source --csv -k regex:'pass_last' -c 1 > gpurun_out/${TAG}_pass_last_source.csv 2>/dev/null
ncu -i /tmp/${TAG}_dp5_step.ncu-rep --page source --csv -k regex:'pass_centre' -c 1 > gpurun_out/${TAG}_pass_centre_source.csv 2>/dev/null
ls -la gpurun_out/${TAG}_* | awk '{print $5, $9}'
