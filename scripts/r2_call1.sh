#!/bin/bash
# round 2, GPU call 1: new parity tests, short-run pass experiment, bench line with the new legs
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > gpurun_out/r2a_box.txt; free -g >> gpurun_out/r2a_box.txt; nproc >> gpurun_out/r2a_box.txt
timeout -s KILL 900 python -m pytest tests/test_gpu_large.py tests/test_gpu_happly.py -x -q > gpurun_out/r2a_pytest_large.log 2>&1; tail -5 gpurun_out/r2a_pytest_large.log
scripts/exp_short_runs.sh gpurun_out/r2a_short_runs.jsonl 2> gpurun_out/r2a_short_runs.err
timeout -s KILL 900 python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err || tail -5 gpurun_out/r2a_bench.err
cut -c1-400 gpurun_out/r2a_bench.json
