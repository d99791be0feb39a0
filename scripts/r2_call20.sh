#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_gpu_propagator.py -x -q -k "hbm" > gpurun_out/r2r_pytest_hbm.log 2>&1; tail -15 gpurun_out/r2r_pytest_hbm.log
timeout -s KILL 600 python scripts/time_propagator_hbm.py 10 12 > gpurun_out/r2r_time_hbm.jsonl 2> gpurun_out/r2r_time_hbm.err; cat gpurun_out/r2r_time_hbm.jsonl; tail -5 gpurun_out/r2r_time_hbm.err
