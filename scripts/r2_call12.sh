#!/bin/bash
# round 2, GPU call 12: multi-segment chain kernel + persistent propagate context: tests, then C2 / C3 timing (M = 4 vs M = 1)
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_gpu_propagator.py tests/test_gpu_evolve.py tests/test_gpu_shard.py -x -q > gpurun_out/r2l_pytest.log 2>&1; tail -4 gpurun_out/r2l_pytest.log
timeout -s KILL 600 python scripts/time_propagator.py > gpurun_out/r2l_time_propagator.json 2> gpurun_out/r2l_time_propagator.err; cat gpurun_out/r2l_time_propagator.json; tail -2 gpurun_out/r2l_time_propagator.err
QSB200_PROP_M=1 timeout -s KILL 600 python scripts/time_propagator.py > gpurun_out/r2l_time_propagator_m1.json 2> gpurun_out/r2l_time_propagator_m1.err; cat gpurun_out/r2l_time_propagator_m1.json
