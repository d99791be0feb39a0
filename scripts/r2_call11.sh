#!/bin/bash
# round 2, GPU call 11 (2 GPUs): pass_centre_pipe - one-GPU tests (ranks as threads), then multi-process parity + timing vs the plain persistent kernel
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_shard.py -x -q -k "pipelined or random_table" > gpurun_out/r2k_pytest_pipe.log 2>&1; tail -4 gpurun_out/r2k_pytest_pipe.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout -s KILL 300 $T --master-port 29612 scripts/shard_parity.py --qubits 24 --method dp5 > gpurun_out/r2k_parity24.log 2>&1; grep '"ok"' gpurun_out/r2k_parity24.log || tail -5 gpurun_out/r2k_parity24.log
for P in 1 0; do
  QSB200_CENTRE_PIPE=$P timeout -s KILL 600 $T --master-port 2962$P scripts/time_shard.py 30 > gpurun_out/r2k_time_cpipe$P.jsonl 2> gpurun_out/r2k_time_cpipe$P.err
  echo "centre_pipe=$P"; grep '^{' gpurun_out/r2k_time_cpipe$P.jsonl | cut -c1-420; grep -o '"parity_max_rel_err": [0-9.e-]*\|"max_abs_diff": [0-9.e-]*\|"strong": {[^}]*}' gpurun_out/r2k_time_cpipe$P.jsonl | tr '\n' ' '; echo
done
