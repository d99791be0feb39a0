#!/bin/bash
# Fused (L2-resident, TMA-fed) strided passes vs separate launches, then DRAM bytes under ncu.
mkdir -p gpurun_out
timeout -s KILL 240 python -m pytest tests/test_gpu_happly.py -x -q -m gpu 2>&1 | tail -15
{
for n in 28; do
  for cfg in "0 1" "20 1" "20 2"; do
    set -- $cfg
    QSB200_FUSE_LB=$1 QSB200_FUSE_LAG=$2 timeout -s KILL 120 python scripts/time_happly.py --n $n --attempts 3 --profile
  done
done
for n in 24 30; do
  QSB200_FUSE_LB=0 timeout -s KILL 120 python scripts/time_happly.py --n $n --attempts 3 --profile
  timeout -s KILL 120 python scripts/time_happly.py --n $n --attempts 3 --profile
done
} 2>&1 | tee gpurun_out/r1m_fused.log
timeout -s KILL 200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum \
  --clock-control none -k regex:"pass_strided_fused|pass_mid|pass_centre" -s 3 -c 3 --csv --log-file gpurun_out/r1m_fused_ncu.csv \
  python scripts/time_happly.py --n 28 --attempts 1 --order 4 > gpurun_out/r1m_fused_ncu.log 2>&1
tail -5 gpurun_out/r1m_fused_ncu.csv | cut -d, -f5,13-
QSB200_FUSE_LAG=2 timeout -s KILL 200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum \
  --clock-control none -k regex:"pass_strided_fused" -s 1 -c 1 --csv --log-file gpurun_out/r1m_fused_ncu_lag2.csv \
  python scripts/time_happly.py --n 28 --attempts 1 --order 4 > gpurun_out/r1m_fused_ncu2.log 2>&1
tail -5 gpurun_out/r1m_fused_ncu_lag2.csv | cut -d, -f5,13-
