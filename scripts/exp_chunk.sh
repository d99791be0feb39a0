set -x
for lb in -1 0 1 2 3; do QSB200_CHUNK_LB=$lb python scripts/time_happly.py --n 28 --attempts 3 --profile; done 2>&1 | grep -v "^+" | tee gpurun_out/r1g_chunk.log
