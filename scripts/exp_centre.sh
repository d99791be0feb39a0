#!/bin/bash
# Where does pass_centre lose its bandwidth?  (a) centre on the LOW strided group (QSB200_CENTRE_LOW=1: the far-strided
# top group goes to pass_mid), (b) the TMA-ring centre kernel on an unsharded plan (QSB200_CENTRE_RING=1).
# Each variant: product-state parity at N = 22 (two strided groups), then per-class timing at N = 28.
OUT=${1:-gpurun_out/exp_centre.jsonl}
: > $OUT
cat > /tmp/par.py <<'PY'
import os, sys
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np
from conftest import random_table
from oracle import product
from quantumsimulations_b200 import engine
n = 22
t = random_table(n, 5, dense_pairs=False); bm = t.bit_major(); phi = product.random_phi(n, 1)
rows = np.random.default_rng(0).integers(0, 1 << n, size=4096, dtype=np.uint64)
with engine.Plan(t) as p:
    got = p.apply_h_product(phi, rows)
want = product.apply_h_rows(bm["h"], bm["J"], phi, rows)
print("parity", engine.tile_layouts(n), float(np.max(np.abs(got - want)) / np.max(np.abs(want))))
PY
for V in "" "QSB200_CENTRE_LOW=1" "QSB200_CENTRE_RING=1"; do
  echo "variant [$V]" >&2
  env $V QSB200_MAX_GROUP=5 timeout -s KILL 120 python /tmp/par.py >&2
  env $V timeout -s KILL 300 python scripts/time_happly.py --n 28 --attempts 2 --profile | sed "s/^{/{\"variant\": \"$V\", /" >> $OUT
done
cat $OUT
