import faulthandler, os, sys
faulthandler.dump_traceback_later(40, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from conftest import random_table
from quantumsimulations_b200 import shard, engine
n, world = 13, 2
table = random_table(n, 1)
print("creating group", flush=True)
grp = shard.LocalShardGroup(table, world)
print("group up", flush=True)
rng = np.random.default_rng(0)
psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
got = grp.apply_h(psi)
print("applied", flush=True)
from oracle import dense
bm = table.bit_major()
want = dense.apply_h(bm["h"], bm["J"], psi)
print("err", np.max(np.abs(got - want)) / np.max(np.abs(want)), flush=True)
grp.close()
