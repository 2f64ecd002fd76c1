import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np
from cudapathtracer_b200 import api
ctx = api.Context(0)
n = 1 << 26
rng = np.random.default_rng(1)
keys = rng.integers(0, 1 << 63, n, dtype=np.uint64); vals = np.arange(n, dtype=np.uint32)
k, v = ctx.debug_sort_pairs(keys, vals, 5)
print("sorted ok", bool(np.all(np.diff((k >> np.uint64(23)).astype(np.int64)) >= 0)))
