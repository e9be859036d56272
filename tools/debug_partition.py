import ctypes as C, sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import presto_b200 as P, oracle_lib as O
from presto_b200.blocks import col_array
N = P.native
ctx = P.default_context()
rng = np.random.default_rng(80)
n = 50001
k1 = rng.integers(-10**6, 10**6, n).astype(np.int64)
k2 = rng.integers(0, 100, n).astype(np.int32)
k2n = (rng.random(n) < 0.05).astype(np.uint8)
keys = [P.Block.long_array(k1), P.Block.int_array(k2, k2n)]
want = O.partition_ids(keys, 8, 0)
dk = [b.to_device() for b in keys]
counts = (C.c_int64 * 8)(); h = C.c_void_p()
N.check(ctx.handle, N.lib().pg_partition_create(ctx.handle, col_array(dk), 2, None, n, 8, 0, counts, C.byref(h)))
ids = np.zeros(n, dtype=np.uint8)
N.check(ctx.handle, N.lib().pg_partition_ids(h, ids.ctypes.data, 0))
bad = np.nonzero(ids != want)[0]
print("mismatches", len(bad), list(counts), [int((want == p).sum()) for p in range(8)])
for i in bad[:10]:
    print(i, k1[i], k2[i], k2n[i], ids[i], want[i], O.hash_rows([P.Block.long_array(k1[i:i+1]), P.Block.int_array(k2[i:i+1], k2n[i:i+1])]))
