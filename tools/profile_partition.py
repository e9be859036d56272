"""One GPU: the exchange's kernels (histogram, scan, all-columns scatter) on an SF50-sized page with 8 local 'receivers', for ncu.
usage: python tools/profile_partition.py [rows] [nparts]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import presto_b200 as P
from presto_b200.blocks import Block, col_array

N = P.native
n = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000_000
nparts = int(sys.argv[2]) if len(sys.argv) > 2 else 8
ctx = P.GpuContext(0)
g = torch.Generator(device="cuda").manual_seed(1)
key = torch.randint(0, 1 << 40, (n,), device="cuda", generator=g)
cols = [Block(N.BIGINT, key, None, None, n), Block(N.SHORT_DECIMAL, torch.arange(n, device="cuda"), None, None, n),
        Block(N.DOUBLE, torch.arange(n, device="cuda"), None, None, n)]
ctx.set_timing(True)
for rep in range(int(os.environ.get("REPS", "2"))):
    if rep == 1:
        ctx.kernel_report(reset=True)
    counts = (C.c_int64 * nparts)()
    h = C.c_void_p()
    N.check(ctx.handle, N.lib().pg_partition_create(ctx.handle, col_array(cols[:1]), 1, None, n, nparts, 0, counts, C.byref(h)))
    bufs = [[torch.empty(max(int(counts[p]), 1), dtype=torch.int64, device="cuda") for p in range(nparts)] for _ in cols]
    pv = (C.c_void_p * (len(cols) * nparts))(*[b.data_ptr() for cb in bufs for b in cb])
    pn = (C.c_void_p * (len(cols) * nparts))(*([None] * (len(cols) * nparts)))
    po = (C.c_int64 * nparts)(*([0] * nparts))
    N.check(ctx.handle, N.lib().pg_partition_scatter_peers_all(h, col_array(cols), len(cols), pv, pn, po))
    torch.cuda.synchronize()
    N.lib().pg_partition_destroy(h)
print("rows per partition:", [int(c) for c in counts])
for k, v in sorted(ctx.kernel_report(reset=True).items(), key=lambda kv: -kv[1][0]):
    print("%-40s %8.3f ms per launch (%d launches)" % (k, v[0] / v[1], v[1]))
