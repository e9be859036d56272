"""Under torchrun: wall-clock breakdown (with synchronisation after every stage) of the partitioned GROUP BY orderkey step.
usage: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/exchange_breakdown.py [total_sf]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import presto_b200 as P
from presto_b200 import tpch
from presto_b200.blocks import Block

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
ctx = P.GpuContext(torch.cuda.current_device())
N = P.native
sf = (float(sys.argv[1]) if len(sys.argv) > 1 else 100.0) / world
li = tpch.lineitem(sf, seed=42 + rank, device="cuda", need=("keys", "double"))
n = li["rows"]
n_orders = int(1_500_000 * sf)
key_off = rank * tpch.order_key(torch.tensor(n_orders, dtype=torch.int64)).item()   # disjoint order-key ranges per rank, as in bench.py
li["orderkey"] += key_off
page = P.Page(Block(N.BIGINT, li["orderkey"], None, None, n), Block(N.SHORT_DECIMAL, li["quantity"], None, None, n),
              Block(N.DOUBLE, li["extendedprice_double"].view(torch.int64), None, None, n), position_count=n)
ex = P.HashPartitionExchange(P.DevicePartitioner(ctx))
aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 2)]
acc = {}


def tick(name, t0):
    torch.cuda.synchronize()
    acc[name] = acc.get(name, 0.0) + (time.perf_counter() - t0)
    return time.perf_counter()


reps = 5
for it in range(reps + 3):
    if it == 3:
        acc.clear()
    dist.barrier(); torch.cuda.synchronize()
    t = time.perf_counter()
    local = ex.exchange(page, [0], slot="agg", lease=True); t = tick("exchange", t)
    op = P.HashAggregationOperator([P.BIGINT], [0], P.SINGLE, aggs, expected_groups=n_orders, ctx=ctx, device_output=True); t = tick("create", t)
    op.add_input(local); t = tick("add_input", t)
    op.finish(); g = op.get_group_count(); t = tick("group_count", t)
    out = op.get_output(max_rows=max(g, 1)); t = tick("get_output", t)
    op.close(); t = tick("close", t)
    del out, local
if rank == 0:
    for k, v in acc.items():
        print("%-12s %8.3f ms" % (k, v / reps * 1e3))
    print("total        %8.3f ms" % (sum(acc.values()) / reps * 1e3), flush=True)
# ---- join build
od = tpch.orders(sf, seed=42 + rank, device="cuda")
od["orderkey"] += key_off
bpage = P.Page(Block(N.BIGINT, od["orderkey"], None, None, n_orders), Block(N.DOUBLE, od["totalprice"].view(torch.int64), None, None, n_orders),
               position_count=n_orders)
from presto_b200.join import PagesIndex
acc.clear()
for it in range(reps + 3):
    if it == 3:
        acc.clear()
    dist.barrier(); torch.cuda.synchronize()
    t = time.perf_counter()
    local = ex.exchange(bpage, [0], slot="build", lease=True); t = tick("b.exchange", t)
    idx = PagesIndex([P.BIGINT, P.DOUBLE], [0], ctx=ctx); t = tick("b.index", t)
    idx.add_page(local); t = tick("b.add_page", t)
    ls = idx.create_lookup_source(); t = tick("b.create_lookup", t)
    del ls, idx; t = tick("b.del", t)
if rank == 0:
    for k, v in acc.items():
        print("%-16s %8.3f ms" % (k, v / reps * 1e3))
    print("total            %8.3f ms" % (sum(acc.values()) / reps * 1e3), flush=True)
import cProfile, pstats
pr = cProfile.Profile()
dist.barrier(); torch.cuda.synchronize()
pr.enable()
for it in range(3):
    local = ex.exchange(bpage, [0], slot="build", lease=True)
    torch.cuda.synchronize()
pr.disable()
if rank == 0:
    pstats.Stats(pr).sort_stats("tottime").print_stats(12)
dist.destroy_process_group()
