"""cProfile of the Q1 step's host side (fused processor, raw columns) + wall-clock per phase."""
import os, sys, time, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import presto_b200 as P
from presto_b200 import tpch

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10
ctx = P.GpuContext(0)
cols = tpch.lineitem(sf, seed=42, device="cuda", need=("raw",))
page = tpch.q1_raw_page(cols)
aggs = tpch.q1_fused_aggregators()
flt, proj = tpch.q1_expressions()
processor = P.PageProcessor(flt, proj, ctx=ctx)
plan = P.PartialFinalAggregation([P.VARCHAR, P.VARCHAR], [0, 1], aggs, ctx, None, expected_groups=100_000, input_processor=processor)
for _ in range(5):
    plan.submit(page)
acc = {}
def tick(name, t0):
    torch.cuda.synchronize()
    acc[name] = acc.get(name, 0.0) + (time.perf_counter() - t0)
reps = 50
for it in range(reps):
    t = time.perf_counter(); op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.SINGLE, aggs, expected_groups=100_000, ctx=ctx, input_processor=processor); tick("create", t)
    t = time.perf_counter(); op.add_input(page); tick("add_input", t)
    t = time.perf_counter(); op.finish(); out = op.get_output(); tick("get_output", t)
    t = time.perf_counter(); op.close(); tick("close", t)
for k, v in acc.items():
    print("%-12s %8.3f ms" % (k, v / reps * 1e3))
print("total        %8.3f ms" % (sum(acc.values()) / reps * 1e3))
torch.cuda.synchronize(); t = time.perf_counter()
for it in range(reps):
    plan.submit(page)
torch.cuda.synchronize(); print("plan.submit back to back: %.3f ms" % ((time.perf_counter() - t) / reps * 1e3))
pr = cProfile.Profile()
pr.enable()
for it in range(200):
    plan.submit(page)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
