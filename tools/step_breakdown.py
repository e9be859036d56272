"""Host-side breakdown of one Q1 operator pass (create / add_input / get_output / close), wall clock with syncs."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import presto_b200 as P
from presto_b200 import tpch

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10
ctx = P.GpuContext(0)
cols = tpch.lineitem(sf, seed=42, device="cuda")
page = tpch.q1_page(cols)
aggs = tpch.q1_aggregators()
acc = {}
def tick(name, t0):
    torch.cuda.synchronize()
    acc[name] = acc.get(name, 0.0) + (time.perf_counter() - t0)
reps = 30
for it in range(reps + 3):
    if it == 3:
        acc.clear()
    t = time.perf_counter(); op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.SINGLE, aggs, expected_groups=100_000, ctx=ctx); tick("create", t)
    t = time.perf_counter(); op.add_input(page); tick("add_input", t)
    t = time.perf_counter(); op.finish(); out = op.get_output(); tick("get_output", t)
    t = time.perf_counter(); op.close(); tick("close", t)
for k, v in acc.items():
    print("%-12s %8.3f ms" % (k, v / reps * 1e3))
print("total        %8.3f ms" % (sum(acc.values()) / reps * 1e3))
ctx.set_timing(True)
ctx.kernel_report(reset=True)
for it in range(10):
    op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.SINGLE, aggs, expected_groups=100_000, ctx=ctx)
    op.add_input(page); op.finish(); out = op.get_output(); op.close()
rep = ctx.kernel_report(reset=True)
tot = 0
for k, v in sorted(rep.items(), key=lambda kv: -kv[1][0]):
    print("%-44s %8.4f ms/step  %5.1f launches/step" % (k, v[0] / 10, v[1] / 10)); tot += v[0] / 10
print("kernel total %.4f ms/step" % tot)
ctx.set_timing(False)
# add_input only, back to back without get_output (isolates the operator's own synchronisations)
torch.cuda.synchronize(); t = time.perf_counter()
for it in range(20):
    op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.SINGLE, aggs, expected_groups=100_000, ctx=ctx)
    op.add_input(page); op.close()
torch.cuda.synchronize(); print("create+add_input+close: %.3f ms" % ((time.perf_counter() - t) / 20 * 1e3))
