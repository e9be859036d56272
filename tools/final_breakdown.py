"""Host-side breakdown of the FINAL stage of the N > 1 Q1 plan: one PARTIAL pass, then FINAL over `world` copies of its rows."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import presto_b200 as P
from presto_b200 import tpch
from presto_b200.blocks import Block
from presto_b200.plans import final_aggregators

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 1
world = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ctx = P.GpuContext(0)
cols = tpch.lineitem(sf, seed=42, device="cuda")
page = tpch.q1_page(cols)
aggs = tpch.q1_aggregators()
fin_aggs = final_aggregators(aggs, 2)
_op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.PARTIAL, aggs, ctx=ctx)
_op.add_input(page); _op.finish(); part = _op.get_output(); _op.close()
blocks = []
for b in part.blocks:
    hb = b.to_host()
    n = part.position_count
    if hb.offsets is not None:
        used = int(hb.offsets[-1])
        offs = np.concatenate([[0]] + [hb.offsets[1:] + r * used for r in range(world)]).astype(np.int32)
        vals = np.concatenate([np.asarray(hb.values).view(np.uint8)[:used]] * world)
    else:
        offs, vals = None, np.concatenate([hb.values] * world)
    nulls = None if hb.nulls is None else np.concatenate([hb.nulls] * world)
    blocks.append(Block(b.type, vals, nulls, offs, n * world))
fpage = P.Page(*blocks, position_count=part.position_count * world)
acc = {}
def tick(name, t0):
    torch.cuda.synchronize()
    acc[name] = acc.get(name, 0.0) + (time.perf_counter() - t0)
reps = 50
for it in range(reps + 3):
    if it == 3:
        acc.clear()
    t = time.perf_counter(); op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, fin_aggs, expected_groups=1000, ctx=ctx); tick("create", t)
    t = time.perf_counter(); op.add_input(fpage); tick("add_input", t)
    t = time.perf_counter(); op.finish(); out = op.get_output(); tick("get_output", t)
    t = time.perf_counter(); op.close(); tick("close", t)
for k, v in acc.items():
    print("%-12s %8.3f ms" % (k, v / reps * 1e3))
print("total        %8.3f ms" % (sum(acc.values()) / reps * 1e3))
ctx.set_timing(True)
ctx.kernel_report(reset=True)
for it in range(10):
    op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, fin_aggs, expected_groups=1000, ctx=ctx)
    op.add_input(fpage); op.finish(); out = op.get_output(); op.close()
rep = ctx.kernel_report(reset=True)
for k, v in sorted(rep.items(), key=lambda kv: -kv[1][0]):
    print("%-44s %8.4f ms/step  %5.1f launches/step" % (k, v[0] / 10, v[1] / 10))
ctx.set_timing(False)
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for it in range(50):
    op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, fin_aggs, expected_groups=1000, ctx=ctx)
    op.add_input(fpage); op.finish(); out = op.get_output(); op.close()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
