"""cfg3 (GROUP BY orderkey) and cfg4 (join build + probe) alone, with per-kernel times and launch counts.
usage: python tools/cfg34.py [sf] [gb|join|both] [reps] [shuffle]      (made for quick A/B runs and for ncu -k captures)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import presto_b200 as P
from presto_b200 import tpch
from presto_b200.blocks import Block

sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10
what = sys.argv[2] if len(sys.argv) > 2 else "both"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
shuffle = len(sys.argv) > 4 and sys.argv[4] == "shuffle"
N = P.native
ctx = P.GpuContext(0)
li = tpch.lineitem(sf, seed=42, device="cuda", need=("keys", "double"), shuffle=shuffle)
n = li["rows"]
n_orders = int(1_500_000 * sf)


def report(name, fn, rows):
    fn()
    torch.cuda.synchronize()
    ctx.set_timing(True)
    ctx.kernel_report(reset=True)
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / reps * 1e3
    rep = ctx.kernel_report(reset=True)
    ctx.set_timing(False)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    ms2 = (time.perf_counter() - t0) / reps * 1e3
    print("%s: %.3f ms/step untimed (%.3f with per-kernel events), %.2f G rows/s" % (name, ms2, ms, rows / ms2 / 1e6))
    tot = 0.0
    for k, v in sorted(rep.items(), key=lambda kv: -kv[1][0]):
        print("    %-40s %8.4f ms/step %5.1f launches/step" % (k, v[0] / reps, v[1] / reps))
        tot += v[0] / reps
    print("    kernel total %.4f ms/step" % tot)


if what in ("gb", "both"):
    page = P.Page(Block(N.BIGINT, li["orderkey"], None, None, n), Block(N.SHORT_DECIMAL, li["quantity"], None, None, n),
                  Block(N.DOUBLE, li["extendedprice_double"].view(torch.int64), None, None, n), position_count=n)
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 2)]

    def gb():
        op = P.HashAggregationOperator([P.BIGINT], [0], P.SINGLE, aggs, expected_groups=n_orders, ctx=ctx, device_output=True)
        op.add_input(page)
        op.finish()
        g = op.get_group_count()
        op.get_output(max_rows=max(g, 1))
        op.close()
    report("groupby_orderkey", gb, n)

if what in ("join", "both"):
    od = tpch.orders(sf, seed=42, device="cuda")
    bpage = P.Page(Block(N.BIGINT, od["orderkey"], None, None, n_orders), Block(N.DOUBLE, od["totalprice"].view(torch.int64), None, None, n_orders),
                   position_count=n_orders)
    keep = {}

    def build():
        idx = P.PagesIndex([P.BIGINT, P.DOUBLE], [0], ctx=ctx)
        idx.add_page(bpage)
        keep["ls"] = idx.create_lookup_source()
    report("join_build", build, n_orders)
    ls = keep["ls"]
    kb = Block(N.BIGINT, li["orderkey"], None, None, n)

    def probe():
        pp, bp = ls.join_page([kb], None, P.JOIN_INNER, device=True, capacity=n)
        ls.append_to(bp, [1], device=True)
    report("join_probe", probe, n)
