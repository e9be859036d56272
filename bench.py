#!/usr/bin/env python
"""bench.py — input rows/s of the hash-aggregation / hash-join hot path (BASELINE.json metric) on N B200s of one node.

Headline (`value`, `e2e`, `roofline`): BASELINE.json configs[1], TPC-H Q1 over lineitem SF10 per GPU (weak scaling), from the raw
scan columns: filter shipdate <= DATE '1998-09-02', project extendedprice*(1-discount) and ...*(1+tax) as exact long decimals,
GROUP BY returnflag, linestatus (2 x varchar(1), 4 groups), sum/avg over short and long decimals and count(*).  The filter and
the projections are fused into the aggregation kernel (pg_hashagg_set_input_processor): derived columns never exist in memory.
  value : whole-job input rows/s with the columns already resident in HBM (one HashAggregationOperator pass per step)
  e2e   : the same pass through the C ABI with HOST (pinned) column buffers, host->device copies inside the timed region
  N > 1 : every rank aggregates its own SF10 shard (PARTIAL), the 4-row intermediate pages are all-gathered over NCCL and merged by a
          FINAL step on every rank (presto-1_b200/plans.py PartialFinalAggregation).
`workloads` (same JSON line): the other half of the metric at TPC-H SF100 *in total* (strong scaling: every rank holds 1/N of the rows,
disjoint order-key ranges): BASELINE.json configs[2] GROUP BY orderkey (150 M groups; dbgen order and a random permutation of the
same rows) and configs[3] lineitem JOIN orders (150 M-row build, 600 M-row probe + gather; dbgen order and shuffled).  At N > 1 the
rows cross the hash-partitioned exchange (presto-1_b200/exchange.py: fused partition + peer-memory stores over NVLink, NCCL
all-to-all as the alternative) inside the timed region.  Every workload checks its own full-size result and raises on a mismatch;
`crosscheck` compares the same plans at a small scale factor with the CPU oracle and, at N > 1, the peer exchange with the NCCL one.
`--impl reference` times the CPU restatement of the reference's operators (oracle/, kind "port": the reference is Java and no JDK
exists here) on all host threads over the same Q1 input.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS_PER_SF = 5_998_605.2
METRIC = "input rows/sec for GROUP BY & hash-join probe at 1/2/4/8 B200; % HBM roofline"

# Algorithmic bytes per row (DESIGN.md §3 states each term)
B_GROUPBY_ORDERKEY = 136      # K 8 + inputs 16 + slot 16 + 2 x state (8 + 24 + 16)
B_JOIN_BUILD = 32             # K 8 + payload 8 + slot 16
B_JOIN_PROBE = 44             # K 8 + slot 16 + payload gather 8 + output (probe idx 4 + payload 8)
# per-kernel models for `roofline.achieved` of each workload's dominant kernel (bytes per row that kernel must move)
KERNEL_BYTES = {
    "wide_locate_kernel": 8 + 16 + 4,            # key, slot, bucket number out
    "wide_update_jit": 4 + 16 + 2 * 48,          # bucket number in, two 8-byte arguments, state read-modify-write
    "wide_agg_jit": 136,                         # fused locate + update
    "j_build_fast_kernel": 8 + 16,               # key, slot write (the payload stays where addPage put it)
    "j_build_kernel": 8 + 16,
    "j_probe_dense_kernel": 8 + 16 + 4 + 8,      # key, slot, probe idx out, build position out
    "j_probe_dense_kernel<true>": 44,            # fused probe + gather (the whole model: pg_lookup_join_page)
    "j_gather_fixed_kernel": 8 + 8 + 8,          # build position in, payload gather, payload out
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sf", type=float, default=10.0, help="TPC-H scale factor of the Q1 lineitem shard per GPU")
    ap.add_argument("--wide-sf", type=float, default=100.0, help="TOTAL scale factor of the GROUP BY orderkey / join workloads")
    ap.add_argument("--wide-steps", type=int, default=5)
    ap.add_argument("--wide-warmup", type=int, default=3)
    ap.add_argument("--cpu-sample-rows", type=int, default=6_000_000)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-e2e-pages", action="store_true")
    ap.add_argument("--page-rows", type=int, default=8192, help="rows per page of the e2e_pages record (the reference's pages hold a few thousand rows)")
    ap.add_argument("--page-batch-rows", type=int, default=1 << 20)
    ap.add_argument("--page-drivers", type=int, default=16)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-workloads", action="store_true", help="headline only (skips the SF100 GROUP BY orderkey / join records)")
    ap.add_argument("--no-crosscheck", action="store_true", help="skip the small-scale oracle cross-checks (profiling runs)")
    ap.add_argument("--only", default="", help="comma list of workloads to run: groupby,join,q3,grid (default all)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clocks / throttle reasons during the timed region (B200_PROFILING.md recipe).  In-process NVML (nvidia_ml_py) when
    it is importable — a query takes microseconds and does not disturb the run; one `nvidia-smi` process per sample (tens of
    milliseconds of host work and driver locks each: measured to cost 25 % of the 8-GPU Q1 step when every rank did it) otherwise."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, device):
        self.device, self.samples, self._stop, self._t = device, [], threading.Event(), None
        self.source = None

    def _nvml_handle(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            idx = self.device
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                idx = int(vis.split(",")[self.device])
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(idx)
        except Exception:
            return None, None

    def _run(self):
        nv, h = self._nvml_handle()
        if nv is not None:
            self.source = "nvml"
            bits = [(0x8, 0), (0x40, 1), (0x20, 2), (0x4, 3)]      # hw_slowdown, hw_thermal, sw_thermal, sw_power_cap
            try:
                mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            except Exception:
                mx = None
            while not self._stop.is_set():
                try:
                    sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                    try:
                        r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                    except Exception:
                        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    row = [str(sm), str(mx)] + ["Active" if r & bit else "Not Active" for bit, _ in bits]
                    self.samples.append(row)
                except Exception:
                    pass
                self._stop.wait(0.005)
            return
        self.source = "nvidia-smi"
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        if self.device is not None:     # None: this rank does not sample (rank 0 prints the line; one sampler per node keeps the host quiet)
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no clock samples (nvml / nvidia-smi unavailable)"]}
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        reasons = [n for i, n in enumerate(self.NAMES) if any(len(s) > 2 + i and s[2 + i].lower().startswith("active") for s in self.samples)]
        mx = [int(s[1]) for s in self.samples if s[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples), "source": self.source}


def q1_config(sf, rows):
    """`config` of the headline, identical in both arms."""
    return {"workload": "TPC-H Q1 on lineitem SF%g per GPU, from the raw columns (returnflag, linestatus, quantity, extendedprice, discount, tax, "
                        "shipdate): filter shipdate <= 1998-09-02; project extendedprice*(1-discount), extendedprice*(1-discount)*(1+tax) as exact "
                        "decimals; GROUP BY returnflag,linestatus; 4x sum, 3x avg (decimal), count(*)" % sf,
            "rows_per_gpu": int(rows), "groups": 4,
            "l2_policy": "inputs (%.2f GB per pass) exceed the 126 MB L2; no flush needed" % (rows * 46 / 1e9)}


def result_digest(out):
    """sha256 over the output rows sorted by key: printed by both arms, equal when they computed the same answer."""
    import hashlib
    hb = [b.to_host() for b in out.blocks]
    rows = sorted(tuple(str(b.get(i)) for b in hb) for i in range(out.position_count))
    return hashlib.sha256(json.dumps(rows).encode()).hexdigest()[:16]


# ------------------------------------------------------------------------------------------- CPU arm (oracle port)
def slice_host_block(P, b, lo, hi):
    """Rows [lo, hi) of a host block as a view (no copy)."""
    from presto_b200.blocks import Block
    if b.type in (P.VARCHAR, P.VARBINARY):
        out = Block(b.type, b.values, None if b.nulls is None else b.nulls[lo:hi], b.offsets[lo:hi + 1], hi - lo)
    else:
        k = 2 if b.type == P.LONG_DECIMAL else 1
        out = Block(b.type, b.values[k * lo:k * hi], None if b.nulls is None else b.nulls[lo:hi], None, hi - lo, b.precision, b.scale)
    out.pinned = getattr(b, "pinned", False)
    return out


class PagedDrivers:
    """The reference's shape of a scan stage: T driver threads, each with its own PARTIAL operator (and context / stream), fed
    reference-sized host pages one addInput at a time; one FINAL operator merges the drivers' intermediate pages.  The page loop
    runs in C (pg_hashagg_add_pages: each page goes through pg_hashagg_add_page's path) so that Python's per-call cost is not what
    is measured; the pinned page accumulator batches the pages (pg_hashagg_set_page_batching)."""

    def __init__(self, P, tpch, torch, device, hpage, nthreads, page_rows, batch_rows, zero_copy=True):
        from presto_b200.context import GpuContext
        from presto_b200.plans import final_aggregators
        self.P, self.torch, self.page_rows, self.batch_rows, self.T = P, torch, page_rows, batch_rows, nthreads
        self.aggs = tpch.q1_fused_aggregators()
        self.fin_aggs = final_aggregators(self.aggs, 2)
        flt, proj = tpch.q1_expressions()
        n = hpage.position_count
        self.slices, self.ctxs, self.procs, self.streams, self.bufs = [], [], [], [], []
        for t in range(nthreads):
            lo, hi = n * t // nthreads, n * (t + 1) // nthreads
            self.slices.append(P.Page(*[slice_host_block(P, b, lo, hi) for b in hpage.blocks], position_count=hi - lo))
            if not zero_copy:       # plain host pages (what a JNI caller has): every page is copied into the pinned accumulator first
                for b in self.slices[-1].blocks:
                    b.pinned = False
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                ctx = GpuContext(device)
            self.streams.append(st)
            self.ctxs.append(ctx)
            self.procs.append(P.PageProcessor(flt, proj, ctx=ctx))
            self.bufs.append(torch.empty(1 << 16, dtype=torch.uint8).pin_memory())
        self.main_ctx = GpuContext(device)
        self.err = None

    def _driver(self, t):
        P = self.P
        try:
            self.torch.cuda.set_device(self.ctxs[t].device)
            op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.PARTIAL, self.aggs, expected_groups=100_000, ctx=self.ctxs[t],
                                           input_processor=self.procs[t])
            op.set_page_batching(self.batch_rows)
            op.add_pages(self.slices[t], self.page_rows)
            op.finish()
            op.get_output_packed(self.bufs[t].data_ptr(), self.bufs[t].numel())
            op.close()
        except BaseException as e:      # surfaced by step()
            self.err = e

    def step(self):
        P = self.P
        threads = [threading.Thread(target=self._driver, args=(t,)) for t in range(self.T)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        if self.err is not None:
            raise self.err
        fin = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, self.fin_aggs, expected_groups=1000, ctx=self.main_ctx)
        fin.add_packed([b.data_ptr() for b in self.bufs])
        fin.finish()
        out = fin.get_output()
        fin.close()
        return out

    def launch_count(self):
        return sum(c.launch_count() for c in self.ctxs) + self.main_ctx.launch_count()


def oracle_q1_pass(O, P, blocks, aggs, nthreads, exprs=None):
    """The reference's plan on host cores: `nthreads` drivers each run ScanFilterAndProject (filter + projections), then getGroupIds
    + accumulators (PARTIAL) over a slice of the pages, then one FINAL merge — restated in oracle/ (C++), called through ctypes
    (GIL released)."""
    n = blocks[0].n
    bounds = [n * i // nthreads for i in range(nthreads + 1)]
    results = [None] * nthreads

    def slice_block(b, lo, hi):
        from presto_b200.blocks import Block
        if b.type in (P.VARCHAR, P.VARBINARY):
            return Block(b.type, b.values, None, b.offsets[lo:hi + 1], hi - lo)
        k = 2 if b.type == P.LONG_DECIMAL else 1
        return Block(b.type, b.values[k * lo:k * hi], None if b.nulls is None else b.nulls[lo:hi], None, hi - lo)

    def work(t):
        lo, hi = bounds[t], bounds[t + 1]
        sl = [slice_block(b, lo, hi) for b in blocks]
        if exprs is not None:
            rc, projected = O.OraclePageProcessor(exprs[0], exprs[1]).process(sl)
            assert rc == 0
            sl = projected.blocks
        gb = O.OracleGroupBy([P.VARCHAR, P.VARCHAR], 100_000)
        ids = gb.get_group_ids(sl[:2])
        g = gb.group_count()
        inter = []
        for a in aggs:
            acc = O.OracleAccumulator(a.kind, a.in_type)
            acc.add_input(ids, g, sl[a.input_channel] if a.input_channel >= 0 else None, sl[a.mask_channel] if a.mask_channel >= 0 else None)
            inter.append(acc.evaluate(0, g, False)[1])
        results[t] = (gb.export(with_hash=False)[0], inter)

    threads = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    fgb = O.OracleGroupBy([P.VARCHAR, P.VARCHAR], 100_000)
    finals = [O.OracleAccumulator(a.kind, a.in_type) for a in aggs]
    for keys, inter in results:
        ids = fgb.get_group_ids(keys)
        for f, st in zip(finals, inter):
            f.add_intermediate(ids, fgb.group_count(), st)
    g = fgb.group_count()
    out = [f.evaluate(0, g, True)[1][0] for f in finals]
    return fgb.export(with_hash=False)[0], out


def host_q1_blocks(tpch, sf, seed):
    cols = tpch.lineitem(sf, seed=seed, device="cpu", need=("raw",))
    page = tpch.q1_raw_page(cols)
    return tpch.page_to_host(page, pinned=False).blocks


def time_cpu_port(P, tpch, sf, nthreads, steps, warmup, seed=42):
    """Returns (rows, per-step seconds, digest of the result)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_lib as O
    blocks = host_q1_blocks(tpch, sf, seed)
    n = blocks[0].n
    aggs = tpch.q1_fused_aggregators()
    exprs = tpch.q1_expressions()
    times = []
    res = None
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        res = oracle_q1_pass(O, P, blocks, aggs, nthreads, exprs)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    keys, outs = res
    return n, times, result_digest(P.Page(*(list(keys) + list(outs))))


# ------------------------------------------------------------------------------------------- GPU arm helpers
class Env:
    """What every workload needs: package, torch, process group, context, measured peak."""

    def __init__(self, P, tpch, torch, dist, ctx, world, rank, peak):
        self.P, self.tpch, self.torch, self.dist, self.ctx, self.world, self.rank, self.peak = P, tpch, torch, dist, ctx, world, rank, peak

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        if self.dist is None:
            return ms
        t = self.torch.tensor([ms], device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def total(self, x):
        """Sum of an int64 (wrapping) over the ranks."""
        if self.dist is None:
            return int(x)
        t = self.torch.tensor([int(x)], device="cuda", dtype=self.torch.int64)
        self.dist.all_reduce(t)
        return int(t.item())

    def timed(self, fn, steps, warmup, drain=None):
        """W untimed + K timed steps bracketed by barrier + synchronize; CUDA events on the launching stream; max over ranks.
        Returns (ms per step, last result, per-kernel report of the timed region, launches in the timed region)."""
        torch, ctx = self.torch, self.ctx
        for _ in range(warmup):
            fn()
        if drain:
            drain()
        self.barrier()
        ctx.kernel_report(reset=True)
        l0 = ctx.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = None
        trace = os.environ.get("PRESTOGPU_BENCH_TRACE") == "1"
        t_prev = time.perf_counter()
        for _ in range(steps):
            out = fn()
            if trace:
                t_now = time.perf_counter()
                print("[trace] rank %d step wall %.3f ms" % (self.rank, (t_now - t_prev) * 1e3), file=sys.stderr, flush=True)
                t_prev = t_now
        if drain:
            out = drain()
        e1.record()
        self.barrier()
        ms = self.max_over_ranks(e0.elapsed_time(e1)) / steps
        return ms, out, ctx.kernel_report(reset=True), ctx.launch_count() - l0


def workload_roofline(env, kern, steps, rows_per_gpu, traffic, name=None):
    """roofline record of a workload's dominant kernel, from the library's per-launch CUDA events inside the timed region."""
    known = {k: v for k, v in kern.items() if k in KERNEL_BYTES}
    pool = known or kern
    if not pool:
        return None
    top = max(pool, key=lambda k: pool[k][0])
    tms, cnt = kern[top]
    launch_ms = tms / max(cnt, 1)
    rows_per_launch = rows_per_gpu * steps / max(cnt, 1)
    bpr = KERNEL_BYTES.get(top)
    rec = {"bound": "hbm", "kernel": top, "launch_ms": round(launch_ms, 4), "launches_timed": cnt, "peak": env.peak, "unit": "GB/s",
           "kernels_ms_per_step": {k: round(v[0] / steps, 3) for k, v in sorted(kern.items(), key=lambda kv: -kv[1][0])[:8]}}
    if bpr is not None and launch_ms > 0:
        ach = rows_per_launch * bpr / (launch_ms / 1e3) / 1e9
        rec.update({"algorithmic_bytes_per_row": bpr, "rows_per_launch": int(rows_per_launch), "achieved": round(ach, 1),
                    "frac": round(ach / env.peak, 4)})
    t = None
    if traffic and env.world == 1 and rows_per_gpu > 140_000_000:      # the captures are one-GPU SF100 launches: per-launch traffic at other sizes is not comparable
        t = traffic.get("%s:%s" % (name, top), traffic.get(top + "@sf100"))
    rec["traffic"] = t
    if t is not None:
        rec["traffic_unit"] = "bytes/launch (ncu dram read+write, SF100 capture on 1 GPU)"
    return rec


def record(env, name, desc, rows_all, ms, brow, kern, steps, rows_per_gpu, traffic, extra=None):
    r = {"workload": desc, "rows": int(rows_all), "ms_per_step": round(ms, 3), "rows_per_s": rows_all / (ms / 1e3),
         "algorithmic_bytes_per_row": brow, "hbm_frac_per_gpu": round(rows_all / env.world * brow / (ms / 1e3) / 1e9 / env.peak, 4),
         "steps": steps, "roofline": workload_roofline(env, kern, steps, rows_per_gpu, traffic, name)}
    if extra:
        r.update(extra)
    return r


# ------------------------------------------------------------------------------------------- SF100 workloads (cfg3, cfg4)
def wide_workloads(args, env, traffic):
    """BASELINE.json configs[2] and [3] at SF100 in total; every rank holds 1/world of the orders (a contiguous key range) and their
    lineitems.  Inputs are device-resident; at world > 1 the hash-partitioned exchange is inside the timed region."""
    P, tpch, torch, ctx, world, rank = env.P, env.tpch, env.torch, env.ctx, env.world, env.rank
    from presto_b200.blocks import Block
    N = P.native
    only = set(x for x in args.only.split(",") if x)
    sf = args.wide_sf / world
    steps, warmup = args.wide_steps, args.wide_warmup
    res = {"scale": "TPC-H SF%g in total, SF%g per GPU (strong scaling: total rows fixed as N grows)" % (args.wide_sf, sf),
           "timing": "%d timed steps after %d warm-up steps per workload; inputs (>= 1.2 GB per column) exceed the 126 MB L2" % (steps, warmup)}
    n_orders = int(1_500_000 * sf)
    li = tpch.lineitem(sf, seed=42 + rank, device="cuda", need=("keys", "double"))
    n = li["rows"]
    key_off = rank * tpch.order_key(torch.tensor(n_orders, dtype=torch.int64)).item()   # disjoint order-key ranges per rank
    li["orderkey"] += key_off
    n_all, o_all = env.total(n), env.total(n_orders)
    exchange = P.HashPartitionExchange(P.DevicePartitioner(ctx)) if world > 1 else None
    how = "; rows hash-partitioned on orderkey across %d GPUs (%s), then one operator per GPU" % (
        world, "fused partition + peer-memory stores over NVLink" if exchange.peer is not None else "NCCL all_to_all_single per column: " + exchange.peer_why) if exchange else ""
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 2)]

    def li_page(c):
        return P.Page(Block(N.BIGINT, c["orderkey"], None, None, n), Block(N.SHORT_DECIMAL, c["quantity"], None, None, n),
                      Block(N.DOUBLE, c["extendedprice_double"].view(torch.int64), None, None, n), position_count=n)

    perm = torch.randperm(n, device="cuda", generator=torch.Generator(device="cuda").manual_seed(7 + rank))
    lis = {k: li[k][perm] for k in ("orderkey", "quantity", "extendedprice_double")}
    del perm
    ctx.set_timing(True)

    # ---------------- cfg3: GROUP BY orderkey
    if not only or "groupby" in only:
        plan = P.PartitionedAggregation([P.BIGINT], [0], aggs, ctx, env.dist, expected_groups=n_orders, exchange=exchange)
        sorted_result = None
        for order, cols in (("dbgen order (clustered by orderkey)", li), ("random permutation of the same rows", lis)):
            page = li_page(cols)
            ms, (groups, out), kern, _ = env.timed(lambda: plan.run(page), steps, warmup)
            g_all = env.total(groups)
            keys, cnt = out.blocks[0].values, out.blocks[1].values
            qty = out.blocks[2].values.view(-1, 2)[:, 0]           # long decimal low words (sums are far below 2^63)
            check = {"groups_equal_orders": g_all == o_all,
                     "count_sum_equals_rows": env.total(cnt.sum().item()) == n_all,
                     "decimal_sum_equals_column_sum": env.total(qty.sum().item()) == env.total(li["quantity"].sum().item()),
                     "key_times_count_checksum": env.total((keys * cnt).sum().item()) == env.total(li["orderkey"].sum().item())}
            if world == 1:
                if sorted_result is None:
                    check["keys_in_first_seen_order"] = bool(torch.equal(keys, torch.unique_consecutive(li["orderkey"])))
                    sorted_result = (keys.clone(), cnt.clone(), qty.clone(), out.blocks[3].values.view(torch.float64).clone())
                else:
                    o = torch.argsort(keys)
                    sk, sc, sq, sd = sorted_result
                    dbl = out.blocks[3].values.view(torch.float64)[o]
                    check["equals_dbgen_order_result_per_group"] = bool(torch.equal(keys[o], sk) and torch.equal(cnt[o], sc) and torch.equal(qty[o], sq))
                    check["double_sums_within_1e-12"] = bool(((dbl - sd).abs() <= 1e-12 * sd.abs()).all().item())
                    del o, dbl
            if not all(check.values()):
                raise SystemExit("bench self-check failed (GROUP BY orderkey, %s): %s" % (order, json.dumps(check)))
            brow = B_GROUPBY_ORDERKEY + (48 if world > 1 else 0)     # + 2 x 24 B partition pass (read + write of key and inputs)
            name = "groupby_orderkey" if cols is li else "groupby_orderkey_shuffled"
            res[name] = record(env, name, "GROUP BY orderkey over lineitem, %s: count(*), sum(quantity decimal), sum(extendedprice double)%s" % (order, how),
                               n_all, ms, brow, kern, steps, n, traffic,
                               {"groups": g_all, "selfcheck": check,
                                "hbm_frac_without_partition_pass": round(n_all / world * B_GROUPBY_ORDERKEY / (ms / 1e3) / 1e9 / env.peak, 4)})
            del out, keys, cnt, qty, page
        del sorted_result, plan
        torch.cuda.empty_cache()

    # ---------------- cfg4: build orders, probe lineitem.orderkey, gather totalprice
    if not only or "join" in only:
        od = tpch.orders(sf, seed=42 + rank, device="cuda")
        od["orderkey"] += key_off
        bpage = P.Page(Block(N.BIGINT, od["orderkey"], None, None, n_orders), Block(N.DOUBLE, od["totalprice"].view(torch.int64), None, None, n_orders),
                       position_count=n_orders)
        join = P.PartitionedHashJoin([P.BIGINT, P.DOUBLE], [0], ctx, env.dist, exchange=exchange)
        bms, ls, bkern, _ = env.timed(lambda: join.build(bpage), steps, warmup)
        res["join_build"] = record(env, "join_build", "hash build orders (orderkey unique, totalprice)" + how, o_all, bms,
                                   B_JOIN_BUILD + (32 if world > 1 else 0), bkern, steps, n_orders, traffic)
        tp_bits = od["totalprice"].view(torch.int64)
        for order, cols in (("dbgen order", li), ("random permutation of the same rows", lis)):
            ppage = P.Page(Block(N.BIGINT, cols["orderkey"], None, None, n), Block(N.SHORT_DECIMAL, cols["quantity"], None, None, n), position_count=n)
            pms, (local, pp, got), pkern, _ = env.timed(lambda: join.probe(ppage, [0], [1]), steps, warmup)
            want = tp_bits[torch.searchsorted(od["orderkey"], cols["orderkey"])]      # every lineitem's order lives on the rank that generated it
            gv = got[0].values.view(torch.int64)
            check = {"matches_equal_probe_rows": env.total(len(pp)) == n_all,
                     "gathered_payload_checksum": env.total(gv.sum().item()) == env.total(want.sum().item())}
            if world == 1:
                check["probe_positions_ascending"] = bool(torch.equal(pp, torch.arange(n, device="cuda", dtype=torch.int32)))
                check["gathered_payload_bit_exact"] = bool(torch.equal(gv, want))
            if not all(check.values()):
                raise SystemExit("bench self-check failed (join probe, %s): %s" % (order, json.dumps(check)))
            name = "join_probe" if cols is li else "join_probe_shuffled"
            res[name] = record(env, name, "INNER join probe lineitem.orderkey -> orders, %s; emits probe positions + gathered totalprice%s" % (order, how),
                               n_all, pms, B_JOIN_PROBE + (32 if world > 1 else 0), pkern, steps, n, traffic,
                               {"matches": n_all, "selfcheck": check})
            del want, gv, got, pp, local, ppage
        del od, bpage, join, ls
    if not only or "q3" in only:
        del li, lis
        torch.cuda.empty_cache()
        res["q3_join_aggregation"] = q3_shape(args, env, exchange)
    ctx.set_timing(False)
    if world == 1 and (not only or "grid" in only):
        res["groupby_hash_grid"] = groupby_hash_grid(args, env)
    if exchange:
        res["exchange"] = {"rows_sent_per_rank": exchange.sent_rows, "bytes_sent_per_rank": exchange.bytes_sent,
                           "collective": ("fused partition + peer-memory stores over NVLink (pg_partition_scatter_peers into symmetric receive buffers)"
                                          if exchange.peer is not None else "NCCL all_to_all_single per column (%s)" % exchange.peer_why)}
    return res


# ------------------------------------------------------------------------------------------- cfg5: TPC-H Q3 shape
def q3_shape(args, env, exchange):
    """BASELINE.json configs[4]: customer (mktsegment = 'BUILDING') JOIN orders (orderdate < 1995-03-15) ON custkey JOIN lineitem
    (shipdate > 1995-03-15) ON orderkey, GROUP BY (l_orderkey, o_orderdate, o_shippriority), sum(extendedprice * (1 - discount));
    1 % NULL o_custkey and l_orderkey.  Filters and the revenue projection are PageProcessors, both joins are
    PartitionedHashJoin (hash exchange on the join key at N > 1), the aggregation is the general (multi-channel) path on each rank's
    join output — its group key starts with the exchanged orderkey, so no further exchange is needed.  SF = --wide-sf in total."""
    P, tpch, torch, ctx, world, rank = env.P, env.tpch, env.torch, env.ctx, env.world, env.rank
    from presto_b200.blocks import Block
    from presto_b200.expressions import Expr as E, PageProcessor
    from presto_b200.operators import take_positions
    N = P.native
    sf = args.wide_sf / world
    cutoff = 9204                                                 # 1995-03-15
    cu = tpch.customer(sf, seed=42 + rank, device="cuda")
    od = tpch.orders(sf, seed=42 + rank, device="cuda")
    li = tpch.lineitem(sf, seed=42 + rank, device="cuda", need=("keys", "raw"))
    n_c, n_o, n_l = cu["rows"], int(od["orderkey"].numel()), li["rows"]
    okey_off = rank * tpch.order_key(torch.tensor(n_o, dtype=torch.int64)).item()
    ckey_off = rank * n_c                                         # disjoint customer / order key ranges per rank
    od["orderkey"] += okey_off; li["orderkey"] += okey_off
    cu["custkey"] += ckey_off; od["custkey"] = (od["custkey"] - 1) % n_c + 1 + ckey_off
    o_null = tpch.null_mask(n_o, 43 + rank, 1, 100, "cuda")
    l_null = tpch.null_mask(n_l, 43 + rank, 2, 100, "cuda")
    cpage = P.Page(Block(N.BIGINT, cu["custkey"], None, None, n_c), Block(N.VARCHAR, cu["mktsegment_bytes"], None, cu["mktsegment_offsets"], n_c), position_count=n_c)
    opage = P.Page(Block(N.BIGINT, od["custkey"], o_null, None, n_o), Block(N.BIGINT, od["orderkey"], None, None, n_o),
                   Block(N.DATE, od["orderdate"], None, None, n_o), Block(N.INTEGER, od["shippriority"], None, None, n_o), position_count=n_o)
    lpage = P.Page(Block(N.BIGINT, li["orderkey"], l_null, None, n_l), Block(N.SHORT_DECIMAL, li["extendedprice"], None, None, n_l, 12, 2),
                   Block(N.SHORT_DECIMAL, li["discount"], None, None, n_l, 12, 2), Block(N.DATE, li["shipdate"], None, None, n_l), position_count=n_l)
    seg_filter = PageProcessor(E.input(1, N.VARCHAR).eq(E.varchar("BUILDING")), [E.input(0, N.BIGINT)], ctx=ctx)
    ord_filter = PageProcessor(E.input(2, N.DATE) < E.date(cutoff), [E.input(0, N.BIGINT), E.input(1, N.BIGINT), E.input(2, N.DATE), E.input(3, N.INTEGER)], ctx=ctx)
    li_filter = PageProcessor(E.input(3, N.DATE) > E.date(cutoff), [E.input(0, N.BIGINT), E.input(1, N.SHORT_DECIMAL, 12, 2), E.input(2, N.SHORT_DECIMAL, 12, 2)], ctx=ctx)
    # the revenue expression sits above the joins, as in the reference's plan (a ProjectNode right below the aggregation): it is the
    # aggregation's input processor and is evaluated for the joined rows only
    revenue = PageProcessor(None, [E.input(0, N.BIGINT), E.input(1, N.DATE), E.input(2, N.INTEGER),
                                   E.input(3, N.SHORT_DECIMAL, 12, 2) * (E.decimal(1, 1, 0) - E.input(4, N.SHORT_DECIMAL, 12, 2))], ctx=ctx)
    aggs = [P.AggregatorFactory(P.AGG_SUM, P.LONG_DECIMAL, 3)]
    join1 = P.PartitionedHashJoin([P.BIGINT], [0], ctx, env.dist, exchange=exchange)
    join2 = P.PartitionedHashJoin([P.BIGINT, P.DATE, P.INTEGER], [0], ctx, env.dist, exchange=exchange)

    stages = None                                                                             # filled by the untimed breakdown pass below

    def mark(name, t0):
        if stages is None:
            return t0
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        stages[name] = round(stages.get(name, 0.0) + (t1 - t0) * 1e3, 3)
        return t1

    def step():
        t = time.perf_counter()
        building = seg_filter.process(cpage, device=True)                                     # custkey of the segment's customers
        orders_f = ord_filter.process(opage, device=True)                                     # custkey, orderkey, orderdate, shippriority
        t = mark("filter_customer_orders", t)
        join1.build(building, slot="q3_b1")
        t = mark("join1_build", t)
        local_o, pp, _ = join1.probe(orders_f, [0], [], slot="q3_p1")
        surviving = P.Page(*[take_positions(local_o.blocks[c], pp) for c in (1, 2, 3)], position_count=len(pp))
        t = mark("join1_probe", t)
        join2.build(surviving, slot="q3_b2")
        t = mark("join2_build", t)
        items = li_filter.process(lpage, device=True)                                         # orderkey, extendedprice, discount
        t = mark("filter_lineitem", t)
        local_l, pp2, gathered = join2.probe(items, [0], [1, 2], slot="q3_p2")
        joined = P.Page(take_positions(local_l.blocks[0], pp2), gathered[0], gathered[1], take_positions(local_l.blocks[1], pp2),
                        take_positions(local_l.blocks[2], pp2), position_count=len(pp2))
        for b in joined.blocks[3:]:
            b.precision, b.scale = 12, 2
        t = mark("join2_probe", t)
        op = P.HashAggregationOperator([P.BIGINT, P.DATE, P.INTEGER], [0, 1, 2], P.SINGLE, aggs, expected_groups=max(len(pp), 1000), ctx=ctx, device_output=True,
                                       input_processor=revenue)
        op.add_input(joined)
        op.finish()
        g = op.get_group_count()
        out = op.get_output(max_rows=max(g, 1))
        op.close()
        mark("aggregation", t)
        return g, out, len(pp), len(pp2), joined

    ms, (groups, out, n_surv, n_joined, joined), kern, _ = env.timed(step, args.wide_steps, args.wide_warmup)
    stages = {}                                                   # one more pass, untimed, synchronised after every stage
    torch.cuda.synchronize()
    if os.environ.get("PRESTOGPU_Q3_CPROFILE") == "1":
        import cProfile, pstats
        prof = cProfile.Profile()
        prof.runcall(step)
        pstats.Stats(prof, stream=sys.stderr).sort_stats("tottime").print_stats(22)
    else:
        step()
    stage_ms, stages = stages, None
    ctx.kernel_report(reset=True)
    # self-check with torch on the same columns (independent of the library): which orders survive, which lineitems join, what they add up to
    if world == 1:
        cust_ok = torch.zeros(n_c + 2, dtype=torch.bool, device="cuda")
        cust_ok[cu["custkey"][cu["segment_code"] == 1]] = True
        o_ok = (od["orderdate"] < cutoff) & (o_null == 0) & cust_ok[od["custkey"].clamp(0, n_c + 1)]
        l_ok = (li["shipdate"] > cutoff) & (l_null == 0)
        order_pos = torch.searchsorted(od["orderkey"], li["orderkey"]).clamp(max=n_o - 1)
        l_join = l_ok & o_ok[order_pos] & (od["orderkey"][order_pos] == li["orderkey"])
        rev = (li["extendedprice"] * (100 - li["discount"]))[l_join]
        want_groups = int(torch.unique(li["orderkey"][l_join]).numel())
        got_rev = out.blocks[3].values.view(-1, 2)[:, 0]              # sums are far below 2^63
        check = {"surviving_orders": n_surv == int(o_ok.sum().item()), "joined_lineitems": n_joined == int(l_join.sum().item()),
                 "groups": groups == want_groups, "revenue_sum_bit_exact": int(got_rev.sum().item()) == int(rev.sum().item())}
    else:
        check = {"groups_positive": env.total(groups) > 0}
    if not all(check.values()):
        raise SystemExit("bench self-check failed (Q3 shape): %s" % json.dumps(check))
    rows_in = env.total(n_c + n_o + n_l)
    return {"workload": "TPC-H Q3 shape (cfg5): customer[mktsegment='BUILDING'] JOIN orders[orderdate<1995-03-15] JOIN lineitem[shipdate>1995-03-15], "
                        "GROUP BY (l_orderkey, o_orderdate, o_shippriority) sum(extendedprice*(1-discount)); 1 % NULL join keys; filters, two hash "
                        "joins, composite-key aggregation", "input_rows": int(rows_in), "ms_per_step": round(ms, 3), "rows_per_s": rows_in / (ms / 1e3),
            "surviving_orders": env.total(n_surv), "joined_lineitems": env.total(n_joined), "groups": env.total(groups), "steps": args.wide_steps,
            "kernels_ms_per_step": {k: round(v[0] / args.wide_steps, 3) for k, v in sorted(kern.items(), key=lambda kv: -kv[1][0])[:10]},
            "stages_ms_untimed_pass": stage_ms, "selfcheck": check}


# ------------------------------------------------------------------------------------------- BenchmarkGroupByHash grid
def groupby_hash_grid(args, env):
    """OPT/BenchmarkGroupByHash.java:71-354 (POSITIONS = 10 M, GROUP_COUNT = 3 M, EXPECTED_SIZE = 10 000: the table grows nine times
    inside the measurement, as in the reference) through the interface-level seam: GroupByHash.addPage and getGroupIds over a
    device-resident page, BIGINT and VARCHAR channels, with and without a precomputed hash channel.  One rank (the grid does not shard)."""
    P, torch, ctx = env.P, env.torch, env.ctx
    from presto_b200.blocks import Block
    from presto_b200.expressions import Expr, PageProcessor
    N = P.native
    positions, groups, expected = 10_000_000, 3_000_000, 10_000
    g = torch.Generator(device="cuda").manual_seed(99)
    v = torch.randint(0, groups, (positions,), device="cuda", generator=g)
    distinct = int(torch.unique(v).numel())
    vb = v.to(torch.int32).contiguous().view(torch.uint8)               # 4-byte little-endian strings, as createVarcharPages builds them
    offs = torch.arange(0, 4 * positions + 1, 4, dtype=torch.int32, device="cuda")
    out = {"what": "OPT/BenchmarkGroupByHash.java: %d positions, %d distinct keys (nextInt(%d)), expectedSize %d; one device page; best of 3 after 2 warm-up runs"
                   % (positions, distinct, groups, expected), "configs": []}
    for dtype, channels, hashed in (("BIGINT", 1, False), ("BIGINT", 1, True), ("BIGINT", 5, True), ("BIGINT", 10, False), ("VARCHAR", 1, True), ("VARCHAR", 5, False)):
        if dtype == "BIGINT":
            blocks = [Block(N.BIGINT, v, None, None, positions) for _ in range(channels)]
            types = [N.BIGINT] * channels
        else:
            blocks = [Block(N.VARCHAR, vb, None, offs, positions) for _ in range(channels)]
            types = [N.VARCHAR] * channels
        if hashed:
            proc = PageProcessor(None, [Expr.hash_of(*[Expr.input(c, types[c]) for c in range(channels)])], ctx=ctx)
            blocks = blocks + [proc.process(P.Page(*blocks, position_count=positions), device=True).blocks[0]]
        page = P.Page(*blocks, position_count=positions)
        rec = {"data_type": dtype, "channels": channels, "hash_channel": hashed,
               "class": "BigintGroupByHash" if (dtype == "BIGINT" and channels == 1) else "MultiChannelGroupByHash"}
        for what in ("add_page", "get_group_ids"):
            best = None
            for it in range(5):
                gb = P.GroupByHash(types, expected, has_hash_channel=hashed, ctx=ctx)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                ids = gb.add_page(page) if what == "add_page" else gb.get_group_ids(page, device_ids=True)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1)
                if it >= 2:
                    best = ms if best is None else min(best, ms)
                ok = gb.get_group_count() == distinct
                if what == "get_group_ids":
                    ok = ok and int(ids.ids.max().item()) == distinct - 1
                gb.close()
                if not ok:
                    raise SystemExit("bench self-check failed (GroupByHash grid %s): group count" % json.dumps(rec))
            rec[what + "_ms"] = round(best, 3)
            rec[what + "_rows_per_s"] = positions / (best / 1e3)
        out["configs"].append(rec)
    return out


# ------------------------------------------------------------------------------------------- cross-checks (small scale factor)
def crosschecks(env):
    """The same plans at a small scale factor against the CPU oracle (order-insensitive where the exchange reorders rows), and at
    N > 1 the peer-memory exchange against the NCCL all-to-all, bit for bit.  Raises on a mismatch.  The oracle is the checker
    here, never the thing measured."""
    import numpy as np
    P, tpch, torch, ctx, world, rank, dist = env.P, env.tpch, env.torch, env.ctx, env.world, env.rank, env.dist
    from presto_b200.blocks import Block
    N = P.native
    sf = 0.02
    n_orders = int(1_500_000 * sf)
    out = {}

    def shard(r, device):
        li = tpch.lineitem(sf, seed=1000 + r, device=device, need=("keys", "double"), shuffle=(r % 2 == 1))
        off = r * tpch.order_key(torch.tensor(n_orders, dtype=torch.int64)).item()
        li["orderkey"] = li["orderkey"] + off
        od = tpch.orders(sf, seed=1000 + r, device=device)
        od["orderkey"] = od["orderkey"] + off
        return li, od

    def wide_decimals(quantity, keys):
        """A decimal(38) column with non-zero high words and both signs: +-quantity x (10^20 + key), sign by row parity."""
        q, k = quantity.cpu().tolist(), keys.cpu().tolist()
        return P.Block.int128_array([(qi * (10**20 + ki)) * (1 if i % 2 else -1) for i, (qi, ki) in enumerate(zip(q, k))])

    li, od = shard(rank, "cuda")
    n = li["rows"]
    page = P.Page(Block(N.BIGINT, li["orderkey"], None, None, n), Block(N.SHORT_DECIMAL, li["quantity"], None, None, n),
                  Block(N.DOUBLE, li["extendedprice_double"].view(torch.int64), None, None, n), wide_decimals(li["quantity"], li["orderkey"]).to_device(),
                  position_count=n)
    exchange = P.HashPartitionExchange(P.DevicePartitioner(ctx)) if world > 1 else None
    if exchange is not None:
        # peer-memory exchange against the NCCL all-to-all on the same page, twice (the second call reuses the symmetric buffers)
        other = P.HashPartitionExchange(P.DevicePartitioner(ctx), peer=False)
        same = True
        for rep in range(2):
            a = exchange.exchange(page, [0], slot="xchk", lease=True)
            b = other.exchange(page, [0])
            same = same and a.position_count == b.position_count
            for x, y in zip(a.blocks, b.blocks):
                same = same and bool(torch.equal(x.values.view(torch.uint8), y.values.view(torch.uint8)))
        t = torch.tensor([1 if same else 0], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        out["peer_exchange_equals_nccl_all_to_all"] = bool(int(t.item()) == 1) if exchange.peer is not None else "peer mode unavailable: " + exchange.peer_why
        if out["peer_exchange_equals_nccl_all_to_all"] is False:
            raise SystemExit("bench cross-check failed: the peer-memory exchange and the NCCL all-to-all deliver different rows")
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 2),
            P.AggregatorFactory(P.AGG_SUM, P.LONG_DECIMAL, 3)]
    plan = P.PartitionedAggregation([P.BIGINT], [0], aggs, ctx, dist, expected_groups=n_orders, exchange=exchange)
    g, res = plan.run(page, slot="xchk")
    mine = [b.values.cpu().numpy() for b in res.blocks]
    join = P.PartitionedHashJoin([P.BIGINT, P.DOUBLE], [0], ctx, dist, exchange=exchange)
    bpage = P.Page(Block(N.BIGINT, od["orderkey"], None, None, n_orders), Block(N.DOUBLE, od["totalprice"].view(torch.int64), None, None, n_orders),
                   position_count=n_orders)
    join.build(bpage, slot="xchk_b")
    ppage = P.Page(Block(N.BIGINT, li["orderkey"], None, None, n), position_count=n)
    local, pp, got = join.probe(ppage, [0], [1], slot="xchk_p")
    jmine = (local.blocks[0].values[pp.long()].cpu().numpy(), got[0].values.view(torch.int64).cpu().numpy())
    if dist is not None:
        gathered = [None] * world
        dist.all_gather_object(gathered, (mine, jmine))
    else:
        gathered = [(mine, jmine)]
    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_lib as O
        shards = [shard(r, "cpu") for r in range(world)]
        keys = np.concatenate([s[0]["orderkey"].numpy() for s in shards])
        qty = np.concatenate([s[0]["quantity"].numpy() for s in shards])
        price = np.concatenate([s[0]["extendedprice_double"].numpy() for s in shards])
        ogb = O.OracleGroupBy([P.BIGINT], n_orders)
        ids = ogb.get_group_ids([P.Block.long_array(keys)])
        G = ogb.group_count()
        okeys = ogb.export(with_hash=False)[0][0].values
        want = [okeys]
        wide = wide_decimals(torch.from_numpy(qty), torch.from_numpy(keys)) if world == 1 else None
        if world > 1:       # row parity is per shard: build the column shard by shard
            from presto_b200.blocks import Block as B_
            parts = [wide_decimals(s[0]["quantity"], s[0]["orderkey"]) for s in shards]
            wide = B_(N.LONG_DECIMAL, np.concatenate([np.asarray(b.values) for b in parts]), None, None, len(keys))
        for a, col in zip(aggs, (None, P.Block.long_array(qty, None, P.SHORT_DECIMAL), P.Block.double_array(price), wide)):
            acc = O.OracleAccumulator(a.kind, a.in_type)
            acc.add_input(ids, G, col)
            want.append(np.asarray(acc.evaluate(0, G, True)[1][0].values))
        got_cols = [np.concatenate([np.asarray(gr[0][c]) for gr in gathered]) for c in range(5)]
        og, ow = np.argsort(got_cols[0], kind="stable"), np.argsort(want[0], kind="stable")
        ok = len(got_cols[0]) == G and np.array_equal(got_cols[0][og], want[0][ow]) and np.array_equal(got_cols[1][og], want[1][ow])
        u64 = lambda a: np.ascontiguousarray(a).view(np.uint64)      # noqa: E731
        ok = ok and np.array_equal(u64(got_cols[2]).reshape(-1, 2)[og], u64(want[2]).reshape(-1, 2)[ow])
        gd, wd = u64(got_cols[3]).view(np.float64)[og], u64(want[3]).view(np.float64)[ow]
        ok = ok and bool((np.abs(gd - wd) <= 1e-12 * np.abs(wd)).all())
        wide_got, wide_want = u64(got_cols[4]).reshape(-1, 2)[og], u64(want[4]).reshape(-1, 2)[ow]
        ok = ok and np.array_equal(wide_got, wide_want)
        out["long_decimal_sums_with_high_words_and_negatives"] = bool(np.array_equal(wide_got, wide_want) and (wide_want[:, 1] & np.uint64(0x7fffffffffffffff)).any()
                                                                   and (wide_want[:, 1] >> np.uint64(63)).any())
        out["groupby_orderkey_equals_oracle"] = bool(ok)
        oj = O.OracleJoin([P.BIGINT, P.DOUBLE], [0])
        for s in shards:
            oj.add_page([P.Block.long_array(s[1]["orderkey"].numpy()), P.Block.double_array(s[1]["totalprice"].numpy())])
        oj.build()
        wp, wb = oj.probe([P.Block.long_array(keys)])
        tp = np.concatenate([s[1]["totalprice"].numpy() for s in shards]).view(np.int64)
        want_pairs = np.stack([keys[wp], tp[wb]], axis=1)
        got_pairs = np.stack([np.concatenate([gr[1][0] for gr in gathered]), np.concatenate([gr[1][1] for gr in gathered])], axis=1)
        srt = lambda a: a[np.lexsort((a[:, 1], a[:, 0]))]      # noqa: E731
        out["join_equals_oracle"] = bool(want_pairs.shape == got_pairs.shape and np.array_equal(srt(want_pairs), srt(got_pairs)))
        out["rows"] = int(len(keys))
        out["what"] = "SF%g per GPU through the same plans vs oracle/ on one host core (group sets, counts, decimal sums, join match sets bit-exact; double sums 1e-12)" % sf
        if not (out["groupby_orderkey_equals_oracle"] and out["join_equals_oracle"]):
            raise SystemExit("bench cross-check against the oracle failed: %s" % json.dumps(out))
    return out


def q1_selfcheck(torch, cols, out, dist):
    """Size-independent check of the timed run's own result at full size: per-group count(*), sum(quantity), sum(extendedprice),
    sum(disc_price) and sum(charge) recomputed with torch int64 arithmetic and scatter-adds over the raw columns (independent of
    both the library and the oracle; every product and every per-group sum stays below 2^63 for TPC-H value ranges) and compared
    bit-exact with the operator's output; counts must also add up to the rows passing the filter."""
    from presto_b200 import tpch
    m = cols["shipdate"] <= tpch.Q1_CUTOFF
    code = (cols["returnflag"].to(torch.int64) * 256 + cols["linestatus"].to(torch.int64))[m]
    disc_price = cols["extendedprice"] * (100 - cols["discount"])
    exp = torch.zeros(5, 65536, dtype=torch.int64, device=code.device)
    exp[0].index_add_(0, code, torch.ones_like(code))
    exp[1].index_add_(0, code, cols["quantity"][m])
    exp[2].index_add_(0, code, cols["extendedprice"][m])
    exp[3].index_add_(0, code, disc_price[m])
    # sum(charge) can pass 2^63 at SF10: accumulate the low 32 bits and the rest separately, recombine in Python
    charge = (disc_price * (100 + cols["tax"]))[m]
    lo = torch.zeros(65536, dtype=torch.int64, device=code.device).index_add_(0, code, charge & 0xffffffff)
    hi = torch.zeros(65536, dtype=torch.int64, device=code.device).index_add_(0, code, charge >> 32)
    passing = torch.tensor([int(m.sum().item())], dtype=torch.int64, device=code.device)
    if dist is not None:
        for t in (exp, lo, hi, passing):
            dist.all_reduce(t)
    exp, lo, hi = exp.cpu().numpy(), lo.cpu().numpy(), hi.cpu().numpy()
    hb = [b.to_host() for b in out.blocks]
    ok, count_sum = True, 0
    for i in range(out.position_count):
        c = hb[0].get(i)[0] * 256 + hb[1].get(i)[0]
        got = (hb[9].get(i), hb[2].get(i), hb[3].get(i), hb[4].get(i), hb[5].get(i))
        want = tuple(int(exp[k][c]) for k in range(4)) + (int(lo[c]) + (int(hi[c]) << 32),)
        ok = ok and got == want
        count_sum += got[0]
    live = int((exp[0] > 0).sum())
    return {"groups": int(out.position_count), "groups_expected": live, "count_sum": int(count_sum), "rows_passing_filter": int(passing.item()),
            "sums_bit_exact_vs_torch_int64": bool(ok and live == out.position_count and count_sum == int(passing.item()))}


def gpu_arm(args):
    import numpy as np
    import torch
    import presto_b200 as P
    from presto_b200 import tpch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    ctx = P.GpuContext(local_rank)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        pass
    env = Env(P, tpch, torch, dist, ctx, world, rank, peak)
    aggs = tpch.q1_fused_aggregators()
    q1_filter, q1_projections = tpch.q1_expressions()
    processor = P.PageProcessor(q1_filter, q1_projections, ctx=ctx)

    cols = tpch.lineitem(args.sf, seed=42 + rank, device="cuda", need=("raw",))
    page = tpch.q1_raw_page(cols)
    rows = page.position_count
    input_bytes = rows * tpch.Q1_RAW_BYTES_PER_ROW    # > 126 MB L2: every step streams its inputs from HBM

    plan = P.PartialFinalAggregation([P.VARCHAR, P.VARCHAR], [0, 1], aggs, ctx, dist, expected_groups=100_000,
                                     reserve_sms=int(os.environ.get("PRESTOGPU_BENCH_RESERVE_SMS", "4" if world <= 2 else "8")), input_processor=processor,
                                     pipeline=os.environ.get("PRESTOGPU_BENCH_PIPELINE", "1") != "0")

    ctx.set_timing(True)
    with ClockSampler(local_rank if rank == 0 else None) as clk:
        l2 = plan.launch_count()
        ms, out, kernels, launches = env.timed(lambda: plan.submit(page), args.steps, args.warmup, drain=plan.drain)
        launches += plan.launch_count() - l2
    ctx.set_timing(False)
    total_rows = rows * world
    value = total_rows / (ms / 1e3)

    # dominant kernel + roofline (per-launch CUDA-event durations recorded inside the library on its launch stream)
    kern_ms = {k: v[0] for k, v in kernels.items()}
    top = max(kern_ms, key=kern_ms.get) if kern_ms else None
    roofline = None
    if top:
        tms, cnt = kernels[top]
        per_launch_ms = tms / max(cnt, 1)
        achieved = input_bytes / (per_launch_ms / 1e3) / 1e9
        roofline = {"bound": "hbm", "kernel": top, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                    "frac": round(achieved / peak, 4), "traffic": traffic.get(top), "peak_source": peak_src,
                    "algorithmic_bytes_per_row": tpch.Q1_RAW_BYTES_PER_ROW, "rows_per_launch": rows,
                    "launch_ms": round(per_launch_ms, 4), "launches_timed": cnt,
                    "kernel_share_of_step": round(tms / args.steps / ms, 4),
                    "step_frac": round(input_bytes / (ms / 1e3) / 1e9 / peak, 4),
                    "kernels_ms_per_step": {k: round(v[0] / args.steps, 4) for k, v in sorted(kernels.items(), key=lambda kv: -kv[1][0])},
                    "traffic_unit": "bytes/launch (ncu dram read+write)", "algorithmic_bytes_per_launch": int(input_bytes)}

    selfcheck = q1_selfcheck(torch, cols, out, dist)
    if not selfcheck["sums_bit_exact_vs_torch_int64"]:
        raise SystemExit("bench self-check failed: %s" % json.dumps(selfcheck))

    # e2e: host (pinned) buffers through the same operator / C ABI call; H2D copies inside the timed region
    e2e_pages = None
    e2e = None
    if not args.no_e2e:
        hpage = tpch.page_to_host(page, pinned=True)
        e_steps = max(2, min(args.steps, 5))
        ems, eout, _, _ = env.timed(lambda: plan.submit(hpage), e_steps, 3, drain=plan.drain)
        out_bytes = sum(int(np.asarray(b.to_host().values).nbytes) for b in eout.blocks)
        e2e = {"value": total_rows / (ems / 1e3), "unit": "rows/s", "h2d_bytes_per_step": int(input_bytes),
               "d2h_bytes_per_step": int(out_bytes), "ms_per_step": round(ems, 3), "steps": e_steps, "warmup": 3}
        # the same end-to-end pass with the input arriving as reference-sized pages on several driver threads
        if not args.no_e2e_pages:
            nthreads = max(1, min(args.page_drivers, (os.cpu_count() or 8) // max(world, 1)))

            def paged(zero_copy):
                drivers = PagedDrivers(P, tpch, torch, local_rank, hpage, nthreads, args.page_rows, args.page_batch_rows, zero_copy=zero_copy)
                for _ in range(2):
                    pout = drivers.step()
                env.barrier()
                l0 = drivers.launch_count()
                t0 = time.perf_counter()
                for _ in range(e_steps):
                    pout = drivers.step()
                torch.cuda.synchronize()
                pms = env.max_over_ranks((time.perf_counter() - t0) * 1e3) / e_steps
                pcheck = q1_selfcheck(torch, cols, pout, None) if world == 1 else {"sums_bit_exact_vs_torch_int64": None}
                if world == 1 and not pcheck["sums_bit_exact_vs_torch_int64"]:
                    raise SystemExit("bench self-check failed (e2e_pages): %s" % json.dumps(pcheck))
                return {"value": total_rows / (pms / 1e3), "unit": "rows/s", "ms_per_step": round(pms, 3),
                        "gpu_launches_per_step": (drivers.launch_count() - l0) // e_steps,
                        "ratio_to_e2e": round((total_rows / (pms / 1e3)) / e2e["value"], 4), "selfcheck": pcheck}

            zc, cp = paged(True), paged(False)
            e2e_pages = dict(zc)
            e2e_pages.update({"steps": e_steps, "warmup": 2, "page_rows": args.page_rows, "pages_per_step": int(math.ceil(rows / args.page_rows)),
                              "driver_threads": nthreads, "batch_rows": args.page_batch_rows, "h2d_bytes_per_step": int(input_bytes),
                              "pages_in": "caller-pinned host memory (PG_MEM_HOST_PINNED): the accumulator copies each page straight to the device",
                              "plain_host_pages": dict(cp, pages_in="plain host memory (PG_MEM_HOST, what a JNI caller hands over): each page is first "
                                                                    "copied into the pinned accumulator"),
                              "how": "host pages of page_rows rows, one addInput (pg_hashagg_add_page path, looped in C) per page on each of driver_threads "
                                     "PARTIAL operators; page accumulator; one FINAL merge; wall clock, max over ranks"})
        del hpage
    plan.close()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        n_s, times, _ = time_cpu_port(P, tpch, args.cpu_sample_rows / ROWS_PER_SF, 1, 3, 1)
        cpu = {"value": n_s / min(times), "unit": "rows/s", "cores": 1, "kind": "port",
               "sample": "oracle/ C++ restatement of PageProcessor + GroupByHash + accumulators (the reference is Java; no JDK on this box), 1 thread, "
                         "%d-row TPC-H-shaped lineitem sample of the same Q1 plan, best of 3" % n_s}

    del page, cols
    torch.cuda.empty_cache()
    workloads, xchk = None, None
    if not args.no_workloads:
        xchk = None if args.no_crosscheck else crosschecks(env)
        workloads = wide_workloads(args, env, traffic)

    if rank == 0:
        line = {
            "metric": METRIC,
            "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int128/int64 (decimal sums), u8 keys", "data": "synthetic (TPC-H-shaped generator, seeded)",
            "config": q1_config(args.sf, rows),
            "plan": "SINGLE" if world == 1 else "PARTIAL -> packed page in device memory -> NCCL all_gather (device to device) -> FINAL on every rank; the FINAL stage of step i overlaps the PARTIAL stage of step i+1",
            "clocks": clk.summary(), "gpu_launches": int(launches), "roofline": roofline, "e2e": e2e, "e2e_pages": e2e_pages,
            "selfcheck": selfcheck, "result_digest": result_digest(out) if world == 1 else None,
        }
        if cpu:
            line["cpu_baseline"] = cpu
        if workloads:
            line["workloads"] = workloads
            line["crosscheck"] = xchk
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def reference_arm(args):
    """CPU arm: the oracle port on all host threads (rank 0 only), over the same Q1 input one GPU gets (SF `--sf`, seed 42)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import presto_b200 as P
    from presto_b200 import tpch
    cores = os.cpu_count() or 1
    # the same generator (counter-based, device-independent) gives this arm the very rows rank 0's GPU aggregates
    warmup = args.warmup
    n_s, times, digest = time_cpu_port(P, tpch, args.sf, cores, args.steps, warmup)
    total = sum(times)
    value = n_s * len(times) / total
    line = {
        "impl": "reference",
        "metric": METRIC,
        "value": value, "unit": "rows/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total / len(times) * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int128/int64 (decimal sums), u8 keys", "data": "synthetic (TPC-H-shaped generator, seeded)",
        "config": q1_config(args.sf, n_s),
        "cpu_baseline": {"value": value, "unit": "rows/s", "cores": cores, "kind": "port",
                         "sample": "oracle/ C++ restatement (reference is Java, no JDK here): %d drivers run PARTIAL over slices of the same "
                                   "%d-row SF%g input one GPU gets (filter + projections, then PARTIAL), one FINAL merge; %d timed passes after %d warm-up"
                                   % (cores, n_s, args.sf, len(times), warmup)},
        "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "result_digest": digest,
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        reference_arm(a)
    else:
        gpu_arm(a)
