#!/usr/bin/env python
"""bench.py — input rows/s of the hash-aggregation hot path (BASELINE.json metric) on N B200s of one node.

Workload at every N: BASELINE.json configs[1], "TPC-H Q1 on lineitem SF10": GROUP BY returnflag, linestatus (2 x varchar(1), 4 groups)
with sum/avg over short and long decimals and count(*), 59.99 M input rows per GPU shard, masked by the shipdate filter.
  value : whole-job input rows/s with the columns already resident in HBM (one HashAggregationOperator pass per step)
  e2e   : the same pass through the C ABI with HOST (pinned) column buffers, host->device copies inside the timed region
  N > 1 : weak scaling — every rank aggregates its own SF10 shard (PARTIAL step), the 4-row intermediate states are
          all-gathered over NCCL and merged by a FINAL-step operator on every rank (the reference's partial->final plan).
`--impl reference` times the CPU restatement of the reference's operators (oracle/, kind "port": the reference is Java and no
JDK exists here) on all host threads over a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS_PER_SF = 5_998_605.2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--extra-shuffle", action="store_true", help="extras: lineitem rows in random order instead of clustered by order key")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sf", type=float, default=10.0, help="TPC-H scale factor of the lineitem shard per GPU")
    ap.add_argument("--cpu-sample-rows", type=int, default=6_000_000)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--extra", action="store_true", help="also time the high-cardinality GROUP BY and join workloads")
    ap.add_argument("--extra-sf", type=float, default=10.0)
    return ap.parse_args()


# ------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.device, self.samples, self._stop, self._t = device, [], threading.Event(), None

    def _run(self):
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
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(s) > 2 + i and s[2 + i].lower().startswith("active") for s in self.samples)]
        mx = [int(s[1]) for s in self.samples if s[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------- CPU arm (oracle port)
def oracle_q1_pass(O, P, blocks, aggs, nthreads):
    """The reference's plan on host cores: `nthreads` drivers each run getGroupIds + accumulators (PARTIAL) over a slice of
    the pages, then one FINAL merge — restated in oracle/ (C++), called through ctypes (GIL released)."""
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


def host_sample_blocks(P, tpch, rows, seed):
    import torch
    sf = rows / ROWS_PER_SF
    cols = tpch.lineitem(sf, seed=seed, device="cpu")
    page = tpch.q1_page(cols)
    return tpch.page_to_host(page, pinned=False).blocks


def time_cpu_port(P, tpch, rows, nthreads, steps, warmup, seed=42):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_lib as O
    blocks = host_sample_blocks(P, tpch, rows, seed)
    n = blocks[0].n
    aggs = tpch.q1_aggregators()
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        oracle_q1_pass(O, P, blocks, aggs, nthreads)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return n, times


# ------------------------------------------------------------------------------------------- GPU arm
def run_single_pass(P, page, aggs, step, ctx, expected=100_000):
    """One HashAggregationOperator pass over `page` (device or host buffers); returns the output page on the host."""
    op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], step, aggs, expected_groups=expected, ctx=ctx)
    op.add_input(page)
    op.finish()
    out = op.get_output()
    op.close()
    return out


_MERGE_BUFS = {}


def merge_partials_nccl(P, dist, torch, part_page, fin_aggs, ctx, world, run_final=True):
    """All-gather the (tiny) intermediate page of every rank over NCCL, then ONE FINAL step over the concatenated rows.
    The page is packed into a pinned buffer: [payload bytes][n][per block: values, offsets, nulls sizes][data]."""
    import numpy as np
    from presto_b200.blocks import Block, _NP_DTYPE
    cap = 1 << 16
    trace = _MERGE_BUFS.setdefault("trace", [0.0] * 5) if os.environ.get("PRESTOGPU_BENCH_TRACE") else None
    t0 = time.perf_counter()
    if "send" not in _MERGE_BUFS:
        _MERGE_BUFS["send_host"] = torch.zeros(cap, dtype=torch.uint8).pin_memory()
        _MERGE_BUFS["send"] = torch.zeros(cap, dtype=torch.uint8, device="cuda")
        _MERGE_BUFS["recv"] = torch.zeros(cap * world, dtype=torch.uint8, device="cuda")
        _MERGE_BUFS["recv_host"] = torch.zeros(cap * world, dtype=torch.uint8).pin_memory()
    blobs = []
    for b in part_page.blocks:
        hb = b.to_host()
        for arr in (hb.values, hb.offsets, hb.nulls):
            blobs.append(np.zeros(0, dtype=np.uint8) if arr is None else np.ascontiguousarray(arr).view(np.uint8).reshape(-1))
    sizes = np.array([part_page.position_count] + [len(a) for a in blobs], dtype=np.int64)
    payload = np.concatenate([sizes.view(np.uint8)] + blobs)
    assert len(payload) <= cap - 8
    sh = _MERGE_BUFS["send_host"].numpy()
    sh[:8] = np.array([len(payload)], dtype=np.int64).view(np.uint8)
    sh[8:8 + len(payload)] = payload
    t1 = time.perf_counter()
    _MERGE_BUFS["send"].copy_(_MERGE_BUFS["send_host"], non_blocking=True)
    dist.all_gather_into_tensor(_MERGE_BUFS["recv"], _MERGE_BUFS["send"])
    _MERGE_BUFS["recv_host"].copy_(_MERGE_BUFS["recv"], non_blocking=True)
    torch.cuda.current_stream().synchronize()
    t2 = time.perf_counter()
    host = _MERGE_BUFS["recv_host"].numpy()
    nb = len(part_page.blocks)
    per_block = [[] for _ in range(nb)]      # (values, offsets, nulls, n) of every rank
    for r in range(world):
        chunk = host[r * cap:(r + 1) * cap]
        ln = int(chunk[:8].view(np.int64)[0])
        pl = chunk[8:8 + ln]
        nsz = 1 + 3 * nb
        sz = pl[:8 * nsz].view(np.int64)
        pos = 8 * nsz
        n = int(sz[0])
        for i in range(nb):
            parts = []
            for j in range(3):
                l = int(sz[1 + 3 * i + j])
                parts.append(pl[pos:pos + l])
                pos += l
            per_block[i].append((parts[0], parts[1], parts[2], n))
    total = sum(t[3] for t in per_block[0])
    blocks = []
    for i, b in enumerate(part_page.blocks):
        offs = None
        if b.type in (P.VARCHAR, P.VARBINARY):
            pieces, vpieces, shift = [np.zeros(1, dtype=np.int32)], [], 0
            for t in per_block[i]:
                o = t[1].view(np.int32)
                used = int(o[-1]) if len(o) else 0
                pieces.append(o[1:] + shift)
                vpieces.append(t[0][:used])          # value buffers may be longer than the bytes in use
                shift += used
            offs = np.concatenate(pieces)
            vals = np.concatenate(vpieces) if shift else np.zeros(1, dtype=np.uint8)
        else:
            vals = np.concatenate([t[0] for t in per_block[i]]).view(_NP_DTYPE[b.type])
        nulls = None
        if any(len(t[2]) for t in per_block[i]):
            nulls = np.concatenate([t[2] if len(t[2]) else np.zeros(t[3], dtype=np.uint8) for t in per_block[i]])
        blocks.append(Block(b.type, vals, nulls, offs, total))
    t3 = time.perf_counter()
    if trace is not None:
        for i, d in enumerate((t1 - t0, t2 - t1, t3 - t2)):
            trace[i] += d
        trace[4] += 1
    if not run_final:
        return blocks, total
    return final_merge(P, blocks, total, fin_aggs, ctx)


def final_merge(P, blocks, total, fin_aggs, ctx):
    """The FINAL step over the gathered intermediate rows."""
    trace = _MERGE_BUFS.get("trace")
    t3 = time.perf_counter()
    fin = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, fin_aggs, expected_groups=1000, ctx=ctx)
    fin.add_input(P.Page(*blocks, position_count=total))
    fin.finish()
    out = fin.get_output()
    fin.close()
    if trace is not None:
        trace[3] += time.perf_counter() - t3
    return out


PACKED_CAP = 1 << 16


def partial_packed(P, page, aggs, ctx, pipe, expected=100_000):
    """PARTIAL pass on the main thread; its output leaves as one packed page (native, one call) in a pinned send buffer."""
    op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.PARTIAL, aggs, expected_groups=expected, ctx=ctx)
    op.add_input(page)
    op.finish()
    buf = pipe.next_send_buffer()
    op.get_output_packed(buf.data_ptr(), PACKED_CAP)
    op.close()
    pipe.submit(buf)


def gather_packed(dist, torch, send_host, world):
    """packed page of this rank -> NCCL all-gather -> the ranks' packed pages in pinned host memory.
    Returns the list of buffer addresses for HashAggregationOperator.add_packed."""
    cap = PACKED_CAP
    trace = _MERGE_BUFS.setdefault("trace", [0.0] * 5) if os.environ.get("PRESTOGPU_BENCH_TRACE") else None
    t0 = time.perf_counter()
    if "send" not in _MERGE_BUFS:
        _MERGE_BUFS["send"] = torch.zeros(cap, dtype=torch.uint8, device="cuda")
        _MERGE_BUFS["recv"] = torch.zeros(cap * world, dtype=torch.uint8, device="cuda")
        # several receive buffers: the FINAL worker of step i may still read its pages while later steps are gathered
        _MERGE_BUFS["recv_host"] = [torch.zeros(cap * world, dtype=torch.uint8).pin_memory() for _ in range(8)]
        _MERGE_BUFS["turn"] = 0
    t1 = time.perf_counter()
    rh = _MERGE_BUFS["recv_host"][_MERGE_BUFS["turn"] % 8]
    _MERGE_BUFS["turn"] += 1
    _MERGE_BUFS["send"].copy_(send_host, non_blocking=True)
    dist.all_gather_into_tensor(_MERGE_BUFS["recv"], _MERGE_BUFS["send"])
    rh.copy_(_MERGE_BUFS["recv"], non_blocking=True)
    torch.cuda.current_stream().synchronize()
    t2 = time.perf_counter()
    if trace is not None:
        trace[0] += t1 - t0
        trace[1] += t2 - t1
        trace[4] += 1
    base = rh.data_ptr()
    return [base + r * cap for r in range(world)]


def final_merge_packed(P, ptrs, fin_aggs, ctx):
    trace = _MERGE_BUFS.get("trace")
    t3 = time.perf_counter()
    fin = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, fin_aggs, expected_groups=1000, ctx=ctx)
    fin.add_packed(ptrs)
    fin.finish()
    out = fin.get_output()
    fin.close()
    if trace is not None:
        trace[3] += time.perf_counter() - t3
    return out


class MergePipeline:
    """Second stage of the N > 1 plan, running beside the first: while the main thread runs the PARTIAL pass of step i + 1,
    a communication thread all-gathers the intermediate rows of step i (NCCL on its own stream, one collective per step, in
    step order on every rank) and FINAL workers, each with its own pg_ctx and high-priority stream, run the FINAL step of
    alternate steps.  Every step's final result is produced; the stages merely overlap, as consecutive pages do in the
    reference's PARTIAL -> exchange -> FINAL plan.  (One FINAL step costs 0.4 ms on an idle GPU but 1.4 ms beside the PARTIAL
    kernel of the next step - it is a chain of small launches and read-backs - so a single worker capped the 8-GPU step at 2 ms.)"""

    FINAL_WORKERS = 2

    def __init__(self, P, dist, torch, fin_aggs, local_rank, world):
        import queue
        self.P, self.dist, self.torch, self.fin_aggs, self.local_rank, self.world = P, dist, torch, fin_aggs, local_rank, world
        self.q = queue.Queue(maxsize=3)      # bounded: the pinned send / receive buffers rotate over 8 slots
        self.fqs = [queue.Queue(maxsize=2) for _ in range(self.FINAL_WORKERS)]     # step i goes to worker i % FINAL_WORKERS (both get warmed up)
        self.results = {}
        self.fctx = []
        self.send_bufs = []
        self.submitted = 0
        self.err, self.ctx = None, None
        self.ready = threading.Event()
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()
        self.ready.wait()
        self.workers = []
        for wi in range(self.FINAL_WORKERS):
            ev = threading.Event()
            t = threading.Thread(target=self._final_worker, args=(ev, self.fqs[wi]), daemon=True)
            t.start()
            ev.wait()
            self.workers.append(t)

    def _run(self):
        torch = self.torch
        torch.cuda.set_device(self.local_rank)
        self.stream = torch.cuda.Stream(priority=-1)
        with torch.cuda.stream(self.stream):
            self.ctx = self.P.GpuContext(self.local_rank)      # bound to this thread's stream
            self.ready.set()
            while True:
                item = self.q.get()
                if item is None:
                    self.q.task_done()
                    return
                try:
                    if self.err is None:
                        step, send_host = item
                        self.fqs[step % self.FINAL_WORKERS].put((step, gather_packed(self.dist, torch, send_host, self.world)))
                except BaseException as e:   # surfaced by drain()
                    self.err = e
                self.q.task_done()

    def _final_worker(self, ready, fq):
        torch = self.torch
        torch.cuda.set_device(self.local_rank)
        stream = torch.cuda.Stream(priority=-1)    # its few small kernels go ahead of the next step's big one
        with torch.cuda.stream(stream):
            ctx = self.P.GpuContext(self.local_rank)
            self.fctx.append(ctx)
            ready.set()
            while True:
                item = fq.get()
                if item is None:
                    fq.task_done()
                    return
                try:
                    if self.err is None:
                        step, ptrs = item
                        self.results[step] = final_merge_packed(self.P, ptrs, self.fin_aggs, ctx)
                        self.results.pop(step - 8, None)
                except BaseException as e:
                    self.err = e
                fq.task_done()

    def next_send_buffer(self):
        if not self.send_bufs:
            self.send_bufs = [self.torch.zeros(PACKED_CAP, dtype=self.torch.uint8).pin_memory() for _ in range(8)]
        b = self.send_bufs[self.submitted % 8]
        return b

    def launch_count(self):
        return self.ctx.launch_count() + sum(c.launch_count() for c in self.fctx)

    def submit(self, page):
        self.q.put((self.submitted, page))
        self.submitted += 1

    def drain(self):
        self.q.join()
        for fq in self.fqs:
            fq.join()
        if self.err is not None:
            raise self.err
        return self.results.get(self.submitted - 1)

    def close(self):
        self.q.put(None)
        self.thread.join(timeout=10)
        for fq in self.fqs:
            fq.put(None)
        for t in self.workers:
            t.join(timeout=10)


def final_aggs_for(P, aggs, first_channel):
    fin, ch = [], first_channel
    for a in aggs:
        fin.append(P.AggregatorFactory(a.kind, a.in_type, ch))
        ch += len(P.intermediate_types(a.kind, a.in_type))
    return fin


def extra_workloads(args, P, tpch, ctx, torch, peak, dist=None, world=1, rank=0):
    """cfg3 (high-cardinality GROUP BY orderkey) and cfg4 (lineitem JOIN orders), device-resident inputs.
    world > 1: every rank holds its own SF shard (weak scaling); rows are hash-partitioned on orderkey and exchanged with an
    NCCL all-to-all (presto-1_b200/exchange.py) before the per-GPU operator, the reference's partitioned plan."""
    from presto_b200.blocks import Block
    N = P.native
    res = {}
    sf = args.extra_sf
    n_orders = int(1_500_000 * sf)
    # disjoint order-key ranges per rank: shard r owns orders [r*n_orders, (r+1)*n_orders)
    li = tpch.lineitem(sf, seed=42 + rank, device="cuda", need=("keys", "double"), shuffle=args.extra_shuffle)
    n = li["rows"]
    key_off = rank * tpch.order_key(torch.tensor(n_orders, dtype=torch.int64)).item()
    li["orderkey"] += key_off
    page = P.Page(Block(N.BIGINT, li["orderkey"], None, None, n), Block(N.SHORT_DECIMAL, li["quantity"], None, None, n),
                  Block(N.DOUBLE, li["extendedprice_double"].view(torch.int64), None, None, n), position_count=n)
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 2)]
    ex = P.HashPartitionExchange(P.DevicePartitioner(ctx)) if world > 1 else None

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        ctx.kernel_report(reset=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1) / steps
        if dist is not None:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, out

    def total(x):
        if dist is None:
            return int(x)
        t = torch.tensor([int(x)], device="cuda", dtype=torch.int64)
        dist.all_reduce(t)
        return int(t.item())

    ctx.set_timing(True)

    def gb_pass():
        local = ex.exchange(page, [0]) if ex else page
        op = P.HashAggregationOperator([P.BIGINT], [0], P.SINGLE, aggs, expected_groups=n_orders, ctx=ctx, device_output=True)
        op.add_input(local)
        op.finish()
        g = op.get_group_count()
        out = op.get_output(max_rows=max(g, 1))     # all groups materialised on the device
        op.close()
        return g

    ms, groups = timed(gb_pass, 3, 1)
    kern = ctx.kernel_report(reset=True)
    n_all, g_all = total(n), total(groups)
    # size-independent properties of the full-size result (outside the timed region): one group per order, the counts
    # add up to the rows, the decimal sums add up to the column sum (linearity), first-seen order = ascending first row
    local = ex.exchange(page, [0]) if ex else page
    chk = P.HashAggregationOperator([P.BIGINT], [0], P.SINGLE, aggs, expected_groups=n_orders, ctx=ctx, device_output=True)
    chk.add_input(local)
    chk.finish()
    gout = chk.get_output(max_rows=max(chk.get_group_count(), 1))
    chk.close()
    cnt_sum = total(gout.blocks[1].values.sum().item())
    qty_sum = total(gout.blocks[2].values.view(-1, 2)[:, 0].sum().item())     # long decimal: low words (sums are far below 2^63)
    gb_check = {"groups_equal_orders": g_all == total(n_orders), "count_sum_equals_rows": cnt_sum == n_all,
                "decimal_sum_equals_column_sum": qty_sum == total(li["quantity"].sum().item())}
    if world == 1 and not args.extra_shuffle:
        gb_check["keys_in_first_seen_order"] = bool(torch.equal(gout.blocks[0].values, torch.unique_consecutive(li["orderkey"])))
    if not all(gb_check.values()):
        raise SystemExit("bench self-check failed (GROUP BY orderkey): %s" % json.dumps(gb_check))
    del gout, local
    brow = 136 + (48 if world > 1 else 0)   # + 2 x 24 B partition pass (read + write of key and inputs) when G > 1
    res["groupby_orderkey"] = {"workload": "GROUP BY orderkey over lineitem SF%g per GPU: count(*), sum(quantity decimal), sum(extendedprice double)%s"
                                           % (sf, "; hash-partitioned all-to-all then SINGLE per GPU" if world > 1 else ""),
                               "rows": n_all, "groups": g_all, "ms_per_step": round(ms, 3), "rows_per_s": n_all / (ms / 1e3),
                               "algorithmic_bytes_per_row": brow, "hbm_frac_per_gpu": round(n_all / world * brow / (ms / 1e3) / 1e9 / peak, 4),
                               "kernels_ms_per_step": {k: round(v[0] / 3, 3) for k, v in sorted(kern.items(), key=lambda kv: -kv[1][0])[:8]},
                               "selfcheck": gb_check}
    del page
    # ---- join: build orders (orderkey unique, totalprice), probe lineitem.orderkey, gather totalprice
    od = tpch.orders(sf, seed=42 + rank, device="cuda")
    od["orderkey"] += key_off
    bpage = P.Page(Block(N.BIGINT, od["orderkey"], None, None, n_orders), Block(N.DOUBLE, od["totalprice"].view(torch.int64), None, None, n_orders),
                   position_count=n_orders)

    def build():
        local = ex.exchange(bpage, [0]) if ex else bpage
        idx = P.PagesIndex([P.BIGINT, P.DOUBLE], [0], ctx=ctx)
        idx.add_page(local)
        return idx.create_lookup_source()

    bms, ls = timed(build, 2, 1)
    bkern = ctx.kernel_report(reset=True)
    ppage = P.Page(Block(N.BIGINT, li["orderkey"], None, None, n), Block(N.SHORT_DECIMAL, li["quantity"], None, None, n), position_count=n)

    def probe():
        local = ex.exchange(ppage, [0]) if ex else ppage
        m = local.position_count
        pp, bp = ls.join_page([local.blocks[0]], None, P.JOIN_INNER, device=True, capacity=max(m, 1))
        out = ls.append_to(bp, [1], device=True)
        return len(pp)

    pms, matches = timed(probe, 3, 1)
    pkern = ctx.kernel_report(reset=True)
    ctx.set_timing(False)
    o_all, m_all = total(n_orders), total(matches)
    # every probe row matches exactly one build row; on one GPU the gathered payload must equal a torch gather
    join_check = {"matches_equal_probe_rows": m_all == n_all}
    if world == 1:
        pp, bp = ls.join_page([ppage.blocks[0]], None, P.JOIN_INNER, device=True, capacity=max(n, 1))
        got = ls.append_to(bp, [1], device=True)[0].values
        want = od["totalprice"].view(torch.int64)[torch.searchsorted(od["orderkey"], li["orderkey"])]
        join_check["probe_positions_ascending"] = bool(torch.equal(pp, torch.arange(n, device="cuda", dtype=torch.int32)))
        join_check["gathered_payload_bit_exact"] = bool(torch.equal(got.view(torch.int64), want))
        del pp, bp, got, want
    if not all(join_check.values()):
        raise SystemExit("bench self-check failed (join): %s" % json.dumps(join_check))
    res["join_build"] = {"workload": "hash build orders SF%g per GPU (orderkey, totalprice)" % sf, "rows": o_all, "ms_per_step": round(bms, 3),
                         "rows_per_s": o_all / (bms / 1e3), "algorithmic_bytes_per_row": 32,
                         "hbm_frac_per_gpu": round(o_all / world * 32 / (bms / 1e3) / 1e9 / peak, 4),
                         "kernels_ms_per_step": {k: round(v[0] / 2, 3) for k, v in sorted(bkern.items(), key=lambda kv: -kv[1][0])[:6]}}
    res["join_probe"] = {"workload": "INNER join probe lineitem.orderkey -> orders SF%g per GPU, emit (probe idx, build pos) + gather totalprice" % sf,
                         "rows": n_all, "matches": m_all, "ms_per_step": round(pms, 3), "rows_per_s": n_all / (pms / 1e3),
                         "algorithmic_bytes_per_row": 44, "hbm_frac_per_gpu": round(n_all / world * 44 / (pms / 1e3) / 1e9 / peak, 4),
                         "kernels_ms_per_step": {k: round(v[0] / 3, 3) for k, v in sorted(pkern.items(), key=lambda kv: -kv[1][0])[:6]},
                         "selfcheck": join_check}
    if ex:
        res["exchange"] = {"rows_sent_per_rank": ex.sent_rows, "bytes_sent_per_rank": ex.bytes_sent, "collective": ("fused partition + peer-memory stores over NVLink (pg_partition_scatter_peers into symmetric receive buffers)" if ex.peer is not None
                                          else "NCCL all_to_all_single per column (%s)" % ex.peer_why)}
    return res


def q1_selfcheck(torch, cols, out, dist):
    """Size-independent check of the timed run's own result at full size: per-group count(*), sum(quantity),
    sum(extendedprice) and sum(disc_price) recomputed with torch int64 scatter-adds (independent of both the library and
    the oracle) and compared bit-exact with the operator's output; counts must also add up to the rows passing the mask."""
    m = cols["q1_mask"].bool()
    code = (cols["returnflag"].to(torch.int64) * 256 + cols["linestatus"].to(torch.int64))[m]
    exp = torch.zeros(4, 65536, dtype=torch.int64, device=code.device)
    exp[0].index_add_(0, code, torch.ones_like(code))
    exp[1].index_add_(0, code, cols["quantity"][m])
    exp[2].index_add_(0, code, cols["extendedprice"][m])
    exp[3].index_add_(0, code, cols["disc_price"][0::2][m])
    passing = torch.tensor([int(m.sum().item())], dtype=torch.int64, device=code.device)
    if dist is not None:
        dist.all_reduce(exp)
        dist.all_reduce(passing)
    exp = exp.cpu().numpy()
    hb = [b.to_host() for b in out.blocks]
    ok, count_sum = True, 0
    for i in range(out.position_count):
        c = hb[0].get(i)[0] * 256 + hb[1].get(i)[0]
        got = (hb[9].get(i), hb[2].get(i), hb[3].get(i), hb[4].get(i))
        ok = ok and got == tuple(int(exp[k][c]) for k in range(4))
        count_sum += got[0]
    live = int((exp[0] > 0).sum())
    return {"groups": int(out.position_count), "groups_expected": live, "count_sum": int(count_sum), "rows_passing_mask": int(passing.item()),
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
    aggs = tpch.q1_aggregators()
    fin_aggs = final_aggs_for(P, aggs, 2)

    cols = tpch.lineitem(args.sf, seed=42 + rank, device="cuda")
    page = tpch.q1_page(cols)
    rows = page.position_count
    input_bytes = rows * tpch.Q1_BYTES_PER_ROW    # > 126 MB L2: every step streams its inputs from HBM

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    pipe = MergePipeline(P, dist, torch, fin_aggs, local_rank, world) if world > 1 else None
    if pipe:
        ctx.reserve_sms(int(os.environ.get("PRESTOGPU_BENCH_RESERVE_SMS", "8")))   # room for the FINAL stage's small kernels

    def step_device():
        if world == 1:
            return run_single_pass(P, page, aggs, P.SINGLE, ctx)
        partial_packed(P, page, aggs, ctx, pipe)     # FINAL of this step overlaps the next PARTIAL
        return None

    def launches_now():
        return ctx.launch_count() + (pipe.launch_count() if pipe else 0)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        if pipe:
            pipe.drain()
        barrier()
        ctx.kernel_report(reset=True)
        launches0 = launches_now()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = None
        for _ in range(steps):
            out = fn()
        if pipe:
            out = pipe.drain()          # the last step's FINAL result must exist before the clock stops
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, out, launches_now() - launches0

    ctx.set_timing(True)
    with ClockSampler(local_rank) as clk:
        ms, out, launches = timed(step_device, args.steps, args.warmup)
    kernels = ctx.kernel_report(reset=True)
    ctx.set_timing(False)
    total_rows = rows * world
    value = total_rows * args.steps / (ms / 1e3)

    # dominant kernel + roofline (per-launch CUDA-event durations recorded inside the library on its launch stream)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    kern_ms = {k: v[0] for k, v in kernels.items()}
    top = max(kern_ms, key=kern_ms.get) if kern_ms else None
    roofline = None
    if top:
        tms, cnt = kernels[top]
        per_launch_ms = tms / max(cnt, 1)
        achieved = input_bytes / (per_launch_ms / 1e3) / 1e9
        roofline = {"bound": "hbm", "kernel": top, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                    "frac": round(achieved / peak, 4), "traffic": None, "peak_source": peak_src,
                    "algorithmic_bytes_per_row": tpch.Q1_BYTES_PER_ROW, "rows_per_launch": rows,
                    "launch_ms": round(per_launch_ms, 4), "launches_timed": cnt,
                    "kernel_share_of_step": round(tms / ms, 4),
                    "kernels_ms_per_step": {k: round(v[0] / args.steps, 4) for k, v in sorted(kernels.items(), key=lambda kv: -kv[1][0])}}
        prof = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(prof):
            try:
                roofline["traffic"] = json.load(open(prof)).get(top)      # bytes per launch, from the committed ncu capture
                roofline["traffic_unit"] = "bytes/launch (ncu dram read+write)"
                roofline["algorithmic_bytes_per_launch"] = int(input_bytes)
            except Exception:
                pass

    selfcheck = q1_selfcheck(torch, cols, out, dist)
    if not selfcheck["sums_bit_exact_vs_torch_int64"]:
        raise SystemExit("bench self-check failed: %s" % json.dumps(selfcheck))

    # e2e: host (pinned) buffers through the same operator / C ABI call; H2D copies inside the timed region
    e2e = None
    if not args.no_e2e:
        hpage = tpch.page_to_host(page, pinned=True)

        def step_host():
            if world == 1:
                return run_single_pass(P, hpage, aggs, P.SINGLE, ctx)
            partial_packed(P, hpage, aggs, ctx, pipe)
            return None

        e_steps = max(2, min(args.steps, 5))
        ems, eout, _ = timed(step_host, e_steps, 2)
        out_bytes = sum(int(np.asarray(b.to_host().values).nbytes) for b in eout.blocks)
        e2e = {"value": total_rows * e_steps / (ems / 1e3), "unit": "rows/s", "h2d_bytes_per_step": int(input_bytes),
               "d2h_bytes_per_step": int(out_bytes), "ms_per_step": round(ems / e_steps, 3), "steps": e_steps}
        del hpage

    # parity spot check of this very run against the oracle on a bounded prefix (the full check lives in tests/)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        n_s, times = time_cpu_port(P, tpch, args.cpu_sample_rows, 1, 3, 1)
        cpu = {"value": n_s / min(times), "unit": "rows/s", "cores": 1, "kind": "port",
               "sample": "oracle/ C++ restatement of GroupByHash+accumulators (the reference is Java; no JDK on this box), 1 thread, "
                         "%d-row TPC-H-shaped lineitem sample of the same Q1 plan, best of 3" % n_s}

    extra = None
    if args.extra:
        del page, cols
        torch.cuda.empty_cache()
        extra = extra_workloads(args, P, tpch, ctx, torch, peak, dist, world, rank)

    if rank == 0:
        line = {
            "metric": "input rows/sec for GROUP BY & hash-join probe at 1/2/4/8 B200; % HBM roofline",
            "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int128/int64 (decimal sums), u8 keys", "data": "synthetic (TPC-H-shaped generator, seeded)",
            "config": {"workload": "TPC-H Q1 on lineitem SF%g per GPU: GROUP BY returnflag,linestatus; 4x sum, 3x avg (decimal), count(*); "
                                   "shipdate filter as mask" % args.sf,
                       "rows_per_gpu": rows, "groups": int(out.position_count), "step": "SINGLE" if world == 1 else "PARTIAL -> NCCL all_gather -> FINAL; the FINAL stage of step i overlaps the PARTIAL stage of step i+1 (second stream)",
                       "l2_policy": "inputs (%.2f GB per pass) exceed the 126 MB L2; no flush needed" % (input_bytes / 1e9)},
            "clocks": clk.summary(), "gpu_launches": int(launches), "roofline": roofline, "e2e": e2e,
            "selfcheck": selfcheck,
        }
        if cpu:
            line["cpu_baseline"] = cpu
        if extra:
            line["extra"] = extra
        print(json.dumps(line), flush=True)
    if pipe:
        pipe.close()
        if rank == 0 and "trace" in _MERGE_BUFS:
            tr = _MERGE_BUFS["trace"]
            print("merge stage ms per call: pack %.3f gather %.3f unpack %.3f final %.3f (calls %d)"
                  % tuple([1e3 * x / max(tr[4], 1) for x in tr[:4]] + [int(tr[4])]), file=sys.stderr)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def reference_arm(args):
    """CPU arm: the oracle port on all host threads (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import presto_b200 as P
    from presto_b200 import tpch
    cores = os.cpu_count() or 1
    n_s, times = time_cpu_port(P, tpch, args.cpu_sample_rows, cores, args.steps, args.warmup)
    total = sum(times)
    value = n_s * len(times) / total
    line = {
        "impl": "reference",
        "metric": "input rows/sec for GROUP BY & hash-join probe at 1/2/4/8 B200; % HBM roofline",
        "value": value, "unit": "rows/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total / len(times) * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int128/int64 (decimal sums), u8 keys", "data": "synthetic (TPC-H-shaped generator, seeded)",
        "config": {"workload": "TPC-H Q1 on lineitem SF%g per GPU: GROUP BY returnflag,linestatus; 4x sum, 3x avg (decimal), count(*); "
                               "shipdate filter as mask" % args.sf, "rows_per_step": n_s},
        "cpu_baseline": {"value": value, "unit": "rows/s", "cores": cores, "kind": "port",
                         "sample": "oracle/ C++ restatement (reference is Java, no JDK here): %d drivers run PARTIAL over slices of a "
                                   "%d-row sample, one FINAL merge" % (cores, n_s)},
        "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        reference_arm(a)
    else:
        gpu_arm(a)
