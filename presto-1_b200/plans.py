"""Multi-GPU plans — the device replacement for the reference's exchange-connected operator pipelines, one process per GPU.

  PartialFinalAggregation   few groups (cfg1/cfg2):  PARTIAL per rank -> intermediate rows all-gathered -> FINAL on every rank.
                            Mirrors AggregationNode.Step PARTIAL/FINAL around a gathering exchange
                            (MAIN/sql/planner/LocalExecutionPlanner.java:3035-3054, OP/HashAggregationOperator.java:381-518).
  PartitionedAggregation    many groups (cfg3): rows hash-partitioned on the group-by channels (HashPartitionExchange), SINGLE step
                            per rank on a disjoint key range — PartitionedOutputOperator + FIXED_HASH_DISTRIBUTION.
  PartitionedHashJoin       cfg4: build and probe sides partitioned by the same function of the join key, one lookup source per
                            rank — the multi-GPU form of PartitionedLookupSourceFactory / PartitionedLookupSource
                            (OP/PartitionedLookupSourceFactory.java:110-296, OP/PartitionedLookupSource.java:92-226): a probe row is
                            routed to the one partition that can hold its key instead of every driver consulting N partitions.

With world_size == 1 every plan degenerates to the plain single-GPU operator (no exchange), so callers need one code path.
"""
import queue
import threading

from . import _native as N
from .accumulators import intermediate_types
from .exchange import DevicePartitioner, HashPartitionExchange
from .join import PagesIndex
from .operators import FINAL, PARTIAL, SINGLE, AggregatorFactory, HashAggregationOperator

PACKED_CAP = 1 << 15      # bytes reserved per rank for one packed intermediate page (64 groups x 10 columns fit; 4 groups are < 1 KB)


def final_aggregators(aggregators, first_channel):
    """The FINAL step's accumulator factories: aggregate i reads its intermediate columns starting where the PARTIAL step's output
    page put them (keys first, then each aggregate's intermediate columns in order)."""
    fin, ch = [], first_channel
    for a in aggregators:
        fin.append(AggregatorFactory(a.kind, a.in_type, ch))
        ch += len(intermediate_types(a.kind, a.in_type))
    return fin


class PartialFinalAggregation:
    """PARTIAL -> all-gather -> FINAL for a GROUP BY with few groups, as a two-stage pipeline.

    submit(page) runs the PARTIAL step of one page set on the calling thread; its intermediate page (a few rows) is written as one
    packed page into device memory (pg_hashagg_get_output_packed_device) and handed to the second stage, which all-gathers the
    pages of every rank over NCCL device to device and runs the FINAL step (pg_hashagg_add_packed_device -> addIntermediate) —
    beside the PARTIAL step of the next submit(), the way consecutive page sets overlap in the reference's pipelined plan.
    drain() waits for every submitted page set and returns the newest FINAL output.  Each stage has its own pg_ctx (stream); the
    second stage's streams have high priority and `reserve_sms` multiprocessors are kept out of the first stage's persistent
    kernels so that its few small kernels are not queued behind them.  pipeline=False runs both stages in sequence on the caller's
    context (no threads)."""

    FINAL_WORKERS = int(__import__('os').environ.get('PRESTOGPU_FINAL_WORKERS', '2'))
    SLOTS = 8

    def __init__(self, group_by_types, group_by_channels, aggregators, ctx, dist=None, expected_groups=100_000, reserve_sms=4,
                 input_processor=None, packed_bytes=PACKED_CAP, pipeline=True):
        import torch
        self.torch, self.dist, self.ctx = torch, dist, ctx
        self.input_processor = input_processor      # fused ScanFilterAndProject in front of the first stage (PageProcessor)
        self.types, self.channels, self.aggs = list(group_by_types), list(group_by_channels), list(aggregators)
        self.expected = expected_groups
        self.world = dist.get_world_size() if dist is not None else 1
        self.fin_aggs = final_aggregators(self.aggs, len(self.types))
        self.submitted = 0
        self.results = {}
        self.err = None
        self._threads = []
        self.packed_bytes = packed_bytes
        self.pipeline = pipeline and self.world > 1
        if self.world == 1:
            return
        self.device = torch.cuda.current_device()
        self.send = [torch.zeros(packed_bytes, dtype=torch.uint8, device="cuda") for _ in range(self.SLOTS)]
        self.recv = [torch.zeros(packed_bytes * self.world, dtype=torch.uint8, device="cuda") for _ in range(self.SLOTS)]
        self.fctx, self.comm_ctx = [], None
        if not self.pipeline:
            return
        self.q = queue.Queue(maxsize=3)           # bounded: the send / receive buffers rotate over SLOTS slots
        self.fqs = [queue.Queue(maxsize=2) for _ in range(self.FINAL_WORKERS)]
        ready = threading.Event()
        t = threading.Thread(target=self._comm_loop, args=(ready,), daemon=True)
        t.start()
        ready.wait()
        self._threads.append(t)
        for wi in range(self.FINAL_WORKERS):
            ev = threading.Event()
            t = threading.Thread(target=self._final_loop, args=(ev, self.fqs[wi]), daemon=True)
            t.start()
            ev.wait()
            self._threads.append(t)
        if reserve_sms:
            ctx.reserve_sms(reserve_sms)

    def _final(self, ctx, recv):
        fin = HashAggregationOperator(self.types, list(range(len(self.types))), FINAL, self.fin_aggs, expected_groups=1000, ctx=ctx)
        fin.add_packed_device(recv.data_ptr(), self.packed_bytes, self.world)
        fin.finish()
        out = fin.get_output()
        fin.close()
        return out

    # ---- stage 1 (calling thread)
    def submit(self, page):
        step = SINGLE if self.world == 1 else PARTIAL
        op = HashAggregationOperator(self.types, self.channels, step, self.aggs, expected_groups=self.expected, ctx=self.ctx,
                                     input_processor=self.input_processor)
        op.add_input(page)
        op.finish()
        k = self.submitted
        self.submitted += 1
        if self.world == 1:
            self.results = {k: op.get_output()}
            op.close()
            return
        slot = k % self.SLOTS
        op.get_output_packed_device(self.send[slot].data_ptr(), self.packed_bytes)      # complete (synchronised) when this returns
        op.close()
        if self.pipeline:
            self.q.put((k, slot))
            return
        self.dist.all_gather_into_tensor(self.recv[slot], self.send[slot])       # stream-ordered before the FINAL step's reads
        self.results = {k: self._final(self.ctx, self.recv[slot])}

    # ---- stage 2a: one collective per page set, in submit order on every rank
    def _comm_loop(self, ready):
        torch = self.torch
        torch.cuda.set_device(self.device)
        stream = torch.cuda.Stream(priority=-1)
        with torch.cuda.stream(stream):
            from .context import GpuContext
            self.comm_ctx = GpuContext(self.device)
            ready.set()
            while True:
                item = self.q.get()
                if item is None:
                    self.q.task_done()
                    return
                try:
                    if self.err is None:
                        k, slot = item
                        self.dist.all_gather_into_tensor(self.recv[slot], self.send[slot])
                        stream.synchronize()
                        self.fqs[k % self.FINAL_WORKERS].put((k, slot))
                except BaseException as e:          # surfaced by drain()
                    self.err = e
                self.q.task_done()

    # ---- stage 2b: FINAL over the gathered intermediate rows
    def _final_loop(self, ready, fq):
        torch = self.torch
        torch.cuda.set_device(self.device)
        stream = torch.cuda.Stream(priority=-1)
        with torch.cuda.stream(stream):
            from .context import GpuContext
            ctx = GpuContext(self.device)
            self.fctx.append(ctx)
            ready.set()
            while True:
                item = fq.get()
                if item is None:
                    fq.task_done()
                    return
                try:
                    if self.err is None:
                        k, slot = item
                        self.results[k] = self._final(ctx, self.recv[slot])
                        self.results.pop(k - self.SLOTS, None)
                except BaseException as e:
                    self.err = e
                fq.task_done()

    def drain(self):
        if self.pipeline:
            self.q.join()
            for fq in self.fqs:
                fq.join()
            if self.err is not None:
                raise self.err
        return self.results.get(self.submitted - 1)

    def launch_count(self):
        """Kernels launched by the second stage's contexts (the first stage's are counted on the caller's ctx)."""
        if not self.pipeline:
            return 0
        return self.comm_ctx.launch_count() + sum(c.launch_count() for c in self.fctx)

    def close(self):
        if not self.pipeline:
            return
        self.q.put(None)
        for fq in self.fqs:
            fq.put(None)
        for t in self._threads:
            t.join(timeout=10)
        self._threads = []
        self.pipeline = False


class PartitionedAggregation:
    """GROUP BY with many groups on N GPUs: exchange on the group-by channels, then a SINGLE step per rank."""

    def __init__(self, group_by_types, group_by_channels, aggregators, ctx, dist=None, expected_groups=100_000, exchange=None,
                 device_output=True):
        self.types, self.channels, self.aggs, self.ctx = list(group_by_types), list(group_by_channels), list(aggregators), ctx
        self.expected, self.device_output = expected_groups, device_output
        self.world = dist.get_world_size() if dist is not None else 1
        self.exchange = exchange if exchange is not None else (HashPartitionExchange(DevicePartitioner(ctx)) if self.world > 1 else None)

    def run(self, page, slot="agg"):
        """One page set through exchange + aggregation; returns (group count on this rank, output page)."""
        local = self.exchange.exchange(page, self.channels, slot=slot, lease=True) if self.exchange else page
        op = HashAggregationOperator(self.types, self.channels, SINGLE, self.aggs, expected_groups=self.expected, ctx=self.ctx,
                                     device_output=self.device_output)
        op.add_input(local)
        op.finish()
        g = op.get_group_count()
        out = op.get_output(max_rows=max(g, 1))
        op.close()
        return g, out


class PartitionedHashJoin:
    """INNER / PROBE_OUTER hash join on N GPUs: both sides exchanged on the join key; each rank builds and probes its partition."""

    def __init__(self, build_types, build_join_channels, ctx, dist=None, exchange=None):
        self.build_types, self.build_join_channels, self.ctx = list(build_types), list(build_join_channels), ctx
        self.world = dist.get_world_size() if dist is not None else 1
        self.exchange = exchange if exchange is not None else (HashPartitionExchange(DevicePartitioner(ctx)) if self.world > 1 else None)
        self.lookup = None

    def build(self, page, slot="build"):
        local = self.exchange.exchange(page, self.build_join_channels, slot=slot, lease=True) if self.exchange else page
        idx = PagesIndex(self.build_types, self.build_join_channels, ctx=self.ctx)
        idx.add_page(local)
        self.lookup = idx.create_lookup_source()
        return self.lookup

    def probe(self, page, probe_join_channels, build_output_channels, join_type=N.JOIN_INNER, slot="probe"):
        """Returns (exchanged probe page, probe positions into it, gathered build blocks)."""
        local = self.exchange.exchange(page, probe_join_channels, slot=slot, lease=True) if self.exchange else page
        m = local.position_count
        pp, out = self.lookup.join_and_append([local.blocks[c] for c in probe_join_channels], build_output_channels, None, join_type, device=True,
                                              capacity=max(m, 1))
        return local, pp, out
