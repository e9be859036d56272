"""Operator shells — host mirror of MAIN/operator/Operator.java:20-104 for the three operators on the hot path:
HashAggregationOperator (:381-518), HashBuilderOperator (:308-350, 428-496) and LookupJoinOperator (:207-378).

Each keeps the reference's pull/push protocol: needs_input() / add_input(page) / get_output() / finish() / is_finished().
"""
import ctypes as C

import numpy as np

from . import _native as N
from .accumulators import final_type, intermediate_types
from .blocks import Block, OutBlock, Page, col_array
from .context import default_context
from .join import LookupSource, PagesIndex

SINGLE, PARTIAL, FINAL, INTERMEDIATE = N.STEP_SINGLE, N.STEP_PARTIAL, N.STEP_FINAL, N.STEP_INTERMEDIATE


class AggregatorFactory:
    """AccumulatorFactory stand-in: (kind, argument type, input channel, mask channel)."""

    def __init__(self, kind, in_type=0, input_channel=-1, mask_channel=-1):
        self.kind, self.in_type, self.input_channel, self.mask_channel = kind, in_type, input_channel, mask_channel


class HashAggregationOperator:
    """GROUP BY operator: addInput pages, then after finish() getOutput returns pages of [keys..., (hash), aggregates...]."""

    def __init__(self, group_by_types, group_by_channels, step, aggregators, hash_channel=None, expected_groups=10_000,
                 max_partial_memory=16 << 20, output_page_rows=8192, ctx=None, device_output=False, input_processor=None):
        self.ctx = ctx or default_context()
        self.step = step
        self.aggregators = list(aggregators)
        self.group_by_types = list(group_by_types)
        self.hash_channel = -1 if hash_channel is None else hash_channel
        self.max_partial_memory = max_partial_memory
        self.output_page_rows = output_page_rows
        self.device_output = device_output
        self._h = C.c_void_p()
        kc = (C.c_int32 * len(group_by_channels))(*group_by_channels)
        kt = (C.c_int32 * len(group_by_types))(*group_by_types)
        specs = (N.PgAggSpec * max(1, len(self.aggregators)))()
        for i, a in enumerate(self.aggregators):
            specs[i].kind, specs[i].in_type, specs[i].input_channel, specs[i].mask_channel = a.kind, a.in_type, a.input_channel, a.mask_channel
        N.check(self.ctx.handle, N.lib().pg_hashagg_create(self.ctx.handle, kc, kt, len(group_by_types), self.hash_channel, specs,
                                                           len(self.aggregators), step, expected_groups, C.byref(self._h)))
        # ScanFilterAndProject fused in front: add_input then takes the processor's INPUT pages and the channels above index its projections
        self.input_processor = input_processor
        if input_processor is not None:
            N.check(self.ctx.handle, N.lib().pg_hashagg_set_input_processor(self._h, input_processor.handle))
        self._finishing = False
        self._finished = False
        self._next_group = 0
        self._flushing = False
        self._unfinished = None      # unfinishedWork (:398-417): a page the memory limit has not let in yet

    def set_page_batching(self, rows):
        """Reference-sized pages: small host pages are accumulated and enter the kernels `rows` rows at a time (0 = off)."""
        N.check(self.ctx.handle, N.lib().pg_hashagg_set_page_batching(self._h, int(rows)))

    def add_pages(self, page, page_rows):
        """The host page as consecutive pages of page_rows rows, each through the per-page addInput path (pg_hashagg_add_pages)."""
        assert self.needs_input(), "Operator is already finishing"
        blocks = page.blocks if isinstance(page, Page) else list(page)
        N.check(self.ctx.handle, N.lib().pg_hashagg_add_pages(self._h, col_array(blocks), len(blocks), blocks[0].n, int(page_rows)))

    def set_memory_limit(self, nbytes):
        """The operator's memory yield: with a limit set, add_input keeps a page that could grow the table past it as unfinished
        work instead of consuming it (pg_hashagg_set_memory_limit)."""
        N.check(self.ctx.handle, N.lib().pg_hashagg_set_memory_limit(self._h, int(nbytes)))

    # ---- Operator interface
    def needs_input(self):       # HashAggregationOperator.java:367-378
        return not self._finishing and not self._flushing and self._unfinished is None

    def _process_unfinished(self):
        """unfinishedWork.process(): True when no page is waiting any more."""
        if self._unfinished is None:
            return True
        blocks = self._unfinished
        rc = N.lib().pg_hashagg_add_page(self._h, col_array(blocks), len(blocks), blocks[0].n)
        if rc == N.PG_NEED_MEMORY:
            return False
        N.check(self.ctx.handle, rc)
        self._unfinished = None
        if self.step == PARTIAL and self.get_estimated_size() > self.max_partial_memory:   # builder full -> flush (:208-215, :501-517)
            self._flushing = True
        return True

    def add_input(self, page):   # :381-440
        assert self.needs_input(), "Operator is already finishing"
        self._unfinished = page.blocks if isinstance(page, Page) else list(page)
        self._process_unfinished()

    def has_unfinished_input(self):
        return self._unfinished is not None

    def finish(self):
        self._finishing = True

    def is_finished(self):
        return self._finished

    def _stats(self):
        g, cap, db, rows = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        N.check(self.ctx.handle, N.lib().pg_hashagg_stats(self._h, C.byref(g), C.byref(cap), C.byref(db), C.byref(rows)))
        return g.value, cap.value, db.value, rows.value

    def get_group_count(self):
        return self._stats()[0]

    def get_capacity(self):
        return self._stats()[1]

    def get_estimated_size(self):
        """InMemoryHashAggregationBuilder.getSizeInMemory: bytes that grow with the group count (slot at fill 0.75, stored
        key row, accumulator states).  The fixed minimum device allocation is reported by get_device_bytes()."""
        per_group = 22 + 8 * (len(self.group_by_types) + 1) + 32 * len(self.aggregators)
        return self._stats()[0] * per_group

    def get_device_bytes(self):
        return self._stats()[2]

    def output_types(self):
        n = C.c_int32()
        buf = (C.c_int32 * 128)()
        N.check(self.ctx.handle, N.lib().pg_hashagg_output_types(self._h, buf, 128, C.byref(n)))
        return list(buf[:n.value])

    def get_output(self, max_rows=None):   # :469-518
        if not self._process_unfinished():     # :473-481: the waiting page first; still blocked -> no output
            return None
        if not (self._finishing or self._flushing) or self._finished:
            return None
        groups = self._stats()[0]
        if self._next_group >= groups:
            if self._flushing and not self._finishing:
                N.check(self.ctx.handle, N.lib().pg_hashagg_reset(self._h))
                self._flushing = False
                self._next_group = 0
            else:
                self._finished = True
            return None
        n = min(max_rows or self.output_page_rows, groups - self._next_group)
        types = self.output_types()
        outs = []
        for i, t in enumerate(types):
            vb = 0
            if t in (N.VARCHAR, N.VARBINARY):
                b = C.c_int64()
                N.check(self.ctx.handle, N.lib().pg_hashagg_varwidth_bytes(self._h, i, C.byref(b)))
                vb = b.value
            outs.append(OutBlock(t, n, vb, device=self.device_output))
        arr = (N.PgColOut * len(outs))()
        for i, o in enumerate(outs):
            arr[i] = o.pg_col_out()
        eg = C.c_int64(-1)
        N.check(self.ctx.handle, N.lib().pg_hashagg_get_output(self._h, self._next_group, n, arr, C.byref(eg)))
        self._next_group += n
        return Page(*[o.block(n) for o in outs], position_count=n)

    def get_output_packed(self, buf_ptr, capacity):
        """All groups as ONE packed host page (the PARTIAL -> FINAL wire format, include/prestogpu.h) written at buf_ptr.
        Returns the number of bytes; the operator is finished afterwards."""
        groups = self._stats()[0]
        nbytes, eg = C.c_int64(), C.c_int64(-1)
        N.check(self.ctx.handle, N.lib().pg_hashagg_get_output_packed(self._h, 0, groups, C.c_void_p(buf_ptr), capacity, C.byref(nbytes), C.byref(eg)))
        self._next_group = groups
        self._finished = True
        return nbytes.value

    def get_output_packed_device(self, device_ptr, capacity):
        """get_output_packed into device memory (the send buffer of an all-gather)."""
        groups = self._stats()[0]
        nbytes, eg = C.c_int64(), C.c_int64(-1)
        N.check(self.ctx.handle, N.lib().pg_hashagg_get_output_packed_device(self._h, 0, groups, C.c_void_p(device_ptr), capacity, C.byref(nbytes),
                                                                             C.byref(eg)))
        self._next_group = groups
        self._finished = True
        return nbytes.value

    def add_packed_device(self, device_ptr, stride, count):
        """`count` packed pages `stride` bytes apart in device memory (an all-gather's receive buffer) as one input page."""
        assert self.needs_input(), "Operator is already finishing"
        N.check(self.ctx.handle, N.lib().pg_hashagg_add_packed_device(self._h, C.c_void_p(device_ptr), stride, count))

    def add_packed(self, buf_ptrs):
        """Packed pages (one per producer) as one input page, rows in buffer order."""
        assert self.needs_input(), "Operator is already finishing"
        arr = (C.c_void_p * len(buf_ptrs))(*buf_ptrs)
        N.check(self.ctx.handle, N.lib().pg_hashagg_add_packed(self._h, arr, len(buf_ptrs)))

    def close(self):
        if self._h:
            N.lib().pg_hashagg_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class HashBuilderOperator:
    """Join build side.  State machine reduced to CONSUMING_INPUT -> LOOKUP_SOURCE_BUILT (no spill on the device path)."""

    def __init__(self, types, hash_channels, precomputed_hash_channel=None, expected_positions=10_000, ctx=None):
        self.index = PagesIndex(types, hash_channels, precomputed_hash_channel, expected_positions, ctx)
        self._lookup = None
        self._finishing = False

    def needs_input(self):
        return not self._finishing

    def add_input(self, page):   # HashBuilderOperator.java:317-350
        assert not self._finishing
        self.index.add_page(page)

    def finish(self):            # :428-496 -> buildLookupSource
        if self._finishing:
            return
        self._finishing = True
        self._lookup = self.index.create_lookup_source()

    def is_finished(self):
        return self._lookup is not None

    def get_output(self):
        return None

    def get_lookup_source(self):
        assert self._lookup is not None, "Lookup source has not been built yet"
        return self._lookup


class LookupJoinOperator:
    """Probe operator: INNER / PROBE_OUTER.  Output page = probe output channels (taken at the emitted probe positions, the
    reference's getPositions view, LookupJoinPageBuilder.java:101-133) followed by the build output channels."""

    def __init__(self, lookup_source, probe_join_channels, probe_output_channels, build_output_channels, join_type=N.JOIN_INNER,
                 probe_hash_channel=None, device_output=None):
        self.lookup = lookup_source
        self.probe_join_channels = list(probe_join_channels)
        self.probe_output_channels = list(probe_output_channels)
        self.build_output_channels = list(build_output_channels)
        self.join_type = join_type
        self.probe_hash_channel = probe_hash_channel
        self.device_output = device_output
        self._pending = None
        self._finishing = False

    def needs_input(self):
        return not self._finishing and self._pending is None

    def add_input(self, page):
        assert self.needs_input()
        self._pending = page

    def finish(self):
        self._finishing = True

    def is_finished(self):
        return self._finishing and self._pending is None

    def get_output(self):
        if self._pending is None:
            return None
        page, self._pending = self._pending, None
        keys = [page.get_block(c) for c in self.probe_join_channels]
        hb = page.get_block(self.probe_hash_channel) if self.probe_hash_channel is not None else None
        on_device = keys[0].device if self.device_output is None else self.device_output
        probe_pos, build_pos = self.lookup.join_page(keys, hb, self.join_type, device=on_device)
        n = len(probe_pos)
        out = [take_positions(page.get_block(c), probe_pos) for c in self.probe_output_channels]
        out += self.lookup.append_to(build_pos, self.build_output_channels, device=on_device)
        return Page(*out, position_count=n)


def take_positions(block, positions):
    """Block.getPositions(int[]) — materialised gather of probe-side columns (host numpy or device torch)."""
    n = len(positions)
    if block.device:
        import torch
        idx = positions.long() if type(positions).__module__.startswith("torch") else torch.from_numpy(np.asarray(positions)).cuda().long()
        if block.type in (N.VARCHAR, N.VARBINARY):
            return take_positions(block.to_host(), idx.cpu().numpy()).to_device()
        if block.type == N.LONG_DECIMAL:
            v = block.values.view(-1, 2)[idx].reshape(-1)
        else:
            v = block.values[idx]
        return Block(block.type, v, None if block.nulls is None else block.nulls[idx], None, n)
    idx = positions.cpu().numpy() if type(positions).__module__.startswith("torch") else np.asarray(positions)
    idx = idx.astype(np.int64)
    if block.type in (N.VARCHAR, N.VARBINARY):
        lens = (block.offsets[1:] - block.offsets[:-1])[idx]
        offs = np.zeros(n + 1, dtype=np.int32)
        np.cumsum(lens, out=offs[1:])
        data = np.zeros(int(offs[-1]), dtype=np.uint8)
        for i, p in enumerate(idx):
            data[offs[i]:offs[i + 1]] = block.values[block.offsets[p]:block.offsets[p + 1]]
        return Block(block.type, data, None if block.nulls is None else block.nulls[idx], offs, n)
    if block.type == N.LONG_DECIMAL:
        v = block.values.reshape(-1, 2)[idx].reshape(-1)
    else:
        v = block.values[idx]
    return Block(block.type, np.ascontiguousarray(v), None if block.nulls is None else np.ascontiguousarray(block.nulls[idx]), None, n)


# ------------------------------------------------------------------------------------------------ f4: MarkDistinct, SetBuilder, HashSemiJoin
class MarkDistinctOperator:
    """OP/MarkDistinctOperator.java:109-140: appends a BOOLEAN channel that is true for the first row of every distinct combination
    of the mark channels (MarkDistinctHash over a GroupByHash)."""

    def __init__(self, mark_types, mark_channels, hash_channel=None, expected_size=10_000, ctx=None):
        from .groupby import GroupByHash
        self.channels, self.hash_channel = list(mark_channels), hash_channel
        self.hash = GroupByHash(list(mark_types), expected_size, has_hash_channel=hash_channel is not None, ctx=ctx)
        self._finishing, self._out = False, None

    def needs_input(self):
        return not self._finishing and self._out is None

    def add_input(self, page):
        assert self.needs_input()
        blocks = page.blocks if isinstance(page, Page) else list(page)
        keys = [blocks[c] for c in self.channels] + ([blocks[self.hash_channel]] if self.hash_channel is not None else [])
        marks = self.hash.mark_distinct(keys)
        from .blocks import Block
        self._out = Page(*(list(blocks) + [Block(N.BOOLEAN, marks.astype(np.uint8), None, None, len(marks))]), position_count=len(marks))

    def get_output(self):
        out, self._out = self._out, None
        return out

    def finish(self):
        self._finishing = True

    def is_finished(self):
        return self._finishing and self._out is None


class DistinctLimitOperator:
    """OP/DistinctLimitOperator.java:115-220: passes through the first row of every new distinct combination of the distinct channels
    (plus the precomputed hash channel) until `limit` rows have been produced.  The first-row test is the one MarkDistinct uses
    (group id == nextDistinctId over the page's GroupByIdBlock; pg_groupby_mark_distinct)."""

    def __init__(self, distinct_types, distinct_channels, limit, hash_channel=None, ctx=None):
        from .groupby import GroupByHash
        if limit < 0:
            raise ValueError("limit must be at least zero")                                   # DistinctLimitOperator.java:120
        self.channels, self.hash_channel = list(distinct_channels), hash_channel
        self.output_channels = self.channels + ([hash_channel] if hash_channel is not None else [])
        self.hash = GroupByHash(list(distinct_types), min(limit, 10_000), has_hash_channel=hash_channel is not None, ctx=ctx)
        self.remaining_limit = limit
        self._finishing, self._page, self._marks = False, None, None

    def needs_input(self):
        return not self._finishing and self.remaining_limit > 0 and self._page is None

    def add_input(self, page):
        assert self.needs_input()
        blocks = page.blocks if isinstance(page, Page) else list(page)
        keys = [blocks[c] for c in self.channels] + ([blocks[self.hash_channel]] if self.hash_channel is not None else [])
        self._marks = self.hash.mark_distinct(keys)
        self._page = blocks

    def get_output(self):
        if self._page is None:
            return None
        positions = np.flatnonzero(self._marks)[:self.remaining_limit]                         # stop at the row that exhausts the limit
        self.remaining_limit -= len(positions)
        blocks, self._page, self._marks = self._page, None, None
        if len(positions) == 0:
            return None
        return Page(*[take_positions(blocks[c], positions) for c in self.output_channels], position_count=len(positions))

    def finish(self):
        self._finishing = True

    def is_finished(self):
        return self._page is None and (self._finishing or self.remaining_limit == 0)


class SetBuilderOperator:
    """OP/SetBuilderOperator.java:123-170: collects the build side's join channel into a ChannelSet (a GroupByHash plus 'contains NULL')."""

    def __init__(self, type_, set_channel, hash_channel=None, expected_positions=10_000, ctx=None):
        from .groupby import GroupByHash
        self.type, self.channel, self.hash_channel = type_, set_channel, hash_channel
        self.hash = GroupByHash([type_], expected_positions, has_hash_channel=hash_channel is not None, ctx=ctx)
        self._finishing = False
        self.contains_null = False

    def needs_input(self):
        return not self._finishing

    def add_input(self, page):
        assert self.needs_input()
        blocks = page.blocks if isinstance(page, Page) else list(page)
        keys = [blocks[self.channel]] + ([blocks[self.hash_channel]] if self.hash_channel is not None else [])
        self.hash.add_page(keys)

    def finish(self):
        """ChannelSetBuilder.build(): containsNull = hash.contains(0, nullBlockPage) (ChannelSet.java:101-109)."""
        if self._finishing:
            return
        self._finishing = True
        from .blocks import Block
        if self.type in (N.VARCHAR, N.VARBINARY):
            null_row = Block(self.type, np.zeros(1, dtype=np.uint8), np.ones(1, dtype=np.uint8), np.zeros(2, dtype=np.int32), 1)
        else:
            from .blocks import _NP_DTYPE
            k = 2 if self.type == N.LONG_DECIMAL else 1
            null_row = Block(self.type, np.zeros(k, dtype=np.uint64 if self.type == N.LONG_DECIMAL else _NP_DTYPE[self.type]), np.ones(1, dtype=np.uint8), None, 1)
        self.contains_null = bool(self.hash.contains([null_row])[0])

    def is_finished(self):
        return self._finishing

    def size(self):
        return self.hash.get_group_count()


class HashSemiJoinOperator:
    """OP/HashSemiJoinOperator.java:176-221: appends the BOOLEAN 'probe key IN (set)' channel with SQL three-valued logic."""

    def __init__(self, set_builder, probe_join_channel, probe_hash_channel=None):
        self.set, self.channel, self.hash_channel = set_builder, probe_join_channel, probe_hash_channel
        self._finishing, self._out = False, None

    def needs_input(self):
        return self.set.is_finished() and not self._finishing and self._out is None

    def add_input(self, page):
        assert self.needs_input()
        blocks = page.blocks if isinstance(page, Page) else list(page)
        hb = blocks[self.hash_channel] if self.hash_channel is not None else None
        vals, nulls = self.set.hash.semi_join(blocks[self.channel], hb, self.set.contains_null)
        from .blocks import Block
        self._out = Page(*(list(blocks) + [Block(N.BOOLEAN, vals, nulls if nulls.any() else None, None, len(vals))]), position_count=len(vals))

    def get_output(self):
        out, self._out = self._out, None
        return out

    def finish(self):
        self._finishing = True

    def is_finished(self):
        return self._finishing and self._out is None


class LookupOuterOperator:
    """OP/LookupOuterOperator.java:140-181: after build and probe have finished, one row per build position that no probe emitted —
    NULL in the probe output channels, the build output channels from the lookup source.  The lookup source must have been
    switched to tracking (LookupSource.track_visited) before the first probe, as JoinBridgeManager wraps it in OuterLookupSource
    for RIGHT / FULL joins."""

    def __init__(self, lookup_source, probe_output_types, build_output_channels, page_rows=8192):
        self.lookup, self.probe_types, self.build_channels, self.page_rows = lookup_source, list(probe_output_types), list(build_output_channels), page_rows
        self._positions, self._at = None, 0

    def needs_input(self):
        return False

    def get_output(self):
        if self._positions is None:
            self._positions = self.lookup.outer_positions()
        if self._at >= len(self._positions):
            return None
        pos = self._positions[self._at:self._at + self.page_rows]
        self._at += len(pos)
        n = len(pos)
        nulls = []
        for t in self.probe_types:
            o = OutBlock(t, n, 1)
            b = o.block(n)
            nulls.append(Block(t, b.values, np.ones(n, dtype=np.uint8), b.offsets, n))
        build = self.lookup.append_to(pos, self.build_channels, device=False)
        return Page(*(nulls + build), position_count=n)

    def is_finished(self):
        return self._positions is not None and self._at >= len(self._positions)
