"""Join build side and LookupSource — host mirror of MAIN/operator/PagesIndex.java (addPage / createLookupSourceSupplier),
MAIN/operator/LookupSource.java:23-50 and MAIN/operator/JoinProbe.java:26-118 over pg_joinbuild_* / pg_lookup_*.

The reference walks one probe row at a time (getJoinPosition / getNextJoinPosition / appendTo); here a whole probe page is
joined per call and the (probe position, build position) pairs come back in exactly the order that walk would emit.
"""
import ctypes as C

import numpy as np

from . import _native as N
from .blocks import Block, OutBlock, Page, col_array
from .context import default_context


class PagesIndex:
    """Accumulates build pages on the device (PagesIndex.addPage :209-240)."""

    def __init__(self, types, join_channels, hash_channel=None, expected_positions=10_000, ctx=None):
        self.ctx = ctx or default_context()
        self.types = list(types)
        self.join_channels = list(join_channels)
        self.hash_channel = -1 if hash_channel is None else int(hash_channel)
        self._h = C.c_void_p()
        t = (C.c_int32 * len(self.types))(*self.types)
        j = (C.c_int32 * len(self.join_channels))(*self.join_channels)
        N.check(self.ctx.handle, N.lib().pg_joinbuild_create(self.ctx.handle, t, len(self.types), j, len(self.join_channels),
                                                             self.hash_channel, expected_positions, C.byref(self._h)))
        self.position_count = 0

    def add_page(self, page):
        blocks = page.blocks if isinstance(page, Page) else list(page)
        n = blocks[0].n
        N.check(self.ctx.handle, N.lib().pg_joinbuild_add_page(self._h, col_array(blocks), n))
        self.position_count += n

    def get_position_count(self):
        return self.position_count

    def create_lookup_source(self):
        """createLookupSourceSupplier(...).get(): builds PagesHash + position links on the device."""
        h = C.c_void_p()
        N.check(self.ctx.handle, N.lib().pg_joinbuild_finish(self._h, C.byref(h)))
        self.position_count = 0
        return LookupSource(self.ctx, h, self.types, self.join_channels)

    def close(self):
        if self._h:
            N.lib().pg_joinbuild_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class LookupSource:
    def __init__(self, ctx, handle, types, join_channels):
        self.ctx, self._h, self.types, self.join_channels = ctx, handle, list(types), list(join_channels)

    def _stats(self):
        r, hs, db, d = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int32()
        N.check(self.ctx.handle, N.lib().pg_lookup_stats(self._h, C.byref(r), C.byref(hs), C.byref(db), C.byref(d)))
        return r.value, hs.value, db.value, bool(d.value)

    def get_channel_count(self):
        return len(self.types)

    def get_join_position_count(self):
        return self._stats()[0]

    def get_in_memory_size_in_bytes(self):
        return self._stats()[2]

    def is_empty(self):
        return self._stats()[0] == 0

    def has_duplicates(self):
        return self._stats()[3]

    def join_page(self, probe_keys, hash_block=None, join_type=N.JOIN_INNER, device=None, capacity=None):
        """JoinProbe over one page: returns (probe positions int32[], build positions int64[]) in reference order."""
        keys = list(probe_keys)
        n = keys[0].n
        on_device = keys[0].device if device is None else device
        hc = hash_block.pg_col() if hash_block is not None else None
        cap = max(n, 1) if capacity is None else capacity
        n_out = C.c_int64()
        while True:
            if on_device:
                import torch
                op = torch.empty(max(cap, 1), dtype=torch.int32, device="cuda")
                ob = torch.empty(max(cap, 1), dtype=torch.int64, device="cuda")
                pp, pb, mem = op.data_ptr(), ob.data_ptr(), N.MEM_DEVICE
            else:
                op = np.empty(max(cap, 1), dtype=np.int32)
                ob = np.empty(max(cap, 1), dtype=np.int64)
                pp, pb, mem = op.ctypes.data, ob.ctypes.data, N.MEM_HOST
            rc = N.lib().pg_lookup_probe(self._h, col_array(keys), len(keys), C.byref(hc) if hc else None, n, join_type, pp, pb, mem, cap,
                                         C.byref(n_out))
            if rc == N.PG_NEED_MEMORY:   # nothing consumed: retry with room for every match
                cap = n_out.value
                continue
            N.check(self.ctx.handle, rc)
            return op[:n_out.value], ob[:n_out.value]

    def join_and_append(self, probe_keys, channels, hash_block=None, join_type=N.JOIN_INNER, device=None, capacity=None):
        """One LookupJoinOperator step: (probe positions int32[], build channels `channels` of the matched rows) — JoinProbe and
        appendTo in one call (pg_lookup_join_page).  Fixed width build channels only (use join_page + append_to otherwise)."""
        keys = list(probe_keys)
        n = keys[0].n
        on_device = keys[0].device if device is None else device
        hc = hash_block.pg_col() if hash_block is not None else None
        cap = max(n, 1) if capacity is None else capacity
        n_out = C.c_int64()
        chs = (C.c_int32 * max(1, len(channels)))(*channels)
        for ch in channels:
            if self.types[ch] in (N.VARCHAR, N.VARBINARY):
                raise N.PrestoGpuError(N.PG_E_INVALID_ARGUMENT, "join_and_append: variable width build channel %d; use join_page + append_to" % ch)
        while True:
            if on_device:
                import torch
                op = torch.empty(max(cap, 1), dtype=torch.int32, device="cuda")
                pp, mem = op.data_ptr(), N.MEM_DEVICE
            else:
                op = np.empty(max(cap, 1), dtype=np.int32)
                pp, mem = op.ctypes.data, N.MEM_HOST
            outs = [OutBlock(self.types[ch], max(cap, 1), 0, device=on_device, zero=False) for ch in channels]
            arr = (N.PgColOut * max(1, len(outs)))()
            for i, o in enumerate(outs):
                arr[i] = o.pg_col_out()
            rc = N.lib().pg_lookup_join_page(self._h, col_array(keys), len(keys), C.byref(hc) if hc else None, n, join_type, chs, len(channels),
                                             pp, mem, arr, cap, C.byref(n_out))
            if rc == N.PG_NEED_MEMORY:   # nothing consumed: retry with room for every match
                cap = n_out.value
                continue
            N.check(self.ctx.handle, rc)
            return op[:n_out.value], [o.block(n_out.value) for o in outs]

    def track_visited(self):
        """OuterLookupSource: remember which build positions get emitted (RIGHT / FULL joins); call before the first probe."""
        N.check(self.ctx.handle, N.lib().pg_lookup_track_visited(self._h))

    def outer_positions(self):
        """OuterPositionIterator: build positions no probe has emitted, ascending (numpy int64)."""
        cap = max(1, self._stats()[0])
        out = np.empty(cap, dtype=np.int64)
        n_out = C.c_int64()
        N.check(self.ctx.handle, N.lib().pg_lookup_outer_positions(self._h, out.ctypes.data, N.MEM_HOST, cap, C.byref(n_out)))
        return out[:n_out.value]

    def append_to(self, build_positions, channels, device=None):
        """appendTo for a vector of build positions: gathers the given build channels (position -1 -> null)."""
        on_device = type(build_positions).__module__.startswith("torch") if device is None else device
        n = len(build_positions)
        if type(build_positions).__module__.startswith("torch"):
            pptr, pmem = (build_positions.data_ptr() if n else None), N.MEM_DEVICE
        else:
            build_positions = np.ascontiguousarray(build_positions, dtype=np.int64)
            pptr, pmem = build_positions.ctypes.data, N.MEM_HOST
        outs = []
        for ch in channels:
            vb = 0
            if self.types[ch] in (N.VARCHAR, N.VARBINARY):
                b = C.c_int64()
                N.check(self.ctx.handle, N.lib().pg_lookup_gather_var_bytes(self._h, pptr, pmem, n, ch, C.byref(b)))
                vb = b.value
            outs.append(OutBlock(self.types[ch], n, vb, device=on_device))
        arr = (N.PgColOut * max(1, len(outs)))()
        for i, o in enumerate(outs):
            arr[i] = o.pg_col_out()
        chs = (C.c_int32 * max(1, len(channels)))(*channels)
        N.check(self.ctx.handle, N.lib().pg_lookup_gather(self._h, pptr, pmem, n, chs, len(channels), arr))
        return [o.block(n) for o in outs]

    def close(self):
        if self._h:
            N.lib().pg_lookup_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
