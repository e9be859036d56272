"""GroupByHash — host mirror of MAIN/operator/GroupByHash.java:31-88 over the C ABI (pg_groupby_*).

Method names follow the reference interface: addPage -> add_page, getGroupIds -> get_group_ids,
appendValuesTo -> append_values_to, contains, getRawHash -> get_raw_hash, getGroupCount, getCapacity.
The reference returns Work<T> objects that may yield on memory pressure; device tables are resized between
kernels, so the work always completes in one process() call here.
"""
import ctypes as C

import numpy as np

from . import _native as N
from .blocks import Block, OutBlock, Page, col_array
from .context import default_context


class GroupByIdBlock:
    """MAIN/operator/GroupByIdBlock.java:30-49: groupCount + one group id per input row."""

    def __init__(self, group_count, ids):
        self.group_count = group_count
        self.ids = ids  # numpy int32 (host) or torch int32 (device)

    def get_group_count(self):
        return self.group_count

    def get_group_id(self, position):
        return int(self.ids[position])

    def get_position_count(self):
        return len(self.ids)


class Work:
    """OP/Work.java: process() -> finished?; get_result() once finished.  A PG_NEED_MEMORY status is the yield."""

    def __init__(self, step):
        self._step, self._done, self._result = step, False, None

    def process(self):
        assert not self._done, "process() called on a finished Work"
        try:
            self._result = self._step()
        except N.PrestoGpuError as e:
            if e.code == N.PG_NEED_MEMORY:
                return False
            raise
        self._done = True
        return True

    def get_result(self):
        assert self._done, "get_result() before the Work finished"
        return self._result


class GroupByHash:
    def __init__(self, types, expected_size, has_hash_channel=False, ctx=None):
        self.ctx = ctx or default_context()
        self.types = list(types)
        self.has_hash_channel = bool(has_hash_channel)
        self._h = C.c_void_p()
        arr = (C.c_int32 * len(self.types))(*self.types)
        N.check(self.ctx.handle, N.lib().pg_groupby_create(self.ctx.handle, arr, len(self.types), int(self.has_hash_channel),
                                                           int(expected_size), C.byref(self._h)))

    # GroupByHash.createGroupByHash (GroupByHash.java:45-59) — the dispatch itself happens in the library
    @staticmethod
    def create_group_by_hash(types, expected_size, has_hash_channel=False, ctx=None):
        return GroupByHash(types, expected_size, has_hash_channel, ctx)

    @property
    def handle(self):
        return self._h

    def _split(self, page):
        """page: Page or list of key blocks, with the precomputed hash block last when has_hash_channel."""
        blocks = page.blocks if isinstance(page, Page) else list(page)
        nk = len(self.types)
        keys = blocks[:nk]
        hash_block = blocks[nk] if self.has_hash_channel and len(blocks) > nk else None
        return keys, hash_block

    def set_memory_limit(self, nbytes):
        """UpdateMemory as a byte limit (OP/UpdateMemory.java): with a limit set, the *_work forms below yield instead of growing past it."""
        N.check(self.ctx.handle, N.lib().pg_groupby_set_memory_limit(self._h, int(nbytes)))

    def add_page_work(self, page):
        """GroupByHash.addPage(page) as the reference returns it: a Work whose process() is False while the memory limit blocks the
        page (nothing consumed) and True once the page is in (OP/GroupByHash.java:73, OP/Work.java)."""
        return Work(lambda: self.add_page(page))

    def get_group_ids_work(self, page, device_ids=None):
        """GroupByHash.getGroupIds(page) as a Work<GroupByIdBlock> (OP/GroupByHash.java:75)."""
        return Work(lambda: self.get_group_ids(page, device_ids))

    def add_page(self, page):
        keys, hb = self._split(page)
        n = keys[0].n
        hc = hb.pg_col() if hb is not None else None
        N.check(self.ctx.handle, N.lib().pg_groupby_add_page(self._h, col_array(keys), C.byref(hc) if hc else None, n))

    def get_group_ids(self, page, device_ids=None):
        keys, hb = self._split(page)
        n = keys[0].n
        on_device = keys[0].device if device_ids is None else device_ids
        hc = hb.pg_col() if hb is not None else None
        gc = C.c_int64()
        if on_device:
            import torch
            ids = torch.empty(max(n, 1), dtype=torch.int32, device="cuda")[:n]
            ptr = ids.data_ptr() if n else None
        else:
            ids = np.empty(n, dtype=np.int32)
            ptr = ids.ctypes.data
        N.check(self.ctx.handle, N.lib().pg_groupby_get_group_ids(self._h, col_array(keys), C.byref(hc) if hc else None, n, ptr,
                                                                  N.MEM_DEVICE if on_device else N.MEM_HOST, C.byref(gc)))
        return GroupByIdBlock(gc.value, ids)

    def mark_distinct(self, page):
        """MarkDistinctHash.markDistinctRows(page) (OP/MarkDistinctHash.java:52-69): numpy bool per row, True for the first row of every
        group this page introduced."""
        keys, hb = self._split(page)
        n = keys[0].n
        hc = hb.pg_col() if hb is not None else None
        out = np.zeros(n, dtype=np.uint8)
        N.check(self.ctx.handle, N.lib().pg_groupby_mark_distinct(self._h, col_array(keys), C.byref(hc) if hc else None, n, out.ctypes.data, N.MEM_HOST))
        return out.astype(bool)

    def semi_join(self, probe_block, hash_block=None, set_contains_null=False):
        """The BOOLEAN column HashSemiJoinOperator appends (OP/HashSemiJoinOperator.java:190-218): (values, nulls) numpy arrays."""
        n = probe_block.n
        hc = hash_block.pg_col() if hash_block is not None else None
        vals, nulls = np.zeros(n, dtype=np.uint8), np.zeros(n, dtype=np.uint8)
        N.check(self.ctx.handle, N.lib().pg_groupby_semi_join(self._h, col_array([probe_block]), C.byref(hc) if hc else None, n, int(bool(set_contains_null)),
                                                              vals.ctypes.data, nulls.ctypes.data, N.MEM_HOST))
        return vals, nulls

    def contains(self, page):
        """contains(position, page, hashChannels) for every position; returns a numpy bool array."""
        keys, hb = self._split(page)
        n = keys[0].n
        hc = hb.pg_col() if hb is not None else None
        out = np.zeros(n, dtype=np.uint8)
        N.check(self.ctx.handle, N.lib().pg_groupby_contains(self._h, col_array(keys), C.byref(hc) if hc else None, n, out.ctypes.data,
                                                             N.MEM_HOST))
        return out.astype(bool)

    def _stats(self):
        g, cap, col, db = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        N.check(self.ctx.handle, N.lib().pg_groupby_stats(self._h, C.byref(g), C.byref(cap), C.byref(col), C.byref(db)))
        return g.value, cap.value, col.value, db.value

    def get_group_count(self):
        return self._stats()[0]

    def get_capacity(self):
        return self._stats()[1]

    def get_estimated_size(self):
        return self._stats()[3]

    def get_types(self):
        return self.types + ([N.BIGINT] if self.has_hash_channel else [])

    def append_values_to(self, first=0, n=None, device=False, with_hash=None):
        """appendValuesTo(groupId, pageBuilder, 0) for groupId in [first, first+n): returns (key blocks, raw hashes|None)."""
        if n is None:
            n = self.get_group_count() - first
        with_hash = self.has_hash_channel if with_hash is None else with_hash
        outs = []
        for k, t in enumerate(self.types):
            vb = 0
            if t in (N.VARCHAR, N.VARBINARY):
                b = C.c_int64()
                N.check(self.ctx.handle, N.lib().pg_groupby_varwidth_bytes(self._h, k, C.byref(b)))
                vb = b.value
            outs.append(OutBlock(t, n, vb, device=device))
        arr = (N.PgColOut * len(outs))()
        for i, o in enumerate(outs):
            arr[i] = o.pg_col_out()
        hashes, hptr, hmem = None, None, N.MEM_HOST
        if with_hash:
            if device:
                import torch
                hashes = torch.zeros(max(n, 1), dtype=torch.int64, device="cuda")[:n]
                hptr, hmem = hashes.data_ptr(), N.MEM_DEVICE
            else:
                hashes = np.zeros(n, dtype=np.int64)
                hptr = hashes.ctypes.data
        N.check(self.ctx.handle, N.lib().pg_groupby_export_keys(self._h, first, n, arr, hptr, hmem))
        return [o.block(n) for o in outs], hashes

    def get_raw_hash(self, group_id):
        _, h = self.append_values_to(group_id, 1, with_hash=True)
        return int(h[0])

    def close(self):
        if self._h:
            N.lib().pg_groupby_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
