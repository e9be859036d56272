"""GroupedAccumulator — host mirror of MAIN/operator/aggregation/GroupedAccumulator.java:22-39 over pg_agg_*.

add_input(GroupByIdBlock, Page) / add_intermediate(GroupByIdBlock, Block...) / evaluate_final / evaluate_intermediate,
for count, count(col), sum, avg, min, max over bigint/integer/date/double/short and long decimal.
"""
import ctypes as C

from . import _native as N
from .blocks import Block, OutBlock
from .context import default_context

KIND_NAMES = {N.AGG_COUNT: "count", N.AGG_COUNT_COL: "count", N.AGG_SUM: "sum", N.AGG_AVG: "avg", N.AGG_MIN: "min", N.AGG_MAX: "max"}


def _is_dec(t):
    return t in (N.SHORT_DECIMAL, N.LONG_DECIMAL)


def final_type(kind, in_type):
    """getFinalType()."""
    if kind in (N.AGG_COUNT, N.AGG_COUNT_COL):
        return N.BIGINT
    if kind == N.AGG_SUM:
        if _is_dec(in_type):
            return N.LONG_DECIMAL   # decimal(38, s), DecimalSumAggregation.java:58-75
        return N.DOUBLE if in_type == N.DOUBLE else N.BIGINT
    if kind == N.AGG_AVG:
        return in_type if _is_dec(in_type) else N.DOUBLE
    return in_type


def intermediate_types(kind, in_type):
    """getIntermediateType() flattened to columns (see the layout table in include/prestogpu.h)."""
    if kind in (N.AGG_COUNT, N.AGG_COUNT_COL):
        return [N.BIGINT]
    if kind in (N.AGG_SUM, N.AGG_AVG) and _is_dec(in_type):
        return [N.VARBINARY]
    if kind == N.AGG_SUM:
        return [N.BIGINT, N.DOUBLE if in_type == N.DOUBLE else N.BIGINT]
    if kind == N.AGG_AVG:
        return [N.DOUBLE, N.BIGINT]
    return [in_type]


class GroupedAccumulator:
    def __init__(self, kind, in_type=0, ctx=None):
        self.ctx = ctx or default_context()
        self.kind = kind
        self.in_type = in_type
        self._h = C.c_void_p()
        N.check(self.ctx.handle, N.lib().pg_agg_create(self.ctx.handle, kind, in_type, C.byref(self._h)))

    @property
    def handle(self):
        return self._h

    def get_final_type(self):
        return final_type(self.kind, self.in_type)

    def get_intermediate_type(self):
        return intermediate_types(self.kind, self.in_type)

    def get_estimated_size(self):
        b = C.c_int64()
        N.check(self.ctx.handle, N.lib().pg_agg_device_bytes(self._h, C.byref(b)))
        return b.value

    @staticmethod
    def _ids(group_ids):
        ids = group_ids.ids
        if type(ids).__module__.startswith("torch"):
            return ids.data_ptr() if len(ids) else None, N.MEM_DEVICE, len(ids)
        return ids.ctypes.data, N.MEM_HOST, len(ids)

    def add_input(self, group_ids, arg=None, mask=None):
        """addInput(GroupByIdBlock groupIdsBlock, Page page): arg = the argument block, mask = optional BOOLEAN block."""
        ptr, mem, n = self._ids(group_ids)
        a = arg.pg_col() if arg is not None else None
        m = mask.pg_col() if mask is not None else None
        N.check(self.ctx.handle, N.lib().pg_agg_add_input(self._h, ptr, mem, group_ids.group_count, C.byref(a) if a else None,
                                                          C.byref(m) if m else None, n))

    def add_intermediate(self, group_ids, state_blocks):
        ptr, mem, n = self._ids(group_ids)
        arr = (N.PgCol * len(state_blocks))()
        for i, b in enumerate(state_blocks):
            arr[i] = b.pg_col()
        N.check(self.ctx.handle, N.lib().pg_agg_add_intermediate(self._h, ptr, mem, group_ids.group_count, arr, n))

    def _evaluate(self, first, n, final, device):
        types = [self.get_final_type()] if final else self.get_intermediate_type()
        outs = [OutBlock(t, n, value_bytes=32 * n if t == N.VARBINARY else 0, device=device) for t in types]
        arr = (N.PgColOut * len(outs))()
        for i, o in enumerate(outs):
            arr[i] = o.pg_col_out()
        eg = C.c_int64(-1)
        N.check(self.ctx.handle, N.lib().pg_agg_evaluate(self._h, first, n, 1 if final else 0, arr, C.byref(eg)))
        return [o.block(n) for o in outs]

    def evaluate_final(self, first, n, device=False):
        """evaluateFinal(groupId, output) for groupId in [first, first+n)."""
        return self._evaluate(first, n, True, device)[0]

    def evaluate_intermediate(self, first, n, device=False):
        return self._evaluate(first, n, False, device)

    def close(self):
        if self._h:
            N.lib().pg_agg_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
