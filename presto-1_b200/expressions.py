"""RowExpressions and the PageProcessor — host mirror of MAIN/sql/relational/{InputReferenceExpression, ConstantExpression,
CallExpression, SpecialForm}.java as MAIN/operator/project/PageProcessor.java:54-57,139-330 evaluates them (filter first, then the
projections over the selected positions), over pg_project_* (include/prestogpu.h).

    e = Expr
    price, disc = e.input(3, P.SHORT_DECIMAL, 12, 2), e.input(4, P.SHORT_DECIMAL, 12, 2)
    disc_price = price * (e.decimal(1, 1, 0) - disc)                  # decimal(25,4), exact, like DecimalOperators.java
    proc = PageProcessor(filter=e.input(5, P.DATE) <= e.date(10471), projections=[e.input(0, P.VARCHAR), disc_price])
    out_page = proc.process(page)                                     # ScanFilterAndProjectOperator's page-at-a-time step
    HashAggregationOperator(..., input_processor=proc)                # the same expressions fused into the aggregation kernel
"""
import ctypes as C

import numpy as np

from . import _native as N
from .blocks import Block, OutBlock, Page, col_array, encode_decimal128
from .context import default_context


class Expr:
    """One node of an expression DAG.  Operators build CallExpressions; result types are computed by the library."""

    def __init__(self, op, type_=0, precision=0, scale=0, channel=-1, args=(), value=None, is_null=False):
        self.op, self.type, self.precision, self.scale, self.channel = op, type_, precision, scale, channel
        self.args, self.value, self.is_null = tuple(args), value, is_null

    # ---- leaves
    @staticmethod
    def input(channel, type_, precision=0, scale=0):
        return Expr(N.EXPR_INPUT, type_, precision, scale, channel)

    @staticmethod
    def const(value, type_, precision=0, scale=0):
        return Expr(N.EXPR_CONST, type_, precision, scale, value=value, is_null=value is None)

    @staticmethod
    def bigint(v):
        return Expr.const(v, N.BIGINT)

    @staticmethod
    def integer(v):
        return Expr.const(v, N.INTEGER)

    @staticmethod
    def date(days):
        return Expr.const(days, N.DATE)

    @staticmethod
    def double(v):
        return Expr.const(v, N.DOUBLE)

    @staticmethod
    def decimal(unscaled, precision, scale):
        return Expr.const(unscaled, N.SHORT_DECIMAL if precision <= 18 else N.LONG_DECIMAL, precision, scale)

    @staticmethod
    def varchar(s):
        return Expr.const(s.encode() if isinstance(s, str) else bytes(s), N.VARCHAR)

    # ---- calls
    def _bin(self, op, other):
        return Expr(op, args=(self, other))

    def __add__(self, o):
        return self._bin(N.EXPR_ADD, o)

    def __sub__(self, o):
        return self._bin(N.EXPR_SUB, o)

    def __mul__(self, o):
        return self._bin(N.EXPR_MUL, o)

    def __lt__(self, o):
        return self._bin(N.EXPR_LT, o)

    def __le__(self, o):
        return self._bin(N.EXPR_LE, o)

    def __gt__(self, o):
        return self._bin(N.EXPR_GT, o)

    def __ge__(self, o):
        return self._bin(N.EXPR_GE, o)

    def eq(self, o):
        return self._bin(N.EXPR_EQ, o)

    def ne(self, o):
        return self._bin(N.EXPR_NE, o)

    def __and__(self, o):
        return self._bin(N.EXPR_AND, o)

    def __or__(self, o):
        return self._bin(N.EXPR_OR, o)

    def __invert__(self):
        return Expr(N.EXPR_NOT, args=(self,))

    def is_null_(self):
        return Expr(N.EXPR_IS_NULL, args=(self,))

    def cast(self, type_, precision=0, scale=0):
        return Expr(N.EXPR_CAST, type_, precision, scale, args=(self,))

    @staticmethod
    def hash_of(*exprs):
        """The $hashvalue projection: combine_hash folded over the channels (HashGenerationOptimizer.java:98,872,926)."""
        h = None
        for x in exprs:
            h = Expr(N.EXPR_HASH, args=(h, x))
        return h


def flatten(roots):
    """Expression DAG -> (ctypes node array, {id(expr): index}, keep-alive list); operands precede their users."""
    order, index = [], {}

    def visit(e):
        if e is None or id(e) in index:
            return
        for a in e.args:
            visit(a)
        index[id(e)] = len(order)
        order.append(e)

    for r in roots:
        visit(r)
    arr = (N.PgExpr * max(1, len(order)))()
    keep = []
    for i, e in enumerate(order):
        x = arr[i]
        x.op, x.type, x.precision, x.scale, x.channel = e.op, e.type, e.precision, e.scale, e.channel
        x.arg0 = index[id(e.args[0])] if len(e.args) > 0 and e.args[0] is not None else -1
        x.arg1 = index[id(e.args[1])] if len(e.args) > 1 and e.args[1] is not None else -1
        x.is_null = 1 if e.is_null else 0
        if e.op == N.EXPR_CONST and e.value is not None:
            if e.type in (N.VARCHAR, N.VARBINARY):
                buf = C.create_string_buffer(e.value, len(e.value))
                keep.append(buf)
                x.bytes, x.nbytes = C.cast(buf, C.c_void_p).value, len(e.value)
            elif e.type == N.LONG_DECIMAL:
                lo, hi = encode_decimal128(int(e.value))
                x.lo, x.hi = np.uint64(lo).astype(np.int64).item(), np.uint64(hi).astype(np.int64).item()
            elif e.type == N.DOUBLE:
                x.lo = np.float64(e.value).view(np.int64).item()
            elif e.type == N.BOOLEAN:
                x.lo = 1 if e.value else 0
            else:
                x.lo = int(e.value)
    return arr, index, keep, len(order)


class PageProcessor:
    """Optional filter + projections.  process(page) is PageProcessor.process: the projections over the rows that pass the filter."""

    def __init__(self, filter=None, projections=(), ctx=None):
        self.ctx = ctx or default_context()
        self.filter, self.projections = filter, list(projections)
        roots = ([filter] if filter is not None else []) + self.projections
        arr, index, keep, n = flatten(roots)
        self._keep = (arr, keep)
        proj = (C.c_int32 * max(1, len(self.projections)))(*[index[id(p)] for p in self.projections])
        self._h = C.c_void_p()
        N.check(self.ctx.handle, N.lib().pg_project_create(self.ctx.handle, arr, n, index[id(filter)] if filter is not None else -1, proj,
                                                           len(self.projections), C.byref(self._h)))
        k = len(self.projections)
        t, p, s, nc = (C.c_int32 * max(1, k))(), (C.c_int32 * max(1, k))(), (C.c_int32 * max(1, k))(), C.c_int32()
        N.check(self.ctx.handle, N.lib().pg_project_output_types(self._h, t, p, s, k, C.byref(nc)))
        self.output_types = [(t[i], p[i], s[i]) for i in range(k)]

    @property
    def handle(self):
        return self._h

    def process(self, page, device=None):
        blocks = page.blocks if isinstance(page, Page) else list(page)
        n = blocks[0].n
        on_device = blocks[0].device if device is None else device
        outs = []
        for (t, p, s), e in zip(self.output_types, self.projections):
            vb = 0
            if t in (N.VARCHAR, N.VARBINARY):
                src = blocks[e.channel]
                vb = int(src.values.numel() if src.device else len(src.values))
            outs.append(OutBlock(t, n, vb, device=on_device, zero=not on_device))     # (the kernels write every selected row of every array)
        arr = (N.PgColOut * max(1, len(outs)))()
        for i, o in enumerate(outs):
            arr[i] = o.pg_col_out()
        n_out = C.c_int64()
        N.check(self.ctx.handle, N.lib().pg_project_process(self._h, col_array(blocks), len(blocks), n, arr, C.byref(n_out)))
        m = n_out.value
        res = []
        for o, (t, p, s) in zip(outs, self.output_types):
            b = o.block(m)
            b.precision, b.scale = p, s
            res.append(b)
        return Page(*res, position_count=m)

    def close(self):
        if self._h:
            N.lib().pg_project_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ScanFilterAndProjectOperator:
    """Operator shell around a PageProcessor (OP/ScanFilterAndProjectOperator.java; page source = whatever calls add_input)."""

    def __init__(self, processor, device_output=None):
        self.processor, self.device_output = processor, device_output
        self._pending, self._finishing = None, False

    def needs_input(self):
        return not self._finishing and self._pending is None

    def add_input(self, page):
        assert self.needs_input()
        self._pending = page

    def get_output(self):
        if self._pending is None:
            return None
        page, self._pending = self._pending, None
        out = self.processor.process(page, device=self.device_output)
        return out if out.position_count > 0 else None

    def finish(self):
        self._finishing = True

    def is_finished(self):
        return self._finishing and self._pending is None
