"""Shared builders for the PageProcessor tests (CPU: oracle vs the reference's vectors; GPU: CUDA vs the oracle)."""
import json
import os

import numpy as np

import presto_b200 as P
from presto_b200 import Expr as E

HERE = os.path.dirname(os.path.abspath(__file__))


def decimal_vectors():
    return json.load(open(os.path.join(HERE, "golden", "decimal_operator_vectors.json")))["cases"]


def dec_type(precision):
    return P.SHORT_DECIMAL if precision <= 18 else P.LONG_DECIMAL


def dec_block(values, precision, scale, nulls=None):
    """values: python ints (unscaled)."""
    if precision <= 18:
        b = P.Block.long_array(np.array([0 if v is None else v for v in values], dtype=np.int64),
                               None if nulls is None and not any(v is None for v in values) else [v is None for v in values] if nulls is None else nulls,
                               P.SHORT_DECIMAL)
    else:
        b = P.Block.int128_array(values, nulls)
    b.precision, b.scale = precision, scale
    return b


def vector_case(c):
    """One reference vector as (page, expression over its two input channels)."""
    a, b = c["a"], c["b"]
    page = P.Page(dec_block([a["unscaled"]], a["precision"], a["scale"]), dec_block([b["unscaled"]], b["precision"], b["scale"]))
    x = E.input(0, dec_type(a["precision"]), a["precision"], a["scale"])
    y = E.input(1, dec_type(b["precision"]), b["precision"], b["scale"])
    return page, {"+": x + y, "-": x - y, "*": x * y}[c["op"]]


def random_page(rng, n, null_rate=0.1):
    """A page with one column of every fixed-width type the expressions handle, plus a short string column."""
    def nulls():
        return (rng.random(n) < null_rate).astype(np.uint8) if null_rate > 0 else None
    big = [int(x) * 10 ** 17 + int(y) for x, y in zip(rng.integers(-10 ** 12, 10 ** 12, n), rng.integers(0, 10 ** 17, n))]
    words = [None if rng.random() < null_rate else ["BUILDING", "MACHINERY", "AUTOMOBILE", "", "B"][i] for i in rng.integers(0, 5, n)]
    return P.Page(
        P.Block.long_array(rng.integers(-10 ** 9, 10 ** 9, n), nulls()),                                    # 0 bigint
        P.Block.int_array(rng.integers(8000, 11000, n).astype(np.int32), nulls(), P.DATE),                  # 1 date
        dec_block([int(v) for v in rng.integers(-10 ** 11, 10 ** 11, n)], 12, 2),                           # 2 decimal(12,2)
        dec_block([int(v) for v in rng.integers(0, 11, n)], 12, 2),                                         # 3 decimal(12,2) small
        dec_block(big, 30, 4, nulls()),                                                                     # 4 decimal(30,4)
        P.Block.double_array(rng.normal(size=n) * 1e3, nulls()),                                            # 5 double
        P.Block.variable_width(words),                                                                      # 6 varchar
        P.Block.int_array(rng.integers(-1000, 1000, n).astype(np.int32), nulls()),                          # 7 integer
        P.Block.byte_array((rng.random(n) < 0.5).astype(np.uint8), nulls()),                                # 8 boolean
    )


def processors():
    """(name, filter, projections) triples over random_page's channels."""
    big, date = E.input(0, P.BIGINT), E.input(1, P.DATE)
    price, disc = E.input(2, P.SHORT_DECIMAL, 12, 2), E.input(3, P.SHORT_DECIMAL, 12, 2)
    wide, dbl = E.input(4, P.LONG_DECIMAL, 30, 4), E.input(5, P.DOUBLE)
    seg, i32, flag = E.input(6, P.VARCHAR), E.input(7, P.INTEGER), E.input(8, P.BOOLEAN)
    one = E.decimal(1, 1, 0)
    disc_price = price * (one - disc)                 # decimal(25,4)
    charge = disc_price * (one + disc)                # decimal(38,6)
    return [
        ("q1_shape", date <= E.date(10471), [seg, price, disc_price, charge, disc]),
        ("no_filter_passthrough", None, [big, seg, wide, dbl, i32, flag]),
        ("varchar_eq_and_date", seg.eq(E.varchar("BUILDING")) & (date < E.date(9500)), [big, date, seg]),
        ("three_valued", (flag | (dbl > E.double(0.0))) & ~(i32 < E.integer(-900)), [flag, dbl, i32, flag.is_null_(), (dbl > E.double(0.0)) | flag]),
        ("decimal_mix", (wide + price) > E.decimal(0, 30, 4), [wide + price, wide - price, wide * disc, price - disc, price + price]),
        ("bigint_double_arith", None, [big + big, big - E.bigint(7), big * E.bigint(3), dbl * dbl, dbl + E.double(1.5), i32 + i32, i32 * E.integer(2)]),
        ("casts_and_hash", big.ne(E.bigint(0)), [i32.cast(P.BIGINT), big.cast(P.DOUBLE), i32.cast(P.LONG_DECIMAL, 20, 3), price.cast(P.LONG_DECIMAL, 20, 6),
                                               E.hash_of(big, date, seg), E.hash_of(wide, dbl, flag, price)]),
        ("filter_passthrough_fixed", date < E.date(9500), [big, date, price, wide, dbl, i32, flag, big]),    # one compaction kernel, widths 1 / 4 / 8 / 16
        ("filter_passthrough_simple_cmp", i32 >= E.integer(0), [flag, wide]),
        ("reject_all", date > E.date(20000), [big, seg]),
        ("null_const", E.const(None, P.BOOLEAN) | (date >= E.date(0)), [big + E.const(None, P.BIGINT), seg]),
    ]


def pages_equal(got, want, rel=0.0):
    """Host pages: same row count, types and values (nulls compared as None)."""
    assert got.position_count == want.position_count, (got.position_count, want.position_count)
    for k, (g, w) in enumerate(zip(got.blocks, want.blocks)):
        g, w = g.to_host(), w.to_host()
        assert g.type == w.type, (k, g.type, w.type)
        gl, wl = g.to_list(), w.to_list()
        if g.type == P.DOUBLE:
            for x, y in zip(gl, wl):
                assert (x is None) == (y is None) and (x is None or x == y or (x != x and y != y)), (k, x, y)
        else:
            assert gl == wl, (k, [(i, x, y) for i, (x, y) in enumerate(zip(gl, wl)) if x != y][:5])
