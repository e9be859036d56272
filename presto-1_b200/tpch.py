"""TPC-H-shaped synthetic column generator, written straight into device memory with torch (plumbing only).

It follows dbgen's value domains for the columns the configs touch: sparse order keys, 1-7 lineitems per
order, quantity 1-50, retail-price formula, discount 0.00-0.10, tax 0.00-0.08, ship/receipt date offsets, returnflag /
linestatus derived from the dates.  It is NOT bit-identical to io.trino.tpch:tpch (an un-vendored dependency of the
reference); the GPU path and the oracle always consume the same generated buffers, so parity does not depend on that.
What it is held to instead: the column statistics the reference ships for the TPC-H connector (row counts, min / max, distinct
counts per column at SF 0.01 and SF 1; tests/golden/tpch_stats.json, checked by tests/test_tpch_stats_cpu.py).
Money columns are unscaled integer cents (short decimal); derived products are long decimals (Int128, sign-magnitude).
"""
import numpy as np
import torch

from . import _native as N
from .blocks import Block, Page

STARTDATE = 8035          # 1992-01-01 as days since 1970-01-01
ENDDATE_ORDER = 10440     # 1998-08-02, last order date
CURRENTDATE = 9298        # 1995-06-17
Q1_CUTOFF = 10471         # 1998-09-02 (date '1998-12-01' - interval '90' day)
LINES_PER_ORDER_MEAN = 4


def order_key(i):
    """dbgen's sparse order keys (mk_sparse): the 1-based order number keeps its low 3 bits and the rest moves up by two bits, so 8 of
    every 32 values are used and order number 1 500 000 x SF gets key 6 000 000 x SF (the o_orderkey max of the reference's statistics
    files).  `i` is the 0-based order index."""
    return (((i + 1) >> 3) << 5) | ((i + 1) & 7)


_M64 = (1 << 64) - 1


def _i64(x):
    """Python int -> the int64 with the same low 64 bits."""
    x &= _M64
    return x - (1 << 64) if x >= (1 << 63) else x


_GOLD, _MIX1, _MIX2 = _i64(0x9E3779B97F4A7C15), _i64(0xBF58476D1CE4E5B9), _i64(0x94D049BB133111EB)


def _mix(seed, stream, idx):
    """Counter-based generator: splitmix64 finaliser of (seed, stream, row index), written with wrapping int64 tensor ops so that
    the CPU and the CUDA build of torch produce the SAME values (torch.Generator streams differ between devices).  Both bench
    arms — the GPU operators and the CPU reference arm — therefore aggregate identical rows."""
    z = idx * _GOLD + _i64(seed * 0xD1342543DE82EF95 + stream * 0x2545F4914F6CDD1D + 0x1234567)
    z = (z ^ ((z >> 30) & ((1 << 34) - 1))) * _MIX1
    z = (z ^ ((z >> 27) & ((1 << 37) - 1))) * _MIX2
    return z ^ ((z >> 31) & ((1 << 33) - 1))


def _randint(seed, stream, idx, lo, hi):
    """Uniform integers in [lo, hi) (int64), one per entry of idx."""
    return lo + ((_mix(seed, stream, idx) >> 11) & ((1 << 53) - 1)) % (hi - lo)


def orders(sf, seed=42, device="cuda", rows=None):
    """orders: o_orderkey (unique, sparse), o_custkey, o_totalprice (double), o_orderdate, o_shippriority, o_orderstatus."""
    n = int(1_500_000 * sf) if rows is None else int(rows)
    idx = torch.arange(n, device=device, dtype=torch.int64)
    okey = order_key(idx)
    custkey = _randint(seed, 1, idx, 1, max(2, int(150_000 * sf)) + 1)
    custkey = torch.where(custkey % 3 == 0, custkey - 1, custkey)             # dbgen: a third of the customers never order (custkey % 3 != 0)
    totalprice = _randint(seed, 2, idx, 85_771, 55_528_517).double() * 0.01   # a product, not a quotient: torch CUDA turns x / c into x * (1 / c), the CPU build does not
    orderdate = _randint(seed, 3, idx, STARTDATE, ENDDATE_ORDER + 1).to(torch.int32)
    shipprio = torch.zeros(n, device=device, dtype=torch.int32)
    return {"orderkey": okey, "custkey": custkey, "totalprice": totalprice, "orderdate": orderdate, "shippriority": shipprio}


SEGMENTS = ("AUTOMOBILE", "BUILDING", "FURNITURE", "MACHINERY", "HOUSEHOLD")


def customer(sf, seed=42, device="cuda"):
    """customer: c_custkey (1..150 000 x SF), c_mktsegment (one of the five TPC-H segments, a VARCHAR column: bytes + offsets)."""
    n = max(2, int(150_000 * sf))
    idx = torch.arange(n, device=device, dtype=torch.int64)
    seg = _randint(seed + 2, 1, idx, 0, len(SEGMENTS))
    lens = torch.tensor([len(x) for x in SEGMENTS], device=device, dtype=torch.int64)
    text = torch.tensor(list("".join(SEGMENTS).encode()), device=device, dtype=torch.uint8)
    starts = torch.cumsum(lens, 0) - lens
    offsets = torch.zeros(n + 1, device=device, dtype=torch.int64)
    torch.cumsum(lens[seg], 0, out=offsets[1:])
    row_of_byte = torch.repeat_interleave(idx, lens[seg])
    within = torch.arange(int(offsets[-1].item()), device=device, dtype=torch.int64) - offsets[:-1][row_of_byte]
    data = text[starts[seg][row_of_byte] + within]
    return {"rows": n, "custkey": idx + 1, "segment_code": seg, "mktsegment_bytes": data, "mktsegment_offsets": offsets.to(torch.int32)}


def null_mask(n, seed, stream, one_in, device):
    """uint8 mask with about one row in `one_in` set (the same rows on CPU and CUDA)."""
    idx = torch.arange(n, device=device, dtype=torch.int64)
    return (_randint(seed, stream, idx, 0, one_in) == 0).to(torch.uint8)


def shuffle_order(n, seed, device):
    """A seeded permutation of range(n), identical on CPU and CUDA (stable argsort of hashed row numbers)."""
    return torch.argsort(_mix(seed, 99, torch.arange(n, device=device, dtype=torch.int64)), stable=True)


def lineitem(sf, seed=42, device="cuda", orders_count=None, need=("q1",), shuffle=False):
    """lineitem columns.  need: subset of {"q1", "keys", "double", "raw", "stats"} selecting which columns are materialised.
    "q1" adds the derived long-decimal columns and the filter mask (the materialised Q1 page); "raw" only the raw columns
    (shipdate, discount, tax, returnflag, linestatus) a fused filter + projection reads."""
    n_orders = int(1_500_000 * sf) if orders_count is None else int(orders_count)
    oidx = torch.arange(n_orders, device=device, dtype=torch.int64)
    counts = _randint(seed + 1, 1, oidx, 1, 8)
    order_idx = torch.repeat_interleave(oidx, counts)
    del counts
    n = int(order_idx.numel())
    out = {"rows": n}
    idx = torch.arange(n, device=device, dtype=torch.int64)     # the row's identity: survives the shuffle
    if shuffle:
        perm = shuffle_order(n, seed + 1, device)
        order_idx, idx = order_idx[perm], idx[perm]
        del perm
    if "keys" in need or "join" in need:
        out["orderkey"] = order_key(order_idx)
    if "stats" in need:                                          # l_linenumber: 1.. within the order (needs the unshuffled row identity)
        first = torch.cumsum(_randint(seed + 1, 1, oidx, 1, 8), 0)
        first = torch.cat([first.new_zeros(1), first[:-1]])
        out["linenumber"] = (idx - first[order_idx] + 1).to(torch.int32)
        del first
    s = seed + 1
    qty = _randint(s, 2, idx, 1, 51)
    partkey = _randint(s, 3, idx, 1, max(2, int(200_000 * sf)) + 1)
    retail = 90_000 + (torch.div(partkey, 10, rounding_mode="floor") % 20_001) + 100 * (partkey % 1000)
    if "stats" in need:
        out["partkey"] = partkey
    out["quantity"] = qty * 100                                  # decimal(12,2)
    out["extendedprice"] = qty * retail                          # decimal(12,2) cents
    del partkey, retail, qty
    if "q1" in need or "raw" in need:
        discount = _randint(s, 4, idx, 0, 11)
        tax = _randint(s, 5, idx, 0, 9)
        orderdate = _randint(s, 6, oidx, STARTDATE, ENDDATE_ORDER + 1)[order_idx]
        shipdate = (orderdate + _randint(s, 7, idx, 1, 122)).to(torch.int32)
        receiptdate = shipdate + _randint(s, 8, idx, 1, 31).to(torch.int32)
        del orderdate
        ra = torch.where((_mix(s, 9, idx) & 1) == 0, ord("R"), ord("A")).to(torch.uint8)
        out["returnflag"] = torch.where(receiptdate <= CURRENTDATE, ra, torch.full_like(ra, ord("N")))
        out["linestatus"] = torch.where(shipdate > CURRENTDATE, ord("O"), ord("F")).to(torch.uint8)
        del ra, receiptdate
        out["shipdate"] = shipdate
        out["discount"] = discount
        out["tax"] = tax
        out["offsets1"] = torch.arange(n + 1, device=device, dtype=torch.int32)   # varchar(1): one byte per row
        if "q1" in need:
            disc_price = out["extendedprice"] * (100 - discount)     # decimal(24,4)
            charge = disc_price * (100 + tax)                        # decimal(36,6); < 2^63 for TPC-H domains
            out["disc_price"] = torch.stack([disc_price, torch.zeros_like(disc_price)], dim=1).reshape(-1)
            out["charge"] = torch.stack([charge, torch.zeros_like(charge)], dim=1).reshape(-1)
            out["q1_mask"] = (shipdate <= Q1_CUTOFF).to(torch.uint8)
    if "double" in need:
        out["extendedprice_double"] = out["extendedprice"].double() * 0.01
    return out


def q1_page(cols):
    """Page for the Q1 configuration: [returnflag, linestatus, quantity, extendedprice, discount, disc_price, charge, mask]."""
    n = cols["rows"]
    off = cols["offsets1"]
    return Page(
        Block(N.VARCHAR, cols["returnflag"], None, off, n),
        Block(N.VARCHAR, cols["linestatus"], None, off, n),
        Block(N.SHORT_DECIMAL, cols["quantity"], None, None, n, 12, 2),
        Block(N.SHORT_DECIMAL, cols["extendedprice"], None, None, n, 12, 2),
        Block(N.SHORT_DECIMAL, cols["discount"], None, None, n, 12, 2),
        Block(N.LONG_DECIMAL, cols["disc_price"], None, None, n, 24, 4),
        Block(N.LONG_DECIMAL, cols["charge"], None, None, n, 36, 6),
        Block(N.BOOLEAN, cols["q1_mask"], None, None, n),
        position_count=n)


Q1_BYTES_PER_ROW = 2 * (1 + 4) + 8 + 8 + 8 + 16 + 16 + 1   # = 67, SURVEY.md §8d "materialised inputs"


def q1_raw_page(cols):
    """The page Q1's table scan delivers: [returnflag, linestatus, quantity, extendedprice, discount, tax, shipdate] — raw lineitem
    columns only; the filter and the derived decimals are expressions (q1_processor)."""
    n = cols["rows"]
    off = cols["offsets1"]
    return Page(
        Block(N.VARCHAR, cols["returnflag"], None, off, n),
        Block(N.VARCHAR, cols["linestatus"], None, off, n),
        Block(N.SHORT_DECIMAL, cols["quantity"], None, None, n, 12, 2),
        Block(N.SHORT_DECIMAL, cols["extendedprice"], None, None, n, 12, 2),
        Block(N.SHORT_DECIMAL, cols["discount"], None, None, n, 12, 2),
        Block(N.SHORT_DECIMAL, cols["tax"], None, None, n, 12, 2),
        Block(N.DATE, cols["shipdate"], None, None, n),
        position_count=n)


# what the fused kernel reads per row: two varchar(1) keys (byte + offset), four 8-byte decimals, the 4-byte ship date
Q1_RAW_BYTES_PER_ROW = 2 * (1 + 4) + 4 * 8 + 4              # = 46 (SURVEY.md §8d "fused derive": 43 with a 1-byte mask instead of the date)


def q1_expressions():
    """(filter, projections) of TPC-H Q1 over q1_raw_page: shipdate <= DATE '1998-09-02';
    [returnflag, linestatus, quantity, extendedprice, extendedprice*(1-discount), extendedprice*(1-discount)*(1+tax), discount]."""
    from .expressions import Expr as E
    qty, price = E.input(2, N.SHORT_DECIMAL, 12, 2), E.input(3, N.SHORT_DECIMAL, 12, 2)
    disc, tax = E.input(4, N.SHORT_DECIMAL, 12, 2), E.input(5, N.SHORT_DECIMAL, 12, 2)
    one = E.decimal(1, 1, 0)
    disc_price = price * (one - disc)            # decimal(25,4)
    charge = disc_price * (one + tax)            # decimal(38,6)
    return E.input(6, N.DATE) <= E.date(Q1_CUTOFF), [E.input(0, N.VARCHAR), E.input(1, N.VARCHAR), qty, price, disc_price, charge, disc]


def q1_fused_aggregators():
    """sum(qty), sum(price), sum(disc_price), sum(charge), avg(qty), avg(price), avg(disc), count(*) over q1_expressions' projections."""
    from .operators import AggregatorFactory as A
    return [A(N.AGG_SUM, N.SHORT_DECIMAL, 2), A(N.AGG_SUM, N.SHORT_DECIMAL, 3), A(N.AGG_SUM, N.LONG_DECIMAL, 4), A(N.AGG_SUM, N.LONG_DECIMAL, 5),
            A(N.AGG_AVG, N.SHORT_DECIMAL, 2), A(N.AGG_AVG, N.SHORT_DECIMAL, 3), A(N.AGG_AVG, N.SHORT_DECIMAL, 6), A(N.AGG_COUNT)]


def q1_aggregators():
    """sum(qty), sum(price), sum(disc_price), sum(charge), avg(qty), avg(price), avg(disc), count(*) — all masked by the filter."""
    from .operators import AggregatorFactory as A
    m = 7
    return [A(N.AGG_SUM, N.SHORT_DECIMAL, 2, m), A(N.AGG_SUM, N.SHORT_DECIMAL, 3, m), A(N.AGG_SUM, N.LONG_DECIMAL, 5, m),
            A(N.AGG_SUM, N.LONG_DECIMAL, 6, m), A(N.AGG_AVG, N.SHORT_DECIMAL, 2, m), A(N.AGG_AVG, N.SHORT_DECIMAL, 3, m),
            A(N.AGG_AVG, N.SHORT_DECIMAL, 4, m), A(N.AGG_COUNT, 0, -1, m)]


def page_to_host(page, pinned=True):
    """Device page -> host page backed by (pinned) numpy arrays: what a JNI caller would hand to the C ABI."""
    def h(t):
        if t is None:
            return None
        c = t.cpu()
        if pinned:
            c = c.pin_memory()
        return c.numpy()
    out = Page(*[Block(b.type, h(b.values), h(b.nulls), h(b.offsets), b.n, b.precision, b.scale) for b in page.blocks],
               position_count=page.position_count)
    for b in out.blocks:
        b.pinned = bool(pinned)      # the numpy arrays view page-locked torch storage that lives as long as they do
    return out
