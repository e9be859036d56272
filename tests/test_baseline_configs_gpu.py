"""Parity for the shapes of BASELINE.json's five configs (SURVEY.md §8c "concrete synthetic inputs"), at scale factors the
oracle finishes in seconds.  The bench measures config 2 at full size; the others are parity cases:
  cfg1  HashAggregationBenchmark: orders, GROUP BY orderstatus, sum(totalprice), 8192-row pages   (HashAggregationBenchmark.java:56-71)
  cfg3  GROUP BY orderkey, dbgen order and shuffled                                               
  cfg4  lineitem JOIN orders + the JMH grid matchRate x buildRowsRepetition                      (BenchmarkHashBuildAndJoinOperators.java:98-105,187-194)
  cfg5  customer JOIN orders JOIN lineitem, composite GROUP BY, 1 % null keys                     (TPC-H Q3 shape)
  grid  BenchmarkGroupByHash: channels x {BIGINT, VARCHAR} x hash channel on/off                  (BenchmarkGroupByHash.java:71-74,258-354)
Every comparison goes through the C ABI on the GPU side and the oracle on the CPU side, on identical buffers."""
import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu

STARTDATE, ENDDATE = 8035, 10591          # 1992-01-01 .. 1998-12-31 as days since the epoch


def drain(op):
    pages = []
    while True:
        p = op.get_output()
        if p is None:
            if op.is_finished() or op.needs_input():
                break
            continue
        pages.append(p)
    return pages


def rows_of(pages):
    rows = []
    for p in pages:
        cols = [b.to_host().to_list() for b in p.blocks]
        rows += [tuple(c[i] for c in cols) for i in range(p.position_count)]
    return rows


def order_key(i):
    return (i // 8) * 32 + (i % 8) + 1      # dbgen's sparse order keys


def gen_orders(rng, n, ncust):
    status = rng.choice(np.array(["F", "O", "P"]), size=n, p=[0.486, 0.488, 0.026])
    return {
        "orderkey": order_key(np.arange(n, dtype=np.int64)),
        "custkey": rng.integers(1, ncust + 1, n).astype(np.int64),
        "totalprice": rng.integers(85_771, 55_528_517, n).astype(np.float64) / 100.0,
        "orderdate": rng.integers(STARTDATE, ENDDATE - 151, n).astype(np.int32),
        "shippriority": np.zeros(n, dtype=np.int32),
        "orderstatus": status,
    }


def gen_lineitem(rng, n_orders, shuffle=False):
    counts = rng.integers(1, 8, n_orders)
    oidx = np.repeat(np.arange(n_orders, dtype=np.int64), counts)
    if shuffle:
        oidx = oidx[rng.permutation(len(oidx))]
    n = len(oidx)
    qty = rng.integers(1, 51, n).astype(np.int64)
    price = qty * rng.integers(90_000, 200_000, n).astype(np.int64)          # decimal(12,2) unscaled
    disc = rng.integers(0, 11, n).astype(np.int64)                            # decimal(12,2): 0.00 .. 0.10
    return {"orderkey": order_key(oidx), "order_idx": oidx, "quantity": qty * 100, "extendedprice": price, "discount": disc,
            "shipdate": rng.integers(STARTDATE + 1, ENDDATE - 30, n).astype(np.int32)}


def oracle_group_rows(P, key_types, pages, key_channels, aggs, final=True, expected=10_000):
    """Oracle rows in group-id order for raw-input pages (lists of host blocks)."""
    ogb = O.OracleGroupBy(key_types, expected)
    accs = [O.OracleAccumulator(a.kind, a.in_type) for a in aggs]
    for blocks in pages:
        ids = ogb.get_group_ids([blocks[c] for c in key_channels])
        g = ogb.group_count()
        for a, oa in zip(aggs, accs):
            oa.add_input(ids, g, blocks[a.input_channel] if a.input_channel >= 0 else None,
                         blocks[a.mask_channel] if a.mask_channel >= 0 else None)
    g = ogb.group_count()
    keys, _ = ogb.export(with_hash=False)
    cols = [k.to_list() for k in keys]
    for oa in accs:
        rc, out, _ = oa.evaluate(0, g, final)
        assert rc == 0
        cols += [o.to_list() for o in out]
    return [tuple(c[i] for c in cols) for i in range(g)]


def assert_rows(got, want):
    assert len(got) == len(want)
    for a, b in zip(got, want):
        assert len(a) == len(b)
        for x, y in zip(a, b):
            if isinstance(x, float) and isinstance(y, float):
                assert x == pytest.approx(y, rel=1e-12, abs=1e-300)      # north_star's tolerance for double sums
            else:
                assert x == y


# ------------------------------------------------------------------------------------------------ cfg1
@pytest.mark.parametrize("device", [False, True])
def test_cfg1_orders_group_by_orderstatus_sum_totalprice(pgpu, device):
    """HashAggregationBenchmark: Step.SINGLE, expectedGroups 100 000, one 8192-row page at a time."""
    P = pgpu
    rng = np.random.default_rng(42)
    o = gen_orders(rng, 150_000, 15_000)                                      # SF0.1
    aggs = [P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 1)]
    op = P.HashAggregationOperator([P.VARCHAR], [0], P.SINGLE, aggs, expected_groups=100_000)
    pages = []
    for lo in range(0, 150_000, 8192):
        hi = min(lo + 8192, 150_000)
        blocks = [P.Block.variable_width(list(o["orderstatus"][lo:hi])), P.Block.double_array(o["totalprice"][lo:hi])]
        pages.append(blocks)
        page = P.Page(*blocks)
        op.add_input(page.to_device() if device else page)
    op.finish()
    got = rows_of(drain(op))
    want = oracle_group_rows(P, [P.VARCHAR], pages, [0], aggs, expected=100_000)
    assert_rows(got, want)
    assert sorted(r[0] for r in got) == [b"F", b"O", b"P"]
    assert op.get_capacity() == 262144                                        # arraySize(100000, 0.75), observable via getCapacity()


def test_cfg1_single_page_takes_the_fused_path(pgpu):
    """The same query fed as one large page (what a device-resident scan hands over) must give the same rows."""
    P = pgpu
    rng = np.random.default_rng(43)
    o = gen_orders(rng, 400_000, 15_000)
    aggs = [P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 1), P.AggregatorFactory(P.AGG_COUNT)]
    blocks = [P.Block.variable_width(list(o["orderstatus"])), P.Block.double_array(o["totalprice"])]
    op = P.HashAggregationOperator([P.VARCHAR], [0], P.SINGLE, aggs, expected_groups=100_000)
    op.add_input(P.Page(*blocks).to_device())
    op.finish()
    assert_rows(rows_of(drain(op)), oracle_group_rows(P, [P.VARCHAR], [blocks], [0], aggs))


# ------------------------------------------------------------------------------------------------ cfg3
@pytest.mark.parametrize("shuffle", [False, True])
def test_cfg3_group_by_orderkey(pgpu, shuffle):
    """count(*), sum(quantity short decimal), sum(extendedprice double); 1-7 rows per group."""
    P = pgpu
    rng = np.random.default_rng(44)
    li = gen_lineitem(rng, 120_000, shuffle)
    n = len(li["orderkey"])
    price = li["extendedprice"].astype(np.float64) / 100.0
    half = n // 2
    pages = [[P.Block.long_array(li["orderkey"][lo:hi]), P.Block.long_array(li["quantity"][lo:hi], None, P.SHORT_DECIMAL),
              P.Block.double_array(price[lo:hi])] for lo, hi in ((0, half), (half, n))]
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 2)]
    op = P.HashAggregationOperator([P.BIGINT], [0], P.SINGLE, aggs, expected_groups=100_000)
    for pg in pages:
        op.add_input(P.Page(*pg).to_device())
    op.finish()
    got = rows_of(drain(op))
    want = oracle_group_rows(P, [P.BIGINT], pages, [0], aggs, expected=100_000)
    assert_rows(got, want)
    assert len(got) == 120_000 and sum(r[1] for r in got) == n               # checksum of counts


# ------------------------------------------------------------------------------------------------ cfg4
def _join(P, build_blocks, types, join_channels, probe_keys, join_type=0, device=True):
    idx = P.PagesIndex(types, join_channels, expected_positions=build_blocks[0].n)
    orc = O.OracleJoin(types, join_channels)
    idx.add_page([b.to_device() for b in build_blocks] if device else build_blocks)
    orc.add_page(build_blocks)
    ls = idx.create_lookup_source()
    orc.build()
    pk = [b.to_device() for b in probe_keys] if device else probe_keys
    gp, gb = ls.join_page(pk, None, join_type)
    wp, wb = orc.probe(probe_keys, None, join_type)
    if device:
        gp, gb = gp.cpu().numpy(), gb.cpu().numpy()
    return ls, gp, gb, wp, wb


@pytest.mark.parametrize("shuffle", [False, True])
def test_cfg4_lineitem_join_orders(pgpu, shuffle):
    """INNER join on orderkey: unique build keys, every probe row matches exactly once; output = probe idx + totalprice."""
    P = pgpu
    rng = np.random.default_rng(45)
    n_orders = 100_000
    o = gen_orders(rng, n_orders, 10_000)
    li = gen_lineitem(rng, n_orders, shuffle)
    build = [P.Block.long_array(o["orderkey"]), P.Block.double_array(o["totalprice"])]
    ls, gp, gb, wp, wb = _join(P, build, [P.BIGINT, P.DOUBLE], [0], [P.Block.long_array(li["orderkey"])])
    assert np.array_equal(gp, wp) and np.array_equal(gb, wb)
    assert len(gp) == len(li["orderkey"]) and np.array_equal(gp, np.arange(len(gp), dtype=np.int32))
    out = ls.append_to(gb, [1])[0].to_host()
    assert np.array_equal(out.values.view(np.float64), o["totalprice"][li["order_idx"]])     # appendTo == a numpy gather


@pytest.mark.parametrize("match_rate", [0.1, 1.0, 2.0])
@pytest.mark.parametrize("repetition", [1, 5])
def test_cfg4_jmh_grid(pgpu, match_rate, repetition):
    """matchRate: fraction of probe rows with a build match (2.0: every probe key appears twice);
    buildRowsRepetition: how many build rows share each key (position links, newest first)."""
    P = pgpu
    rng = np.random.default_rng(46)
    nbuild, nprobe = 50_000, 200_000
    distinct = nbuild // repetition
    bkeys = np.tile(np.arange(distinct, dtype=np.int64), repetition)
    build = [P.Block.long_array(bkeys), P.Block.long_array(np.arange(nbuild, dtype=np.int64))]
    if match_rate <= 1.0:
        hit = rng.random(nprobe) < match_rate
        pkeys = np.where(hit, rng.integers(0, distinct, nprobe), rng.integers(distinct, 2 * distinct + 1, nprobe)).astype(np.int64)
    else:
        pkeys = np.repeat(rng.integers(0, distinct, nprobe // 2), 2).astype(np.int64)
    for join_type in (P.JOIN_INNER, P.JOIN_PROBE_OUTER):
        _, gp, gb, wp, wb = _join(P, build, [P.BIGINT, P.BIGINT], [0], [P.Block.long_array(pkeys)], join_type)
        assert np.array_equal(gp, wp) and np.array_equal(gb, wb)
        if match_rate >= 1.0:
            assert len(gp) == nprobe * repetition


# ------------------------------------------------------------------------------------------------ cfg5
def test_cfg5_q3_shape_two_joins_then_composite_group_by(pgpu):
    """customer(mktsegment = 'BUILDING') JOIN orders(orderdate < 1995-03-15) ON custkey JOIN lineitem(shipdate > 1995-03-15)
    ON orderkey, GROUP BY (l_orderkey, o_orderdate, o_shippriority), sum(extendedprice * (1 - discount)) as decimal(24,4);
    1 % NULL o_custkey and l_orderkey.  Each operator's output is compared with the oracle's before it feeds the next."""
    P = pgpu
    rng = np.random.default_rng(47)
    ncust, n_orders = 15_000, 150_000
    cutoff = 9204                                                              # 1995-03-15
    segment = rng.choice(np.array(["AUTOMOBILE", "BUILDING", "FURNITURE", "MACHINERY", "HOUSEHOLD"]), ncust)
    cust_keys = np.arange(1, ncust + 1, dtype=np.int64)[segment == "BUILDING"]
    o = gen_orders(rng, n_orders, ncust)
    osel = o["orderdate"] < cutoff
    o_custnull = (np.random.default_rng(43).random(n_orders) < 0.01).astype(np.uint8)
    li = gen_lineitem(rng, n_orders)
    lsel = li["shipdate"] > cutoff
    l_keynull = (np.random.default_rng(43).random(len(lsel)) < 0.01).astype(np.uint8)

    # join 1: build = filtered customer, probe = filtered orders on custkey (null custkeys never match)
    ocust = P.Block.long_array(o["custkey"][osel], o_custnull[osel])
    _, gp, gb, wp, wb = _join(P, [P.Block.long_array(cust_keys)], [P.BIGINT], [0], [ocust])
    assert np.array_equal(gp, wp) and np.array_equal(gb, wb)
    ord_rows = np.flatnonzero(osel)[gp]                                        # orders that survive, in probe order
    # join 2: build = surviving orders (orderkey, orderdate, shippriority), probe = filtered lineitem on orderkey
    build2 = [P.Block.long_array(o["orderkey"][ord_rows]), P.Block.int_array(o["orderdate"][ord_rows], None, P.DATE),
              P.Block.int_array(o["shippriority"][ord_rows])]
    lkey = P.Block.long_array(li["orderkey"][lsel], l_keynull[lsel])
    ls2, gp2, gb2, wp2, wb2 = _join(P, build2, [P.BIGINT, P.DATE, P.INTEGER], [0], [lkey])
    assert np.array_equal(gp2, wp2) and np.array_equal(gb2, wb2)
    assert 0 < len(gp2) < int(lsel.sum())
    gathered = ls2.append_to(gb2, [1, 2])
    lrows = np.flatnonzero(lsel)[gp2]
    assert np.array_equal(gathered[0].to_host().values, o["orderdate"][ord_rows][gb2])
    # aggregation: revenue = extendedprice * (100 - discount) as a long decimal (scale 4)
    revenue = [int(p) * (100 - int(d)) for p, d in zip(li["extendedprice"][lrows], li["discount"][lrows])]
    blocks = [P.Block.long_array(li["orderkey"][lrows]), gathered[0].to_host(), gathered[1].to_host(), P.Block.int128_array(revenue)]
    aggs = [P.AggregatorFactory(P.AGG_SUM, P.LONG_DECIMAL, 3)]
    op = P.HashAggregationOperator([P.BIGINT, P.DATE, P.INTEGER], [0, 1, 2], P.SINGLE, aggs, expected_groups=10_000)
    op.add_input(P.Page(*blocks).to_device())
    op.finish()
    got = rows_of(drain(op))
    want = oracle_group_rows(P, [P.BIGINT, P.DATE, P.INTEGER], [blocks], [0, 1, 2], aggs)
    assert_rows(got, want)
    assert sum(r[3] for r in got) == sum(revenue)                              # linearity: group sums add up to the column sum


def test_cfg5_null_group_and_probe_outer(pgpu):
    """The null-injection variant on the aggregation side: NULL l_orderkey rows form the null group of a composite key,
    and a probe-outer join keeps them with NULL build columns."""
    P = pgpu
    rng = np.random.default_rng(48)
    n = 300_000
    okey = rng.integers(1, 40_000, n).astype(np.int64)
    knull = (np.random.default_rng(43).random(n) < 0.01).astype(np.uint8)
    date = (STARTDATE + okey % 2000).astype(np.int32)
    prio = np.zeros(n, dtype=np.int32)
    val = rng.integers(1, 10**9, n).astype(np.int64)
    blocks = [P.Block.long_array(okey, knull), P.Block.int_array(date, knull, P.DATE), P.Block.int_array(prio), P.Block.long_array(val, None, P.SHORT_DECIMAL)]
    aggs = [P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 3), P.AggregatorFactory(P.AGG_COUNT)]
    op = P.HashAggregationOperator([P.BIGINT, P.DATE, P.INTEGER], [0, 1, 2], P.SINGLE, aggs, expected_groups=10_000)
    op.add_input(P.Page(*blocks))
    op.finish()
    got = rows_of(drain(op))
    assert_rows(got, oracle_group_rows(P, [P.BIGINT, P.DATE, P.INTEGER], [blocks], [0, 1, 2], aggs))
    null_rows = [r for r in got if r[0] is None]
    assert len(null_rows) == 1 and null_rows[0][4] == int(knull.sum())
    build = [P.Block.long_array(np.arange(1, 20_000, dtype=np.int64))]
    _, gp, gb, wp, wb = _join(P, build, [P.BIGINT], [0], [blocks[0]], P.JOIN_PROBE_OUTER)
    assert np.array_equal(gp, wp) and np.array_equal(gb, wb)
    assert len(gp) == n and int((gb < 0).sum()) >= int(knull.sum())


# ------------------------------------------------------------------------------------------------ BenchmarkGroupByHash grid
@pytest.mark.parametrize("channels", [1, 5, 10])
@pytest.mark.parametrize("data_type", ["BIGINT", "VARCHAR"])
@pytest.mark.parametrize("hash_enabled", [False, True])
def test_groupby_hash_benchmark_grid(pgpu, channels, data_type, hash_enabled):
    """uniform nextInt(groupCount) keys replicated over `channels` columns, 4-byte strings for VARCHAR; group ids,
    group count and exported keys must equal the oracle's, page by page."""
    P = pgpu
    rng = np.random.default_rng(49 + channels)
    n, groups, npages = 60_000, 30_000, 3
    types = [P.BIGINT if data_type == "BIGINT" else P.VARCHAR] * channels
    gb = P.GroupByHash(types, expected_size=groups, has_hash_channel=hash_enabled)
    ogb = O.OracleGroupBy(types, groups, hash_enabled)
    for pg in range(npages):
        v = rng.integers(0, groups, n)
        if data_type == "BIGINT":
            blocks = [P.Block.long_array(v.astype(np.int64)) for _ in range(channels)]
        else:
            strs = [int(x).to_bytes(4, "little") for x in v]
            blocks = [P.Block.variable_width(strs) for _ in range(channels)]
        if hash_enabled:
            blocks = blocks + [P.Block.long_array(O.hash_rows(blocks))]
        ids = gb.get_group_ids([b.to_device() for b in blocks])
        want = ogb.get_group_ids(blocks)
        assert np.array_equal(ids.ids.cpu().numpy(), want)
        assert ids.get_group_count() == ogb.group_count() == gb.get_group_count()
    assert gb.get_capacity() == ogb.capacity()
