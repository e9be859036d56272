"""GPU parity for the fused GROUP BY operator (pg_hashagg_*) through the HashAggregationOperator mirror:
OPT/TestHashAggregationOperator.java testHashAggregation :162-220 (40 000 groups x 3 pages, expected rows
(i, 3, 3i, i, ..., 3, 3), order-insensitive), partial -> final, testMultiplePartialFlushes :513-593, and a Q1-shaped
decimal aggregation checked against the oracle."""
import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


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


def seq_page(P, n, starts):
    """RowPagesBuilder.addSequencePage(n, starts...) for (VARCHAR, VARCHAR, VARCHAR, BIGINT, BOOLEAN)."""
    return P.Page(
        P.Block.variable_width([str(starts[0] + i) for i in range(n)]),
        P.Block.variable_width([str(starts[1] + i) for i in range(n)]),
        P.Block.variable_width([str(starts[2] + i) for i in range(n)]),
        P.Block.long_array(np.arange(starts[3], starts[3] + n, dtype=np.int64)),
        P.Block.byte_array(((np.arange(n) + starts[4]) % 2).astype(np.uint8)),
    )


@pytest.mark.parametrize("hash_enabled", [False, True])
@pytest.mark.parametrize("device", [False, True])
def test_hash_aggregation_reference_rows(pgpu, hash_enabled, device):
    P = pgpu
    n = 40_000
    pages = [seq_page(P, n, (100, 0, s, 0, 500)) for s in (100_000, 200_000, 300_000)]
    hash_channel = None
    if hash_enabled:
        pages = [P.Page(*(p.blocks + [P.Block.long_array(O.hash_rows([p.blocks[1]]))])) for p in pages]
        hash_channel = 5
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.BIGINT, 3), P.AggregatorFactory(P.AGG_AVG, P.BIGINT, 3),
            P.AggregatorFactory(P.AGG_MAX, P.BIGINT, 3), P.AggregatorFactory(P.AGG_COUNT_COL, P.VARCHAR, 0),
            P.AggregatorFactory(P.AGG_COUNT_COL, P.BOOLEAN, 4)]
    op = P.HashAggregationOperator([P.VARCHAR], [1], P.SINGLE, aggs, hash_channel, expected_groups=100_000)
    for p in pages:
        assert op.needs_input()
        op.add_input(p.to_device() if device else p)
    op.finish()
    out = drain(op)
    assert len(out) > 1, "Expected more than one output page"
    rows = rows_of(out)
    if hash_enabled:
        rows = [r[:1] + r[2:] for r in rows]   # drop the hash column like assertPagesEqualIgnoreOrder does
    expected = [(str(i).encode(), 3, 3 * i, float(i), i, 3, 3) for i in range(n)]
    assert sorted(rows) == sorted(expected)
    assert op.is_finished()


def test_partial_then_final(pgpu):
    P = pgpu
    rng = np.random.default_rng(3)
    n, groups = 200_000, 5000
    k1 = rng.integers(0, groups, n).astype(np.int64)
    k2 = rng.integers(0, 3, n).astype(np.int32)
    v_dec = rng.integers(-10**9, 10**9, n).astype(np.int64)
    v_dbl = rng.random(n) * 1000
    nulls = (rng.random(n) < 0.1).astype(np.uint8)
    def page(lo, hi):
        return P.Page(P.Block.long_array(k1[lo:hi]), P.Block.int_array(k2[lo:hi]), P.Block.long_array(v_dec[lo:hi], nulls[lo:hi], P.SHORT_DECIMAL),
                      P.Block.double_array(v_dbl[lo:hi]))
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 2), P.AggregatorFactory(P.AGG_AVG, P.SHORT_DECIMAL, 2),
            P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 3), P.AggregatorFactory(P.AGG_AVG, P.DOUBLE, 3), P.AggregatorFactory(P.AGG_MIN, P.DOUBLE, 3),
            P.AggregatorFactory(P.AGG_MAX, P.SHORT_DECIMAL, 2), P.AggregatorFactory(P.AGG_COUNT_COL, P.SHORT_DECIMAL, 2)]
    types = [P.BIGINT, P.INTEGER]
    single = P.HashAggregationOperator(types, [0, 1], P.SINGLE, aggs)
    partials = [P.HashAggregationOperator(types, [0, 1], P.PARTIAL, aggs) for _ in range(2)]
    for i, (lo, hi) in enumerate([(0, 70_000), (70_000, 140_000), (140_000, n)]):
        single.add_input(page(lo, hi))
        partials[i % 2].add_input(page(lo, hi).to_device())
    single.finish()
    # final: intermediate columns start at channel 2; channels per aggregate follow intermediate_types()
    fin_aggs, ch = [], 2
    for a in aggs:
        fin_aggs.append(P.AggregatorFactory(a.kind, a.in_type, ch))
        ch += len(P.intermediate_types(a.kind, a.in_type))
    final = P.HashAggregationOperator(types, [0, 1], P.FINAL, fin_aggs)
    for p in partials:
        p.finish()
        for out in drain(p):
            final.add_input(out)
    final.finish()
    got_single = sorted(rows_of(drain(single)))
    got_final = sorted(rows_of(drain(final)))
    assert len(got_single) == len(got_final)
    for a, b in zip(got_single, got_final):
        for x, y in zip(a, b):
            if isinstance(x, float):
                assert x == pytest.approx(y, rel=1e-12)
            else:
                assert x == y
    # and against numpy for the exact columns
    import collections
    cnt = collections.Counter(zip(k1.tolist(), k2.tolist()))
    sums = collections.defaultdict(int)
    for a, b, v, isn in zip(k1.tolist(), k2.tolist(), v_dec.tolist(), nulls.tolist()):
        if not isn:
            sums[(a, b)] += v
    for r in got_single:
        assert r[2] == cnt[(r[0], r[1])]
        assert r[3] == (sums[(r[0], r[1])] if (r[0], r[1]) in sums else None)


def test_multiple_partial_flushes(pgpu):
    """testMultiplePartialFlushes :513-593: a partial aggregation over its memory limit flushes and keeps accepting input."""
    P = pgpu
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.BIGINT, 0)]
    op = P.HashAggregationOperator([P.BIGINT], [0], P.PARTIAL, aggs, max_partial_memory=1)   # always "full" after a page
    total = []
    for start in (0, 1000, 500):
        assert op.needs_input()
        op.add_input(P.Page(P.Block.long_array(np.arange(start, start + 1000, dtype=np.int64))))
        assert not op.needs_input()            # full -> must be drained first
        total += rows_of(drain(op))
        assert op.needs_input()
    op.finish()
    total += rows_of(drain(op))
    assert op.is_finished()
    # intermediate rows: (key, count, [count, sum]) -> key, 1, 1, key ; keys of overlapping pages appear twice
    assert len(total) == 3000
    assert sorted(r[0] for r in total) == sorted(list(range(0, 2000)) + list(range(500, 1500)))
    assert all(r[1] == 1 and r[2] == 1 and r[3] == r[0] for r in total)


def test_q1_shaped_decimal_aggregation_matches_oracle(pgpu):
    """cfg2 shape at small scale: GROUP BY returnflag, linestatus (2 x varchar(1)); sum/avg over short and long decimals."""
    P = pgpu
    rng = np.random.default_rng(1)
    n = 300_000
    rf = rng.choice([b"A", b"N", b"R"], n)
    ls = rng.choice([b"F", b"O"], n)
    qty = rng.integers(1, 51, n).astype(np.int64) * 100
    price = rng.integers(90_000, 10_500_000, n).astype(np.int64)
    disc = rng.integers(0, 11, n).astype(np.int64)
    tax = rng.integers(0, 9, n).astype(np.int64)
    disc_price = [int(p) * (100 - int(d)) for p, d in zip(price, disc)]
    charge = [dp * (100 + int(t)) for dp, t in zip(disc_price, tax)]
    blocks = [P.Block.variable_width(list(rf)), P.Block.variable_width(list(ls)), P.Block.long_array(qty, None, P.SHORT_DECIMAL),
              P.Block.long_array(price, None, P.SHORT_DECIMAL), P.Block.long_array(disc, None, P.SHORT_DECIMAL),
              P.Block.int128_array(disc_price), P.Block.int128_array(charge)]
    aggs = [P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 2), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 3),
            P.AggregatorFactory(P.AGG_SUM, P.LONG_DECIMAL, 5), P.AggregatorFactory(P.AGG_SUM, P.LONG_DECIMAL, 6),
            P.AggregatorFactory(P.AGG_AVG, P.SHORT_DECIMAL, 2), P.AggregatorFactory(P.AGG_AVG, P.SHORT_DECIMAL, 3),
            P.AggregatorFactory(P.AGG_AVG, P.SHORT_DECIMAL, 4), P.AggregatorFactory(P.AGG_COUNT)]
    op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.SINGLE, aggs)
    op.add_input(P.Page(*blocks).to_device())
    op.finish()
    got = sorted(rows_of(drain(op)))
    # oracle: group ids then accumulators
    ogb = O.OracleGroupBy([P.VARCHAR, P.VARCHAR], 10_000)
    ids = ogb.get_group_ids(blocks[:2])
    g = ogb.group_count()
    keys, _ = ogb.export(with_hash=False)
    cols = [keys[0].to_list(), keys[1].to_list()]
    for a in aggs:
        oa = O.OracleAccumulator(a.kind, a.in_type)
        oa.add_input(ids, g, blocks[a.input_channel] if a.input_channel >= 0 else None)
        rc, out, _ = oa.evaluate(0, g, True)
        assert rc == 0
        cols.append(out[0].to_list())
    want = sorted(tuple(c[i] for c in cols) for i in range(g))
    assert got == want
    assert len(got) == 6
    # and against exact python integers
    for r in got:
        sel = [i for i in range(n) if rf[i] == r[0] and ls[i] == r[1]]
        assert r[2] == sum(int(qty[i]) for i in sel)
        assert r[4] == sum(disc_price[i] for i in sel)
        assert r[9] == len(sel)


@pytest.mark.parametrize("key_kind", ["varchar", "bigint", "int_date"])
def test_fused_low_cardinality_page_with_late_and_unpackable_keys(pgpu, key_kind):
    """Pages >= 1M rows take the fused kernel (packed-key look-up + private partial states).  Groups that first appear
    after the prefix, null keys and strings too long for the packed form must come out exactly like the general path."""
    P = pgpu
    rng = np.random.default_rng(9)
    n = 1_300_000
    late = n - 200_000
    if key_kind == "varchar":
        base = np.array(["A", "N", "R", "", "xy"], dtype=object)
        ks = base[rng.integers(0, 5, n)]
        ks[late:late + 50_000] = "late"                 # 4 bytes: not packable and new after the prefix
        ks[late + 60_000:late + 70_000] = "Q"           # packable but unknown to the prefix
        ks[rng.random(n) < 0.001] = None
        key_blocks = [P.Block.variable_width(list(ks)), P.Block.variable_width(list(base[rng.integers(0, 2, n)]))]
        types = [P.VARCHAR, P.VARCHAR]
    elif key_kind == "bigint":
        k = rng.integers(0, 7, n).astype(np.int64)
        k[late:late + 1000] = 12345
        nulls = (rng.random(n) < 0.001).astype(np.uint8)
        key_blocks = [P.Block.long_array(k, nulls)]
        types = [P.BIGINT]
    else:
        a = rng.integers(0, 3, n).astype(np.int32)
        d = rng.integers(9000, 9003, n).astype(np.int32)
        d[late:] += 7
        key_blocks = [P.Block.int_array(a), P.Block.int_array(d, None, P.DATE)]
        types = [P.INTEGER, P.DATE]
    nk = len(key_blocks)
    v_short = rng.integers(-10**10, 10**10, n).astype(np.int64)
    v_nulls = (rng.random(n) < 0.05).astype(np.uint8)
    v_long = P.Block.int128_array([int(x) * 10**15 for x in rng.integers(-10**6, 10**6, 50_000)] * (n // 50_000))
    v_dbl = rng.random(n)
    mask = (rng.random(n) < 0.9).astype(np.uint8)
    blocks = key_blocks + [P.Block.long_array(v_short, v_nulls, P.SHORT_DECIMAL), v_long, P.Block.double_array(v_dbl), P.Block.byte_array(mask)]
    c = nk
    aggs = [P.AggregatorFactory(P.AGG_COUNT, 0, -1, c + 3), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, c, c + 3),
            P.AggregatorFactory(P.AGG_AVG, P.SHORT_DECIMAL, c, c + 3), P.AggregatorFactory(P.AGG_SUM, P.LONG_DECIMAL, c + 1),
            P.AggregatorFactory(P.AGG_COUNT_COL, P.SHORT_DECIMAL, c), P.AggregatorFactory(P.AGG_MIN, P.SHORT_DECIMAL, c),
            P.AggregatorFactory(P.AGG_MAX, P.DOUBLE, c + 2), P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, c + 2, c + 3)]
    op = P.HashAggregationOperator(types, list(range(nk)), P.SINGLE, aggs)
    op.add_input(P.Page(*blocks).to_device())
    op.finish()
    got = rows_of(drain(op))
    ogb = O.OracleGroupBy(types, 10_000)
    ids = ogb.get_group_ids(blocks[:nk])
    g = ogb.group_count()
    keys, _ = ogb.export(with_hash=False)
    cols = [k.to_list() for k in keys]
    for a in aggs:
        oa = O.OracleAccumulator(a.kind, a.in_type)
        oa.add_input(ids, g, blocks[a.input_channel] if a.input_channel >= 0 else None, blocks[a.mask_channel] if a.mask_channel >= 0 else None)
        rc, out, _ = oa.evaluate(0, g, True)
        assert rc == 0
        cols.append(out[0].to_list())
    want = [tuple(col[i] for col in cols) for i in range(g)]
    assert len(got) == len(want) == g
    keyf = lambda r: tuple((x is None, x) for x in r[:nk])
    for a, b in zip(sorted(got, key=keyf), sorted(want, key=keyf)):
        for x, y in zip(a, b):
            if isinstance(x, float):
                assert x == pytest.approx(y, rel=1e-12)
            else:
                assert x == y


def _oracle_rows(P, key_block, agg_pages, aggs, step_final=True, key_types=None, key_channels=None):
    """Oracle result in group-id (= first-seen) order: list of tuples (key..., agg0, agg1, ...)."""
    key_types = key_types or [P.BIGINT]
    key_channels = key_channels or [0]
    ogb = O.OracleGroupBy(key_types, 10_000)
    accs = [O.OracleAccumulator(a.kind, a.in_type) for a in aggs]
    for blocks in agg_pages:
        ids = ogb.get_group_ids([blocks[c] for c in key_channels])
        g = ogb.group_count()
        for a, oa in zip(aggs, accs):
            oa.add_input(ids, g, blocks[a.input_channel] if a.input_channel >= 0 else None, blocks[a.mask_channel] if a.mask_channel >= 0 else None)
    g = ogb.group_count()
    keys, _ = ogb.export(with_hash=False)
    cols = [k.to_list() for k in keys]
    for oa in accs:
        rc, out, _ = oa.evaluate(0, g, step_final)
        assert rc == 0
        cols += [o.to_list() for o in out]
    return [tuple(c[i] for c in cols) for i in range(g)]


def _assert_rows_equal(got, want):
    assert len(got) == len(want)
    for a, b in zip(got, want):
        assert len(a) == len(b)
        for x, y in zip(a, b):
            if isinstance(x, float) and isinstance(y, float):
                assert x == pytest.approx(y, rel=1e-12, abs=1e-300)     # double sums: order dependent, tolerance of north_star
            else:
                assert x == y


@pytest.mark.parametrize("split,reorder", [("1", False), ("0", False), ("0", True)])
@pytest.mark.parametrize("expected_groups", [100, 400_000])
@pytest.mark.parametrize("clustered", [False, True])
def test_many_groups_bigint_key_matches_oracle_in_first_seen_order(pgpu, monkeypatch, expected_groups, clustered, split, reorder):
    """cfg3 shape: GROUP BY one BIGINT key with many groups takes the bucket-resident-state path (hashagg_wide.cu).
    Output order must be the reference's first-seen group-id order, across pages; key 0, the null key, negative values,
    null arguments, a mask, long decimals and table growth (expected_groups=100) are all in play."""
    P = pgpu
    rng = np.random.default_rng(21)
    n, distinct = 700_000, 180_000
    universe = np.concatenate([np.array([0, -1, 2**62, -2**63 + 5], dtype=np.int64), rng.integers(-10**12, 10**12, distinct).astype(np.int64)])
    keys = universe[rng.integers(0, len(universe), n)]
    if clustered:
        keys = np.repeat(keys[: n // 4 + 1], 4)[:n]
    knull = (rng.random(n) < 0.001).astype(np.uint8)
    dec = rng.integers(-10**15, 10**15, n).astype(np.int64)
    dnull = (rng.random(n) < 0.05).astype(np.uint8)
    dbl = rng.random(n) * 1e6            # positive like TPC-H prices: the 1e-12 bound is relative to the sum, not to cancellation residue
    big = rng.integers(-2**40, 2**40, n).astype(np.int64)
    long_dec = [int(x) * 10**20 * (1 if i % 3 else -1) for i, x in enumerate(rng.integers(1, 10**9, n))]
    mask = (rng.random(n) < 0.8).astype(np.uint8)
    def page(lo, hi):
        return [P.Block.long_array(keys[lo:hi], knull[lo:hi]), P.Block.long_array(dec[lo:hi], dnull[lo:hi], P.SHORT_DECIMAL),
                P.Block.double_array(dbl[lo:hi]), P.Block.long_array(big[lo:hi]), P.Block.int128_array(long_dec[lo:hi]),
                P.Block.byte_array(mask[lo:hi])]
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 2),
            P.AggregatorFactory(P.AGG_AVG, P.BIGINT, 3), P.AggregatorFactory(P.AGG_SUM, P.LONG_DECIMAL, 4, 5), P.AggregatorFactory(P.AGG_SUM, P.BIGINT, 3),
            P.AggregatorFactory(P.AGG_COUNT_COL, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_AVG, P.SHORT_DECIMAL, 1, 5)]
    bounds = [(0, 400_000), (400_000, 400_007), (400_007, n)]
    pages = [page(lo, hi) for lo, hi in bounds]
    monkeypatch.setenv("PRESTOGPU_WIDE", "1")      # clustered input would otherwise stay on the general path
    monkeypatch.setenv("PRESTOGPU_WIDE_SPLIT", split)   # locate + update kernels (default) / one fused kernel
    if reorder:      # pages rewritten partition-major by hash prefix before the fused kernel (normally only for >= 4 Mi rows, tables >= 128 MB)
        monkeypatch.setenv("PRESTOGPU_WIDE_REORDER_MIN_ROWS", "5")
        monkeypatch.setenv("PRESTOGPU_WIDE_REORDER_SLICE_MB", "0")
    else:
        monkeypatch.setenv("PRESTOGPU_WIDE_REORDER", "0")
    op = P.HashAggregationOperator([P.BIGINT], [0], P.SINGLE, aggs, expected_groups=expected_groups)
    op.ctx.kernel_report(reset=True)
    op.ctx.set_timing(True)
    for i, blocks in enumerate(pages):
        pg_ = P.Page(*blocks)
        op.add_input(pg_.to_device() if i % 2 == 0 else pg_)
    op.finish()
    seen = op.ctx.kernel_report(reset=True)
    op.ctx.set_timing(False)
    assert "wide_agg_jit" in seen or "wide_update_jit" in seen, "the many-groups path was not taken: %s" % sorted(seen)
    assert ("wide_reorder_scatter_kernel" in seen) == reorder, sorted(seen)
    got = rows_of(drain(op))
    want = _oracle_rows(P, None, pages, aggs)
    _assert_rows_equal(got, want)
    assert op.get_group_count() == len(want)


@pytest.mark.parametrize("split", ["1", "0"])
@pytest.mark.parametrize("expected_groups", [100, 150_000, 2_000_000])
def test_many_groups_sorted_keys_take_the_range_chunked_table(pgpu, monkeypatch, expected_groups, split):
    """Rows sorted by key (lineitem by order key) take the many-groups path with the range-chunked front region
    (hashagg_wide.cu).  The first page fixes the plan (k0, chunk width, span); later pages bring a denser key range
    (full chunks spill into the hashed table), keys below k0 and far above the planned span, an unsorted page, key 0 and
    null keys.  Results and first-seen order must equal the oracle's whatever the plan was."""
    P = pgpu
    rng = np.random.default_rng(33)
    def sparse(i):                      # dbgen's order keys: 8 used, 24 skipped
        return ((i >> 3) << 5) | (i & 7)
    k1 = np.repeat(sparse(np.arange(1, 100_001, dtype=np.int64)), rng.integers(1, 8, 100_000))        # sorted runs of 1..7
    k2 = np.repeat(np.arange(k1[-1] + 1, k1[-1] + 120_001, dtype=np.int64), 2)                         # 4x denser than the sample
    k3 = np.concatenate([np.arange(-5000, 0, dtype=np.int64), np.arange(10**12, 10**12 + 20_000, dtype=np.int64)])   # outside the span
    k4 = rng.permutation(np.concatenate([k1[:50_000], k3[:1000], np.zeros(10, dtype=np.int64)]))      # unsorted revisit
    keys = np.concatenate([k1, k2, k3, k4])
    n = len(keys)
    knull = (rng.random(n) < 0.0005).astype(np.uint8)
    dec = rng.integers(-10**15, 10**15, n).astype(np.int64)
    dbl = rng.random(n) * 1e6
    def page(lo, hi):
        return [P.Block.long_array(keys[lo:hi], knull[lo:hi]), P.Block.long_array(dec[lo:hi], None, P.SHORT_DECIMAL), P.Block.double_array(dbl[lo:hi])]
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 2)]
    cuts = [0, len(k1), len(k1) + len(k2), len(k1) + len(k2) + len(k3), n]
    pages = [page(cuts[i], cuts[i + 1]) for i in range(4)]
    monkeypatch.setenv("PRESTOGPU_WIDE_SPLIT", split)
    op = P.HashAggregationOperator([P.BIGINT], [0], P.SINGLE, aggs, expected_groups=expected_groups)
    op.ctx.kernel_report(reset=True)
    op.ctx.set_timing(True)
    for blocks in pages:
        op.add_input(P.Page(*blocks).to_device())
    op.finish()
    seen = op.ctx.kernel_report(reset=True)
    op.ctx.set_timing(False)
    assert "wide_agg_jit" in seen or "wide_update_jit" in seen, "the many-groups path was not taken: %s" % sorted(seen)
    got = rows_of(drain(op))
    want = _oracle_rows(P, None, pages, aggs)
    _assert_rows_equal(got, want)
    assert op.get_group_count() == len(want)


def test_many_groups_partial_then_final(pgpu):
    """PARTIAL step on the many-groups path ships real counts in its intermediate rows; FINAL merges them."""
    P = pgpu
    rng = np.random.default_rng(22)
    n = 500_000
    keys = rng.integers(1, 150_000, n).astype(np.int64)
    dec = rng.integers(-10**9, 10**9, n).astype(np.int64)
    dbl = rng.random(n)
    blocks = [P.Block.long_array(keys), P.Block.long_array(dec, None, P.SHORT_DECIMAL), P.Block.double_array(dbl)]
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_AVG, P.DOUBLE, 2)]
    part = P.HashAggregationOperator([P.BIGINT], [0], P.PARTIAL, aggs, max_partial_memory=1 << 40)
    part.add_input(P.Page(*blocks).to_device())
    part.finish()
    inter = drain(part)
    want_inter = _oracle_rows(P, None, [blocks], aggs, step_final=False)
    _assert_rows_equal(rows_of(inter), want_inter)
    fin_aggs, ch = [], 1
    for a in aggs:
        fin_aggs.append(P.AggregatorFactory(a.kind, a.in_type, ch))
        ch += len(P.intermediate_types(a.kind, a.in_type))
    final = P.HashAggregationOperator([P.BIGINT], [0], P.FINAL, fin_aggs)
    for p in inter:
        final.add_input(p)
    final.finish()
    _assert_rows_equal(rows_of(drain(final)), _oracle_rows(P, None, [blocks], aggs))


def test_packed_pages_carry_partial_output_to_final(pgpu):
    """pg_hashagg_get_output_packed / pg_hashagg_add_packed: the PARTIAL -> FINAL wire format.  Three producers (Q1-shaped plans
    over different row ranges, VARCHAR keys, decimal sum/avg states as VARBINARY, counts) each leave one packed page; a FINAL
    operator fed with the three buffers must give exactly what a FINAL operator fed with the same pages block by block gives,
    and what the oracle computes over all rows in one SINGLE step."""
    P = pgpu
    from presto_b200 import tpch
    cols = tpch.lineitem(60_000 / 5_998_605.2, seed=9, device="cuda")
    page = tpch.page_to_host(tpch.q1_page(cols), pinned=False)
    aggs = tpch.q1_aggregators()
    n = page.position_count
    cuts = [0, n // 3, n // 3 + 11, n]
    bufs, pages = [], []
    import torch
    STRIDE = 1 << 15
    dev_pages = torch.zeros(3 * STRIDE, dtype=torch.uint8, device="cuda")
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        blocks = [b.slice(lo, hi) if hasattr(b, "slice") else None for b in page.blocks]
        if any(b is None for b in blocks):
            pytest.skip("Block.slice not available")
        for packed in (True, False):
            op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.PARTIAL, aggs, max_partial_memory=1 << 40)
            op.add_input(P.Page(*blocks, position_count=hi - lo))
            op.finish()
            if packed:
                buf = np.zeros(1 << 16, dtype=np.uint8)
                with pytest.raises(P.native.PrestoGpuError):
                    op.get_output_packed(buf.ctypes.data, 16)            # too small: PG_NEED_MEMORY, nothing consumed
                nb = op.get_output_packed(buf.ctypes.data, buf.nbytes)
                assert 0 < nb <= buf.nbytes and int(buf[:8].view(np.int64)[0]) == nb
                bufs.append(buf)
                # the same page written into device memory (slot len(bufs)-1 of an all-gather style receive buffer): same size and
                # header (bytes behind the used part of a variable width value area are undefined, so contents are compared below
                # through what a FINAL operator makes of them)
                op._finished, op._next_group = False, 0
                slot = dev_pages[(len(bufs) - 1) * STRIDE:]
                assert op.get_output_packed_device(slot.data_ptr(), STRIDE) == nb
                assert np.array_equal(slot[:24].cpu().numpy(), buf[:24])
            else:
                pages += drain(op)
            op.close()
    fin_aggs, ch = [], 2
    for a in aggs:
        fin_aggs.append(P.AggregatorFactory(a.kind, a.in_type, ch))
        ch += len(P.intermediate_types(a.kind, a.in_type))
    f1 = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, fin_aggs)
    f1.add_packed([b.ctypes.data for b in bufs])
    f1.finish()
    f2 = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, fin_aggs)
    for p in pages:
        f2.add_input(p)
    f2.finish()
    got, want = rows_of(drain(f1)), rows_of(drain(f2))
    assert got == want and len(got) > 0
    f3 = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, fin_aggs)
    f3.add_packed_device(dev_pages.data_ptr(), STRIDE, 3)
    f3.finish()
    assert rows_of(drain(f3)) == want
    single = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.SINGLE, aggs)
    single.add_input(page)
    single.finish()
    assert sorted(got) == sorted(rows_of(drain(single)))


def test_partial_output_travels_as_serialized_pages(pgpu):
    """f3: PARTIAL output pages (VARCHAR keys, decimal sum / avg states as VARBINARY, counts, with nulls in an argument) written in
    the reference's PagesSerde framing (presto-1_b200/pages_serde.py), read back and fed to a FINAL operator: the same rows as feeding the
    pages directly, and as one SINGLE step."""
    P = pgpu
    from presto_b200 import pages_serde, tpch
    cols = tpch.lineitem(40_000 / 5_998_605.2, seed=12, device="cuda")
    page = tpch.page_to_host(tpch.q1_page(cols), pinned=False)
    aggs = tpch.q1_aggregators()
    n = page.position_count
    wire, direct = [], []
    for lo, hi in ((0, n // 2), (n // 2, n)):
        op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.PARTIAL, aggs, max_partial_memory=1 << 40)
        op.add_input(P.Page(*[b.slice(lo, hi) for b in page.blocks], position_count=hi - lo))
        op.finish()
        out = drain(op)
        op.close()
        direct += out
        wire.append(pages_serde.write_pages(out))
    inter_types = [P.VARCHAR, P.VARCHAR]
    fin_aggs, ch = [], 2
    for a in aggs:
        fin_aggs.append(P.AggregatorFactory(a.kind, a.in_type, ch))
        inter_types += list(P.intermediate_types(a.kind, a.in_type))
        ch += len(P.intermediate_types(a.kind, a.in_type))
    assert [b.type for b in direct[0].blocks] == inter_types
    f1 = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, fin_aggs)
    for buf in wire:
        for pg in pages_serde.read_pages(buf, inter_types):
            f1.add_input(pg)
    f1.finish()
    f2 = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.FINAL, fin_aggs)
    for pg in direct:
        f2.add_input(pg)
    f2.finish()
    got, want = rows_of(drain(f1)), rows_of(drain(f2))
    assert got == want and len(got) == 4
    single = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.SINGLE, aggs)
    single.add_input(page)
    single.finish()
    assert sorted(got) == sorted(rows_of(drain(single)))


def test_operator_yields_on_memory_limit_and_counts_every_row_once(pgpu):
    """HashAggregationOperator's unfinishedWork (HashAggregationOperator.java:398-417, 473-481): with a memory limit the page that
    would grow the table is kept, needsInput() is false and getOutput() gives nothing until the memory is granted; then every row
    counts exactly once (sums bit-exact against the oracle)."""
    P = pgpu
    rng = np.random.default_rng(5)
    n = 2_500_000
    keys = rng.permutation(n).astype(np.int64) % 1_800_000        # more distinct keys than the starting table's fill limit
    vals = rng.integers(-10**12, 10**12, n).astype(np.int64)
    blocks = [P.Block.long_array(keys), P.Block.long_array(vals, None, P.SHORT_DECIMAL)]
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1)]
    op = P.HashAggregationOperator([P.BIGINT], [0], P.SINGLE, aggs, expected_groups=10)
    limit = op.get_device_bytes()             # what the empty operator holds: any growth needs a grant
    op.set_memory_limit(limit)
    op.add_input(P.Page(*blocks))
    assert op.has_unfinished_input() and not op.needs_input()
    assert op.get_group_count() == 0
    op.finish()
    assert op.get_output() is None and op.has_unfinished_input()      # still blocked: no output, nothing consumed
    grants = 0
    while op.has_unfinished_input():
        limit *= 2
        grants += 1
        op.set_memory_limit(limit)
        op._process_unfinished()
        assert grants < 40
    assert grants >= 1
    got = rows_of(drain(op))
    want = _oracle_rows(P, None, [blocks], aggs)
    _assert_rows_equal(got, want)


@pytest.mark.parametrize("shape", ["q1_varchar_keys", "bigint_key_many_groups"])
def test_reference_sized_pages_through_the_accumulator_equal_one_big_page(pgpu, shape):
    """pg_hashagg_set_page_batching: 1 000-row host pages (nulls first appearing in a later page, a nullable mask, variable width
    keys) accumulate and enter the kernels a batch at a time; group ids keep their first-seen order and every aggregate equals
    the oracle's — also when the stream ends in the middle of a batch and when stats are read half way."""
    P = pgpu
    rng = np.random.default_rng(11)
    n = 57_321
    if shape == "q1_varchar_keys":
        flags = np.array(["A", "N", "R"])[rng.integers(0, 3, n)]
        status = np.array(["F", "O"])[rng.integers(0, 2, n)]
        keys = [P.Block.variable_width(list(flags)), P.Block.variable_width(list(status))]
        key_types, key_channels = [P.VARCHAR, P.VARCHAR], [0, 1]
    else:
        keys = [P.Block.long_array(rng.integers(0, 20_000, n).astype(np.int64))]
        key_types, key_channels = [P.BIGINT], [0]
    nk = len(keys)
    dec = rng.integers(-10**12, 10**12, n).astype(np.int64)
    dnull = (rng.random(n) < 0.1).astype(np.uint8)
    dnull[:20_000] = 0                       # the first pages arrive without a null mask
    dbl = rng.random(n) * 1e4
    mask = (rng.random(n) < 0.7).astype(np.uint8)
    cols = keys + [P.Block.long_array(dec, dnull, P.SHORT_DECIMAL), P.Block.double_array(dbl), P.Block.byte_array(mask)]
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, nk), P.AggregatorFactory(P.AGG_AVG, P.SHORT_DECIMAL, nk, nk + 2),
            P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, nk + 1), P.AggregatorFactory(P.AGG_MAX, P.SHORT_DECIMAL, nk)]

    def page(lo, hi):
        out = []
        for b in cols:
            if b.type in (P.VARCHAR, P.VARBINARY):
                out.append(P.Block(b.type, b.values, None, b.offsets[lo:hi + 1], hi - lo))
            else:
                nulls = None if b.nulls is None or not b.nulls[lo:hi].any() and hi <= 20_000 else b.nulls[lo:hi]
                out.append(P.Block(b.type, b.values[lo:hi], nulls, None, hi - lo))
        return out

    pages = [page(lo, min(lo + 1000, n)) for lo in range(0, n, 1000)]
    op = P.HashAggregationOperator(key_types, key_channels, P.SINGLE, aggs, expected_groups=100)
    op.set_page_batching(16_000)
    op.ctx.kernel_report(reset=True)
    op.ctx.set_timing(True)
    for i, pg_ in enumerate(pages):
        op.add_input(P.Page(*pg_))
        if i == 30:
            assert op._stats()[3] == 31_000        # input rows: reading the stats hands the accumulated pages in
    op.finish()
    got = rows_of(drain(op))
    launches = sum(v[1] for v in op.ctx.kernel_report(reset=True).values())
    op.ctx.set_timing(False)
    want = _oracle_rows(P, None, [cols], aggs, key_types=key_types, key_channels=key_channels) if shape == "q1_varchar_keys" else _oracle_rows(P, None, [cols], aggs)
    _assert_rows_equal(got, want)
    assert launches < 40 * 6, "the pages were not batched: %d kernel launches for %d pages" % (launches, len(pages))
    # the bulk form of the same thing
    op2 = P.HashAggregationOperator(key_types, key_channels, P.SINGLE, aggs, expected_groups=100)
    op2.set_page_batching(16_000)
    op2.add_pages(P.Page(*cols), 1000)
    op2.finish()
    _assert_rows_equal(rows_of(drain(op2)), want)


def test_caller_pinned_pages_go_to_the_device_without_a_host_copy(pgpu):
    """PG_MEM_HOST_PINNED pages through the accumulator's device mode (variable width keys with re-based offsets, a null mask that
    first appears mid-stream, a batch boundary inside the stream, plain host pages mixed in) give the oracle's rows."""
    import torch
    P = pgpu
    rng = np.random.default_rng(13)
    n = 41_003
    flags = np.array(["A", "NN", "R"])[rng.integers(0, 3, n)]
    keyb = P.Block.variable_width(list(flags))
    dec = rng.integers(-10**12, 10**12, n).astype(np.int64)
    dnull = (rng.random(n) < 0.1).astype(np.uint8)
    dnull[:15_000] = 0
    dbl = rng.random(n) * 1e3
    cols = [keyb, P.Block.long_array(dec, dnull, P.SHORT_DECIMAL), P.Block.double_array(dbl)]
    aggs = [P.AggregatorFactory(P.AGG_COUNT), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, 1), P.AggregatorFactory(P.AGG_SUM, P.DOUBLE, 2),
            P.AggregatorFactory(P.AGG_COUNT_COL, P.SHORT_DECIMAL, 1)]

    def pin(a):
        return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()

    pinned = [P.Block(b.type, pin(b.values), pin(b.nulls), pin(b.offsets), b.n) for b in cols]
    keep = pinned        # the pinned arrays stay alive until the operator has been drained (the PG_MEM_HOST_PINNED promise)

    def page(src, lo, hi, is_pinned):
        out = []
        for b in src:
            if b.type in (P.VARCHAR, P.VARBINARY):
                nb = P.Block(b.type, b.values, None, b.offsets[lo:hi + 1], hi - lo)
            else:
                nulls = None if (b.nulls is None or hi <= 15_000) else b.nulls[lo:hi]
                nb = P.Block(b.type, b.values[lo:hi], nulls, None, hi - lo)
            nb.pinned = is_pinned
            out.append(nb)
        return out

    op = P.HashAggregationOperator([P.VARCHAR], [0], P.SINGLE, aggs, expected_groups=100)
    op.set_page_batching(12_000)
    for i, lo in enumerate(range(0, n, 1500)):
        hi = min(lo + 1500, n)
        plain = i in (9, 10)                    # two plain host pages in the middle: the accumulator switches modes
        op.add_input(P.Page(*page(cols if plain else pinned, lo, hi, not plain)))
    op.finish()
    got = rows_of(drain(op))
    want = _oracle_rows(P, None, [cols], aggs, key_types=[P.VARCHAR], key_channels=[0])
    _assert_rows_equal(got, want)
    assert keep
