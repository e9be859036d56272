"""GPU parity: join build + probe (pg_joinbuild_*, pg_lookup_*) vs the oracle and vs the expected rows of
OPT/TestHashJoinOperator.java (testInnerJoin :165-202, testInnerJoinWithNullProbe :547-583, ...WithNullBuild :585-622,
...WithNullOnBothSides :624-667, testProbeOuterJoin :668-710, empty lookup source :1036-1200).
Join output is compared in ORDER (the reference's assertOperatorEquals is ordered)."""
import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


def seq_varchar(P, n, start):
    return P.Block.variable_width([str(start + i) for i in range(n)])


def seq_bigint(P, n, start):
    return P.Block.long_array(np.arange(start, start + n, dtype=np.int64))


def page_rows(page):
    cols = [b.to_host().to_list() for b in page.blocks]
    return [tuple(c[i].decode() if isinstance(c[i], bytes) else c[i] for c in cols) for i in range(page.position_count)]


def run_join(P, build_pages, probe_pages, types, join_type, hash_build=False, hash_probe=False, device=False):
    """Drive HashBuilderOperator + LookupJoinOperator like JoinTestUtils: join channel 0, all channels output."""
    btypes = list(types)
    hb = None
    if hash_build:
        build_pages = [P.Page(*(p.blocks + [P.Block.long_array(O.hash_rows([p.blocks[0]]))])) for p in build_pages]
        btypes = btypes + [P.BIGINT]
        hb = len(types)
    builder = P.HashBuilderOperator(btypes, [0], hb)
    for p in build_pages:
        builder.add_input(p.to_device() if device else p)
    builder.finish()
    ls = builder.get_lookup_source()
    hp = None
    if hash_probe:
        probe_pages = [P.Page(*(p.blocks + [P.Block.long_array(O.hash_rows([p.blocks[0]]))])) for p in probe_pages]
        hp = len(probe_pages[0].blocks) - 1
    nprobe_out = len(probe_pages[0].blocks) - (1 if hash_probe else 0)
    op = P.LookupJoinOperator(ls, [0], list(range(nprobe_out)), list(range(len(types))), join_type, hp)
    rows = []
    for p in probe_pages:
        op.add_input(p.to_device() if device else p)
        out = op.get_output()
        rows += page_rows(out)
    op.finish()
    assert op.is_finished()
    return rows


HASH_GRID = [(False, False), (True, False), (False, True), (True, True)]


@pytest.mark.parametrize("hash_probe,hash_build", HASH_GRID)
@pytest.mark.parametrize("device", [False, True])
def test_inner_join_reference_rows(pgpu, hash_probe, hash_build, device):
    P = pgpu
    types = [P.VARCHAR, P.BIGINT, P.BIGINT]
    build = [P.Page(seq_varchar(P, 10, 20), seq_bigint(P, 10, 30), seq_bigint(P, 10, 40))]
    probe = [P.Page(seq_varchar(P, 1000, 0), seq_bigint(P, 1000, 1000), seq_bigint(P, 1000, 2000))]
    rows = run_join(P, build, probe, types, P.JOIN_INNER, hash_build, hash_probe, device)
    expected = [(str(20 + i), 1020 + i, 2020 + i, str(20 + i), 30 + i, 40 + i) for i in range(10)]
    assert rows == expected


@pytest.mark.parametrize("hash_probe,hash_build", HASH_GRID)
def test_inner_join_nulls(pgpu, hash_probe, hash_build):
    P = pgpu
    V = lambda xs: P.Page(P.Block.variable_width(xs))
    t = [P.VARCHAR]
    # testInnerJoinWithNullProbe
    assert run_join(P, [V(["a", "b", "c"])], [V(["a", None, None, "a", "b"])], t, P.JOIN_INNER, hash_build, hash_probe) == \
        [("a", "a"), ("a", "a"), ("b", "b")]
    # testInnerJoinWithNullBuild: build a,null,null,a,b ; probe a,b,c -> a matches twice (newest build row first)
    assert run_join(P, [V(["a", None, None, "a", "b"])], [V(["a", "b", "c"])], t, P.JOIN_INNER, hash_build, hash_probe) == \
        [("a", "a"), ("a", "a"), ("b", "b")]
    # testInnerJoinWithNullOnBothSides
    assert run_join(P, [V(["a", None, None, "a", "b"])], [V(["a", "b", None, "c"])], t, P.JOIN_INNER, hash_build, hash_probe) == \
        [("a", "a"), ("a", "a"), ("b", "b")]


@pytest.mark.parametrize("hash_probe,hash_build", HASH_GRID)
def test_probe_outer_join(pgpu, hash_probe, hash_build):
    P = pgpu
    types = [P.VARCHAR, P.BIGINT, P.BIGINT]
    build = [P.Page(seq_varchar(P, 10, 20), seq_bigint(P, 10, 30), seq_bigint(P, 10, 40))]
    probe = [P.Page(seq_varchar(P, 15, 20), seq_bigint(P, 15, 1020), seq_bigint(P, 15, 2020))]
    rows = run_join(P, build, probe, types, P.JOIN_PROBE_OUTER, hash_build, hash_probe)
    expected = [(str(20 + i), 1020 + i, 2020 + i, str(20 + i), 30 + i, 40 + i) for i in range(10)]
    expected += [(str(30 + i), 1030 + i, 2030 + i, None, None, None) for i in range(5)]   # testProbeOuterJoin :668-710
    assert rows == expected
    V = lambda xs: P.Page(P.Block.variable_width(xs))
    # testOuterJoinWithNullProbe / NullBuild / NullOnBothSides :760-900
    assert run_join(P, [V(["a", "b", "c"])], [V(["a", None, None, "a", "b"])], [P.VARCHAR], P.JOIN_PROBE_OUTER, hash_build, hash_probe) == \
        [("a", "a"), (None, None), (None, None), ("a", "a"), ("b", "b")]
    assert run_join(P, [V(["a", None, None, "a", "b"])], [V(["a", "b", "c"])], [P.VARCHAR], P.JOIN_PROBE_OUTER, hash_build, hash_probe) == \
        [("a", "a"), ("a", "a"), ("b", "b"), ("c", None)]


def test_empty_lookup_source(pgpu):
    P = pgpu
    V = lambda xs: P.Page(P.Block.variable_width(xs))
    assert run_join(P, [], [V(["a", "b", None])], [P.VARCHAR], P.JOIN_INNER) == []
    assert run_join(P, [], [V(["a", "b", None])], [P.VARCHAR], P.JOIN_PROBE_OUTER) == [("a", None), ("b", None), (None, None)]


def _key_block(P, rng, t, n, card, null_rate):
    nulls = (rng.random(n) < null_rate).astype(np.uint8) if null_rate else None
    if t == P.BIGINT:
        v = rng.integers(0, card, n).astype(np.int64)
        return P.Block.long_array(v, nulls)
    if t in (P.INTEGER, P.DATE):
        return P.Block.int_array(rng.integers(0, card, n).astype(np.int32), nulls, t)
    if t == P.DOUBLE:
        v = rng.integers(0, card, n).astype(np.float64) / 2
        v[rng.random(n) < 0.02] = np.nan
        v[rng.random(n) < 0.02] = -0.0
        return P.Block.double_array(v, nulls)
    if t == P.VARCHAR:
        return P.Block.variable_width([None if (nulls is not None and nulls[i]) else "s%d" % x for i, x in enumerate(rng.integers(0, card, n))])
    if t == P.LONG_DECIMAL:
        return P.Block.int128_array([int(x) * 10**19 for x in rng.integers(-card, card, n)], nulls)
    raise AssertionError


_payload_log = []

JOIN_CASES = [
    ("bigint_unique", lambda P: [P.BIGINT], None, 0.0),
    ("bigint_dups", lambda P: [P.BIGINT], 500, 0.05),
    ("bigint_heavy_skew", lambda P: [P.BIGINT], 3, 0.0),
    ("varchar", lambda P: [P.VARCHAR], 300, 0.05),
    ("composite", lambda P: [P.BIGINT, P.DATE, P.INTEGER], 12, 0.03),
    ("double_dec", lambda P: [P.DOUBLE, P.LONG_DECIMAL], 20, 0.02),
]


@pytest.mark.parametrize("name,types_fn,card,null_rate", JOIN_CASES, ids=[c[0] for c in JOIN_CASES])
@pytest.mark.parametrize("join_type", [0, 1])
def test_random_join_matches_oracle_in_order(pgpu, name, types_fn, card, null_rate, join_type):
    P = pgpu
    rng = np.random.default_rng(11)
    ktypes = types_fn(P)
    nk = len(ktypes)
    nbuild, nprobe = 5000, (300 if name == "bigint_heavy_skew" else 20000)
    types = ktypes + [P.DOUBLE, P.VARCHAR]   # payload channels
    idx = PIdx = P.PagesIndex(types, list(range(nk)))
    orc = O.OracleJoin(types, list(range(nk)))
    _payload_log.clear()
    for pg in range(3):
        if card is None:
            keys = [P.Block.long_array(rng.permutation(nbuild * 4)[:nbuild].astype(np.int64) + pg * nbuild * 4)]
        else:
            keys = [_key_block(P, rng, t, nbuild, card, null_rate) for t in ktypes]
        pay = rng.random(nbuild)
        _payload_log.append(pay)
        payload = [P.Block.double_array(pay), P.Block.variable_width([None if i % 7 == 0 else "p%d" % (i % 50) for i in range(nbuild)])]
        blocks = keys + payload
        idx.add_page([b.to_device() for b in blocks] if pg == 1 else blocks)
        orc.add_page(blocks)
    ls = idx.create_lookup_source()
    orc.build()
    assert ls.get_join_position_count() == orc.rows() == 3 * nbuild
    pcard = card if card is not None else nbuild * 12
    pkeys = [_key_block(P, rng, t, nprobe, pcard, null_rate) for t in ktypes]
    for device in (False, True):
        pk = [b.to_device() for b in pkeys] if device else pkeys
        gp, gb = ls.join_page(pk, None, join_type)
        wp, wb = orc.probe(pkeys, None, join_type)
        gp = gp.cpu().numpy() if device else gp
        gb = gb.cpu().numpy() if device else gb
        assert np.array_equal(gp, wp), name
        assert np.array_equal(gb, wb), name   # order-sensitive: chains are newest build row first
    # appendTo: gather payload + keys and compare with numpy gathers on the host copy
    wb = wb[:20000]
    out = ls.append_to(wb, list(range(len(types))))
    assert out[nk].n == len(wb)
    dev_out = ls.append_to(__import__("torch").from_numpy(wb).cuda(), list(range(len(types))))
    for a, b in zip(out, dev_out):
        a, b = a.to_host(), b.to_host()
        assert np.array_equal(a.values, b.values) and a.n == b.n
        assert (a.nulls is None and b.nulls is None) or np.array_equal(a.nulls, b.nulls)
    # the gathered payload equals the oracle-side build rows: check the double payload channel through numpy
    allpay = np.concatenate([pg for pg in _payload_log])
    sel = wb >= 0
    assert np.array_equal(out[nk].to_host().values[sel], allpay.view(np.int64)[wb[sel]])


def test_large_skewed_segment_sorted_descending(pgpu):
    """one key repeated 5000 times exercises the block-wide segment sorter"""
    P = pgpu
    n = 5000
    keys = np.zeros(n + 100, dtype=np.int64)
    keys[n:] = np.arange(1, 101)
    idx = P.PagesIndex([P.BIGINT], [0])
    idx.add_page([P.Block.long_array(keys)])
    ls = idx.create_lookup_source()
    assert ls.has_duplicates()
    gp, gb = ls.join_page([P.Block.long_array([0, 5, 0])])
    want = list(range(n - 1, -1, -1))
    assert gb.tolist() == want + [n + 4] + want
    assert gp.tolist() == [0] * n + [1] + [2] * n


RANGE_CASES = ["tpch_sparse_sorted", "dense_sorted_with_dups", "two_far_clusters", "heavy_skew_overflow", "extreme_span", "with_nulls_and_sentinel"]


@pytest.mark.parametrize("case", RANGE_CASES)
@pytest.mark.parametrize("join_type", [0, 1])
def test_bigint_join_range_chunked_table_matches_oracle(pgpu, case, join_type):
    """Single BIGINT join key, enough build rows for the range-chunked slot placement (join.cu: JLayout).  Key distributions that
    suit it (dbgen's sparse sorted order keys, dense keys), that overflow single lines into the hashed region (clusters, skew:
    the build then falls back to the plain table when the overflow region fills up), keys at the int64 extremes, nulls on both
    sides and the EMPTY_KEY bit pattern.  Match lists are compared with the oracle in order (newest build row first)."""
    P = pgpu
    rng = np.random.default_rng(77)
    n = 60_000
    i = np.arange(1, n + 1, dtype=np.int64)
    sparse = ((i >> 3) << 5) | (i & 7)
    knull = None
    if case == "tpch_sparse_sorted":
        bkeys = sparse
    elif case == "dense_sorted_with_dups":
        bkeys = np.sort(rng.integers(1000, 1000 + n // 2, n).astype(np.int64))
    elif case == "two_far_clusters":
        bkeys = np.concatenate([np.arange(n // 2, dtype=np.int64), 10**15 + np.arange(n // 2, dtype=np.int64) * 3])
    elif case == "heavy_skew_overflow":
        bkeys = np.concatenate([np.arange(50_000, dtype=np.int64) * 7, rng.integers(0, 2**40, n - 50_000).astype(np.int64)])
        bkeys = rng.permutation(bkeys)
    elif case == "extreme_span":
        bkeys = np.concatenate([rng.integers(-2**63, 2**63 - 1, n - 2, dtype=np.int64), np.array([-2**63, 2**63 - 1], dtype=np.int64)])
    else:
        bkeys = sparse.copy()
        bkeys[::1000] = -0x7FFFFFFFFFFFFFFF          # the bit pattern of the table's empty marker (0x8000000000000001)
        knull = (rng.random(n) < 0.01).astype(np.uint8)
    pay = rng.random(n)
    types = [P.BIGINT, P.DOUBLE]
    idx = P.PagesIndex(types, [0])
    orc = O.OracleJoin(types, [0])
    cuts = [0, n // 3, n // 3 + 5, n]
    for a, b in zip(cuts[:-1], cuts[1:]):
        blocks = [P.Block.long_array(bkeys[a:b], None if knull is None else knull[a:b]), P.Block.double_array(pay[a:b])]
        idx.add_page([x.to_device() for x in blocks])
        orc.add_page(blocks)
    ls = idx.create_lookup_source()
    orc.build()
    m = 150_000
    pkeys = np.concatenate([bkeys[rng.integers(0, n, m // 2)], rng.integers(int(bkeys.min()) // 2 - 5, int(bkeys.max()) // 2 + 5, m - m // 2).astype(np.int64)])
    pkeys[:: 5000] = -0x7FFFFFFFFFFFFFFF
    pkeys = np.sort(pkeys) if case in ("tpch_sparse_sorted", "dense_sorted_with_dups") else pkeys
    pnull = (rng.random(m) < 0.01).astype(np.uint8)
    pk = P.Block.long_array(pkeys, pnull)
    gp, gb = ls.join_page([pk.to_device()], None, join_type)
    wp, wb = orc.probe([pk], None, join_type)
    assert np.array_equal(gp.cpu().numpy(), wp), case
    assert np.array_equal(gb.cpu().numpy(), wb), case
    out = ls.append_to(__import__("torch").from_numpy(wb).cuda(), [1])
    sel = wb >= 0
    assert np.array_equal(out[0].to_host().values[sel], pay.view(np.int64)[wb[sel]])
    # the fused step (pg_lookup_join_page) must give the same probe positions and the same gathered channel, nulls included
    fp, fout = ls.join_and_append([pk.to_device()], [1, 0], None, join_type)
    assert np.array_equal(fp.cpu().numpy(), wp), case
    got = fout[0].to_host()
    assert np.array_equal(got.values[sel], pay.view(np.int64)[wb[sel]]), case
    assert np.array_equal(np.asarray(got.nulls) != 0, ~sel), case
    gk = fout[1].to_host()
    assert np.array_equal(gk.values[sel], bkeys[wb[sel]]), case


@pytest.mark.parametrize("join_type", [0, 1])
@pytest.mark.parametrize("all_match", [True, False])
def test_fused_join_step_unique_build_keys(pgpu, join_type, all_match):
    """pg_lookup_join_page on the foreign-key shape (unique build keys): the one-kernel path when every probe row emits, the
    compacting fallback when INNER rows miss; host outputs take the two-step path.  Compared with the oracle's probe + appendTo."""
    P = pgpu
    rng = np.random.default_rng(91)
    n, m = 40_000, 130_001
    i = np.arange(1, n + 1, dtype=np.int64)
    bkeys = ((i >> 3) << 5) | (i & 7)
    pay = rng.integers(-10**15, 10**15, n).astype(np.int64)
    pnulls = (rng.random(n) < 0.05).astype(np.uint8)
    big = [int(x) * 10**21 * (-1 if x % 2 else 1) for x in rng.integers(1, 10**9, n)]
    types = [P.BIGINT, P.SHORT_DECIMAL, P.LONG_DECIMAL]
    blocks = [P.Block.long_array(bkeys), P.Block.long_array(pay, pnulls, P.SHORT_DECIMAL), P.Block.int128_array(big)]
    idx = P.PagesIndex(types, [0])
    orc = O.OracleJoin(types, [0])
    idx.add_page([b.to_device() for b in blocks])
    orc.add_page(blocks)
    ls = idx.create_lookup_source()
    orc.build()
    pkeys = bkeys[rng.integers(0, n, m)]
    if not all_match:
        pkeys[::7] += 8            # keys dbgen never uses
    pk = P.Block.long_array(pkeys, None if all_match else (rng.random(m) < 0.01).astype(np.uint8))
    wp, wb = orc.probe([pk], None, join_type)
    sel = wb >= 0
    bigv = np.asarray(blocks[2].values).reshape(n, 2)
    for device in (True, False):
        fp, fout = ls.join_and_append([pk.to_device() if device else pk], [2, 1], None, join_type, device=device)
        fp = fp.cpu().numpy() if device else fp
        assert np.array_equal(fp, wp)
        g2, g1 = fout[0].to_host(), fout[1].to_host()
        assert np.array_equal(np.asarray(g2.values).reshape(-1, 2)[sel], bigv[wb[sel]])
        assert np.array_equal(np.asarray(g2.nulls) != 0, ~sel)
        want_null = ~sel
        want_null[sel] = pnulls[wb[sel]] != 0
        assert np.array_equal(np.asarray(g1.nulls) != 0, want_null)
        ok = ~want_null
        assert np.array_equal(np.asarray(g1.values)[ok], pay[wb[ok]])


@pytest.mark.parametrize("misses", ["sampled", "unsampled", "none"])
def test_fused_join_step_samples_large_pages(pgpu, misses):
    """Large INNER probe pages are sampled (start, middle, end) before the one-kernel path is tried: misses inside a sample go straight to
    the compacting path, misses elsewhere are caught by the kernel's own count.  Either way the output equals the oracle's."""
    P = pgpu
    rng = np.random.default_rng(17)
    n, m = 50_000, 300_000
    bkeys = np.arange(1, n + 1, dtype=np.int64) * 3
    pay = rng.integers(-10**12, 10**12, n).astype(np.int64)
    types = [P.BIGINT, P.BIGINT]
    blocks = [P.Block.long_array(bkeys), P.Block.long_array(pay)]
    idx = P.PagesIndex(types, [0])
    orc = O.OracleJoin(types, [0])
    idx.add_page([b.to_device() for b in blocks])
    orc.add_page(blocks)
    ls = idx.create_lookup_source()
    orc.build()
    pkeys = bkeys[rng.integers(0, n, m)]
    if misses == "sampled":
        pkeys[::11] += 1
    elif misses == "unsampled":
        pkeys[20_000:30_000:3] += 1
    pk = P.Block.long_array(pkeys)
    wp, wb = orc.probe([pk], None, 0)
    fp, fout = ls.join_and_append([pk.to_device()], [1], None, 0)
    assert np.array_equal(fp.cpu().numpy(), wp)
    assert np.array_equal(np.asarray(fout[0].to_host().values), pay[wb])
